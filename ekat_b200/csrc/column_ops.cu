// column_ops.cu -- the column-ops neighbours of the Pack glue for every column of a batch:
// ekat::adj_diff, ekat::shift_left / shift_right, ekat::index_and_shift.
//
// Replaces the per-column, per-index loop around ekat::adj_diff<Forward>(view, index, index_end) (reference
// src/pack/ekat_pack_kokkos.hpp:100-164; call form tests/pack/pack_kokkos_utils.cpp:1055-1080) -- the finite-difference
// step column physics runs right before it fills (dl, d, du):
//   Forward : out[k] = (k < n-1 ? in[k+1] : 0) - in[k]          (:149-152; packs: :118-121)
//   backward: out[k] = in[k] - (k > 0 ? in[k-1] : 0)            (:153-157; packs: :122-126)
// over the n scalars of the (scalarized) column; for a view of Pack<Real,N> n = npack*N, padding lanes included, exactly
// as the reference's shift_left / shift_right with a 0 fill behave. One rounded subtraction per element: bit-identical.
//
// Roofline: HBM, 2 * n * word bytes per column (one read, one write).
//   COL_MAJOR in[c*pitch + k]: 16-byte vectors along the column, the neighbour element across the vector boundary is a
//     second (L1-resident) scalar load;
//   LEV_MAJOR in[k*pitch + c]: one thread per column walks the levels with the previous row in a register and eight
//     rows in flight, so every row is read from HBM once.
#include "tridiag_lm_common.cuh"  // Vec

namespace ekb {

// One work item = one 16-byte vector (VEC scalars) of one column, or one scalar when the operands are not 16-byte
// aligned; items are dealt out flat over the whole batch so that no thread idles whatever the column length is.
template <typename R, bool FWD, bool VECTOR>
__global__ void __launch_bounds__(256)
k_adj_diff_cm(const R* __restrict__ in, R* __restrict__ out, int64_t ncol, int n, int64_t pitch_in, int64_t pitch_out) {
  constexpr int VEC = VECTOR ? 16 / sizeof(R) : 1;
  const unsigned per_col = (unsigned)(n / VEC);  // n % VEC == 0 when VECTOR
  const int64_t items = ncol * per_col;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < items; i += (int64_t)gridDim.x * 256) {
    const int64_t c = items < ((int64_t)1 << 32) ? (int64_t)((unsigned)i / per_col) : i / per_col;
    const int k0 = (int)(i - c * per_col) * VEC;
    const R* x = in + (size_t)c * pitch_in;
    R* o = out + (size_t)c * pitch_out;
    if (VECTOR) {
      Vec<R, VEC> v = *reinterpret_cast<const Vec<R, VEC>*>(x + k0), r;
      if (FWD) {
        const R nxt = k0 + VEC < n ? __ldg(x + k0 + VEC) : (R)0;
#pragma unroll
        for (int j = 0; j < VEC; ++j) r.v[j] = x_sub(j + 1 < VEC ? v.v[j + 1] : nxt, v.v[j]);
      } else {
        const R prv = k0 > 0 ? __ldg(x + k0 - 1) : (R)0;
#pragma unroll
        for (int j = 0; j < VEC; ++j) r.v[j] = x_sub(v.v[j], j > 0 ? v.v[j - 1] : prv);
      }
      *reinterpret_cast<Vec<R, VEC>*>(o + k0) = r;
    } else {
      if (FWD) o[k0] = x_sub(k0 < n - 1 ? __ldg(x + k0 + 1) : (R)0, __ldg(x + k0));
      else o[k0] = x_sub(__ldg(x + k0), k0 > 0 ? __ldg(x + k0 - 1) : (R)0);
    }
  }
}

template <typename R, bool FWD>
__global__ void __launch_bounds__(128)
k_adj_diff_lm(const R* __restrict__ in, R* __restrict__ out, int64_t ncol, int n, int64_t pitch_in, int64_t pitch_out) {
  const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= ncol) return;
  const R* x = in + c;
  R* o = out + c;
  if (FWD) {
    R cur = n > 0 ? ld_stream(x) : (R)0;
    int k = 0;
    for (; k + 8 < n; k += 8) {  // rows k+1 .. k+8 exist
      R v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = ld_stream(x + (size_t)(k + 1 + u) * pitch_in);
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        st_stream(o + (size_t)(k + u) * pitch_out, x_sub(v[u], cur));
        cur = v[u];
      }
    }
    for (; k < n; ++k) {
      const R nxt = k < n - 1 ? ld_stream(x + (size_t)(k + 1) * pitch_in) : (R)0;
      st_stream(o + (size_t)k * pitch_out, x_sub(nxt, cur));
      cur = nxt;
    }
  } else {
    R prv = (R)0;
    int k = 0;
    for (; k + 8 <= n; k += 8) {
      R v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = ld_stream(x + (size_t)(k + u) * pitch_in);
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        st_stream(o + (size_t)(k + u) * pitch_out, x_sub(v[u], prv));
        prv = v[u];
      }
    }
    for (; k < n; ++k) {
      const R v = ld_stream(x + (size_t)k * pitch_in);
      st_stream(o + (size_t)k * pitch_out, x_sub(v, prv));
      prv = v;
    }
  }
}

template <typename R>
int adj_diff(const R* in, R* out, int64_t ncol, int n, int64_t pitch_in, int64_t pitch_out, int col_major, int forward,
             cudaStream_t st) {
  if (ncol == 0 || n == 0) return EKAT_B200_OK;
  if (!col_major) {
    const unsigned blocks = (unsigned)((ncol + 127) / 128);
    if (forward) k_adj_diff_lm<R, true><<<blocks, 128, 0, st>>>(in, out, ncol, n, pitch_in, pitch_out);
    else k_adj_diff_lm<R, false><<<blocks, 128, 0, st>>>(in, out, ncol, n, pitch_in, pitch_out);
    note_launch(sizeof(R) == 8 ? "adj_diff_lm<f64>" : "adj_diff_lm<f32>");
    return check_cuda(cudaGetLastError());
  }
  constexpr int VEC = 16 / sizeof(R);
  const int vec_ok = n % VEC == 0 && pitch_in % VEC == 0 && pitch_out % VEC == 0 &&
                     (reinterpret_cast<uintptr_t>(in) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0;
  const int64_t items = ncol * (vec_ok ? n / VEC : n);
  const unsigned blocks = (unsigned)tmin<int64_t>((items + 255) / 256, (int64_t)148 * 8 * 64);
  if (vec_ok) {
    if (forward) k_adj_diff_cm<R, true, true><<<blocks, 256, 0, st>>>(in, out, ncol, n, pitch_in, pitch_out);
    else k_adj_diff_cm<R, false, true><<<blocks, 256, 0, st>>>(in, out, ncol, n, pitch_in, pitch_out);
  } else {
    if (forward) k_adj_diff_cm<R, true, false><<<blocks, 256, 0, st>>>(in, out, ncol, n, pitch_in, pitch_out);
    else k_adj_diff_cm<R, false, false><<<blocks, 256, 0, st>>>(in, out, ncol, n, pitch_in, pitch_out);
  }
  note_launch(sizeof(R) == 8 ? "adj_diff_cm<f64>" : "adj_diff_cm<f32>");
  return check_cuda(cudaGetLastError());
}

// ------------------------------------------------------------------------------------------------------------
// ekat::shift_right / ekat::shift_left over a whole column (reference src/pack/ekat_pack.hpp:521-555): applied pack
// after pack along a scalarized column, shift_right(v(k-1), v(k)) moves every scalar one place up the column and
// shift_left(v(k+1), v(k)) one place down, the first / last place taking the scalar `fill` (the ScalarType overloads,
// :531-538, :549-555 -- what adj_diff passes at the ends, ekat_pack_kokkos.hpp:119,124). Generalised to any shift s:
//   out[k] = (0 <= k - s < n) ? in[k - s] : fill        s = +1: shift_right, s = -1: shift_left
// A copy, so bit-identical by construction. Roofline: HBM, 2 * n * word bytes per column.
template <typename R, bool COLMAJ>
__global__ void __launch_bounds__(256)
k_shift(const R* __restrict__ in, R* __restrict__ out, int64_t ncol, int n, int64_t pitch_in, int64_t pitch_out, int shift,
        R fill) {
  const int64_t total = ncol * n;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (int64_t)gridDim.x * 256) {
    int64_t c;
    int k;
    if (COLMAJ) {  // lanes along the column: element (c, k) at c*pitch + k
      c = i / n;
      k = (int)(i - c * n);
    } else {  // lanes along the batch: element (c, k) at k*pitch + c
      k = (int)(i / ncol);
      c = i - (int64_t)k * ncol;
    }
    const int ks = k - shift;
    const bool ok = ks >= 0 && ks < n;
    const size_t src = COLMAJ ? (size_t)c * pitch_in + (ok ? ks : 0) : (size_t)(ok ? ks : 0) * pitch_in + c;
    const size_t dst = COLMAJ ? (size_t)c * pitch_out + k : (size_t)k * pitch_out + c;
    const R v = ld_stream(in + src);
    st_stream(out + dst, ok ? v : fill);
  }
}

template <typename R>
int shift(const R* in, R* out, int64_t ncol, int n, int64_t pitch_in, int64_t pitch_out, int col_major, int s, R fill,
          cudaStream_t st) {
  if (ncol == 0 || n == 0) return EKAT_B200_OK;
  const int64_t total = ncol * n;
  const unsigned blocks = (unsigned)tmin<int64_t>((total + 255) / 256, (int64_t)kNumSms * 8 * 16);
  if (col_major) k_shift<R, true><<<blocks, 256, 0, st>>>(in, out, ncol, n, pitch_in, pitch_out, s, fill);
  else k_shift<R, false><<<blocks, 256, 0, st>>>(in, out, ncol, n, pitch_in, pitch_out, s, fill);
  note_launch(sizeof(R) == 8 ? "shift<f64>" : "shift<f32>");
  return check_cuda(cudaGetLastError());
}

// ------------------------------------------------------------------------------------------------------------
// ekat::index_and_shift<Shift>(a, idx, index, index_shift) for every index pack of every column (reference
// src/pack/ekat_pack_kokkos.hpp:71-98; the gather pair y1(k1), y1(k1+1) of a column remap):
//   out[c][j] = a[c][idx[c][j]],   out_shift[c][j] = a[c][idx[c][j] + shift]
// Where idx or idx + shift falls outside [0, n) the reference's release build reads out of bounds; its debug build
// leaves `index` as the caller had it and sets `index_shift` to ekat::invalid (a quiet NaN, :85-93) -- that defined
// behaviour is the one kept here. Gathered copies: bit-identical by construction.
// Roofline: HBM, (n + m) * word + m * 4 bytes read and 2 * m * word bytes written per column (a is read once: the
// gathers of a column hit L1/L2).
template <typename R, bool COLMAJ>
__global__ void __launch_bounds__(256)
k_index_and_shift(const R* __restrict__ a, const int* __restrict__ idx, R* __restrict__ out, R* __restrict__ out_shift,
                  int64_t ncol, int n, int m, int64_t pitch_a, int64_t pitch_idx, int64_t pitch_out, int shift) {
  const int64_t total = ncol * m;
  const R invalid = sizeof(R) == 8 ? (R)__longlong_as_double(0x7ff8000000000000ll) : (R)__int_as_float(0x7fc00000);
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (int64_t)gridDim.x * 256) {
    int64_t c;
    int j;
    if (COLMAJ) {
      c = i / m;
      j = (int)(i - c * m);
    } else {
      j = (int)(i / ncol);
      c = i - (int64_t)j * ncol;
    }
    const size_t io = COLMAJ ? (size_t)c * pitch_idx + j : (size_t)j * pitch_idx + c;
    const size_t oo = COLMAJ ? (size_t)c * pitch_out + j : (size_t)j * pitch_out + c;
    const int k = __ldg(idx + io), ks = k + shift;
    if (k >= 0 && k < n && ks >= 0 && ks < n) {
      const size_t a0 = COLMAJ ? (size_t)c * pitch_a + k : (size_t)k * pitch_a + c;
      const size_t a1 = COLMAJ ? (size_t)c * pitch_a + ks : (size_t)ks * pitch_a + c;
      out[oo] = __ldg(a + a0);
      out_shift[oo] = __ldg(a + a1);
    } else {
      out_shift[oo] = invalid;
    }
  }
}

template <typename R>
int index_and_shift(const R* a, const int* idx, R* out, R* out_shift, int64_t ncol, int n, int m, int64_t pitch_a,
                    int64_t pitch_idx, int64_t pitch_out, int col_major, int s, cudaStream_t st) {
  if (ncol == 0 || m == 0) return EKAT_B200_OK;
  const int64_t total = ncol * m;
  const unsigned blocks = (unsigned)tmin<int64_t>((total + 255) / 256, (int64_t)kNumSms * 8 * 16);
  if (col_major) k_index_and_shift<R, true><<<blocks, 256, 0, st>>>(a, idx, out, out_shift, ncol, n, m, pitch_a, pitch_idx, pitch_out, s);
  else k_index_and_shift<R, false><<<blocks, 256, 0, st>>>(a, idx, out, out_shift, ncol, n, m, pitch_a, pitch_idx, pitch_out, s);
  note_launch(sizeof(R) == 8 ? "index_and_shift<f64>" : "index_and_shift<f32>");
  return check_cuda(cudaGetLastError());
}

template int shift<double>(const double*, double*, int64_t, int, int64_t, int64_t, int, int, double, cudaStream_t);
template int shift<float>(const float*, float*, int64_t, int, int64_t, int64_t, int, int, float, cudaStream_t);
template int index_and_shift<double>(const double*, const int*, double*, double*, int64_t, int, int, int64_t, int64_t, int64_t, int, int, cudaStream_t);
template int index_and_shift<float>(const float*, const int*, float*, float*, int64_t, int, int, int64_t, int64_t, int64_t, int, int, cudaStream_t);
template int adj_diff<double>(const double*, double*, int64_t, int, int64_t, int64_t, int, int, cudaStream_t);
template int adj_diff<float>(const float*, float*, int64_t, int, int64_t, int64_t, int, int, cudaStream_t);

}  // namespace ekb
