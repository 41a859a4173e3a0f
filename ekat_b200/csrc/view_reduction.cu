// view_reduction.cu -- ekat::ReductionUtils::view_reduction for every column of a batch.
//
// Replaces the per-column loop around ekat::ReductionUtils<ExeSpace,Serialize>::view_reduction(team, begin, end, view)
// (reference src/algorithm/ekat_reduction_utils.hpp:67-144, :156-179): the vertical sum over the scalar range
// [begin, end) of one column, in one of the two orders the reference defines:
//   packn == 0  Serialize = true (ekatBFB builds): one element after the other, in level order (:33-37, :98-103);
//   packn >= 1  Serialize = false over Pack<Real,packn>, as a team evaluates it lane by lane: the leading partial pack
//               element by element (:113-119), the full packs summed per pack lane in pack order from the identity
//               (:123-131), the lane sums added in lane order (reduce_sum<false>, src/pack/ekat_pack_math.hpp:55-83),
//               then the trailing partial pack (:135-141). The code keeps that structure literally.
// Both orders are fixed, so results are bit-identical to the reference (host build, 1-thread teams) for any
// launch geometry. Pack padding (scalars >= end, NaN in the reference's tests) is never read as data.
//
// Roofline: HBM. Algorithmic bytes = (end - begin) * word per column (the [ncol] result is noise).
//   LEV_MAJOR in[k * pitch + c]: one thread per column, a level step of a warp is one 256-byte row segment; eight
//     levels are loaded before the first dependent add so that eight rows are in flight per warp.
//   COL_MAJOR in[c * pitch + k] (the reference's [ncol][npack*N] rows): a tile of 32 columns is staged in shared
//     memory by 8/4-byte cp.async (coalesced along the column, rows padded to an odd number of words), two tiles
//     per CTA in flight; then lane l sums column l from shared memory in the required order, conflict-free.
#include "common.cuh"

namespace ekb {

// The reference's order over one column held as x(k), k a scalar index. PACKED = false: Serialize = true.
template <typename R, bool PACKED, class F>
__device__ __forceinline__ R reduce_in_order(const F& x, int begin, int end, int N) {
  R result = (R)0;
  if (!PACKED) {
    for (int k = begin; k < end; ++k) result = x_add(result, x(k));
    return result;
  }
  const bool gb = begin % N != 0, ge = end % N != 0;
  const int plb = gb ? begin / N + 1 : begin / N, ple = end / N;
  if (gb)
    for (int j = begin % N; j < N; ++j) result = x_add(result, x((plb - 1) * N + j));
  if (plb != ple) {
    // lane sums, then the sum over lanes: evaluated lane by lane so that no per-lane array is needed
    R s = (R)0;
    for (int j = 0; j < N; ++j) {
      R lane = (R)0;
      for (int k = plb; k < ple; ++k) lane = x_add(lane, x(k * N + j));
      s = x_add(s, lane);
    }
    result = x_add(result, s);
  }
  if (ge)
    for (int j = 0; j < end % N; ++j) result = x_add(result, x(ple * N + j));
  return result;
}

// ------------------------------------------------------------------------------------------- level-major
template <typename R, bool PACKED>
__global__ void __launch_bounds__(128)
k_view_reduction_lm(const R* __restrict__ in, R* __restrict__ out, int64_t ncol, int begin, int end, int64_t pitch, int N) {
  const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= ncol) return;
  const R* col = in + c;
  if (!PACKED) {
    R result = (R)0;
    int k = begin;
    for (; k + 8 <= end; k += 8) {
      R v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = ld_stream(col + (size_t)(k + u) * pitch);
#pragma unroll
      for (int u = 0; u < 8; ++u) result = x_add(result, v[u]);
    }
    for (; k < end; ++k) result = x_add(result, ld_stream(col + (size_t)k * pitch));
    out[c] = result;
  } else {
    // pack lanes are levels here: keep up to 16 lane sums in registers when the pack allows it, else the generic order
    if (N <= 16 && begin % N == 0 && end % N == 0 && begin < end) {
      R lane[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) lane[j] = (R)0;
      for (int k = begin; k < end; k += N) {
#pragma unroll
        for (int j = 0; j < 16; ++j)
          if (j < N) lane[j] = x_add(lane[j], ld_stream(col + (size_t)(k + j) * pitch));
      }
      R s = (R)0;
#pragma unroll
      for (int j = 0; j < 16; ++j)
        if (j < N) s = x_add(s, lane[j]);
      out[c] = x_add((R)0, s);
    } else {
      out[c] = reduce_in_order<R, true>([&](int k) { return ld_stream(col + (size_t)k * pitch); }, begin, end, N);
    }
  }
}

// ------------------------------------------------------------------------------------------- column-major
// CTA = one warp with a pair of tile buffers (three CTAs share an SM at 128 f64 levels); tile = 32 columns x n levels,
// rows padded to P words (P odd). Persistent: CTA b walks tiles b, b + gridDim.x, ...
template <typename R, bool PACKED>
__global__ void __launch_bounds__(32)
k_view_reduction_cm(const R* __restrict__ in, R* __restrict__ out, int64_t ncol, int begin, int end, int64_t pitch, int N,
                    int lo, int hi, int P, int64_t ntiles) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x;
  const int n = hi - lo;  // staged scalar range [lo, hi) covers everything the order touches
  R* buf = reinterpret_cast<R*>(smem_raw);
  const int64_t stride = (int64_t)gridDim.x;
  auto stage = [&](int64_t tile, int b) {
    R* dst = buf + (size_t)b * 32 * P;
    const int64_t c0 = tile * 32;
    for (int r = 0; r < 32; ++r) {
      const int64_t c = c0 + r;
      if (c < ncol) {
        const R* src = in + (size_t)c * pitch + lo;
        for (int k = lane; k < n; k += 32) cp_async<sizeof(R)>(dst + r * P + k, src + k);
      }
    }
    cp_async_commit();
  };
  int64_t tile = (int64_t)blockIdx.x;
  int b = 0;
  if (tile < ntiles) stage(tile, 0);
  for (; tile < ntiles; tile += stride, b ^= 1) {
    if (tile + stride < ntiles) {
      stage(tile + stride, b ^ 1);
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncwarp();
    const R* row = buf + (size_t)b * 32 * P + lane * P - lo;  // row[k] = element k of column (tile*32 + lane)
    const int64_t c = tile * 32 + lane;
    if (c < ncol) out[c] = reduce_in_order<R, PACKED>([&](int k) { return row[k]; }, begin, end, N);
    __syncwarp();  // every lane is done with the buffer before it is staged again
  }
}

// Aligned rows (the usual case: pitch and range even): the tile is staged by one 1D bulk copy (TMA) per column,
// issued by the lane that owns the column, completion counted in bytes on the buffer's mbarrier. Rows are padded by
// 16 bytes (bulk copies need 16-byte aligned destinations), which leaves a 2-way bank conflict on the column reads
// (f32: 4-way) -- a few hundred cycles per tile against the ~1,500 its bytes cost at the HBM roofline. The reading
// loop keeps eight loads in flight ahead of the dependent adds; the packed order keeps its NP lane sums in registers.
// NP = 0: Serialize = true; NP = 1, 2, 4, 8, 16: Pack<Real,NP>.
template <typename R, int NP>
__global__ void __launch_bounds__(32)
k_view_reduction_cm_tma(const R* __restrict__ in, R* __restrict__ out, int64_t ncol, int begin, int end, int64_t pitch,
                        int lo, int hi, int P, int64_t ntiles) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int lane = threadIdx.x;
  const int n = hi - lo;
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem_raw);  // [2]
  R* buf = reinterpret_cast<R*>(smem_raw + 16);
  const int64_t stride = (int64_t)gridDim.x;
  if (lane == 0) {
    mbar_init(&bar[0], 1);
    mbar_init(&bar[1], 1);
    fence_mbar_init();
  }
  __syncwarp();
  const unsigned row_bytes = (unsigned)n * sizeof(R);
  auto stage = [&](int64_t tile, int b) {
    const int64_t c0 = tile * 32;
    const int rows = (int)tmin<int64_t>(32, ncol - c0);
    if (lane == 0) mbar_expect_tx(&bar[b], (unsigned)rows * row_bytes);
    if (lane < rows) tma_load_1d(buf + ((size_t)b * 32 + lane) * P, in + (size_t)(c0 + lane) * pitch + lo, row_bytes, &bar[b]);
  };
  int64_t tile = (int64_t)blockIdx.x;
  int b = 0;
  unsigned ph = 0;  // phase bit per buffer
  if (tile < ntiles) stage(tile, 0);
  for (; tile < ntiles; tile += stride, b ^= 1) {
    if (tile + stride < ntiles) stage(tile + stride, b ^ 1);
    mbar_wait(&bar[b], (ph >> b) & 1u);
    ph ^= 1u << b;
    const R* row = buf + ((size_t)b * 32 + lane) * P - lo;  // row[k] = element k of column (tile*32 + lane)
    const int64_t c = tile * 32 + lane;
    if (c < ncol) {
      R result = (R)0;
      if (NP == 0) {
        int k = begin;
        for (; k + 8 <= end; k += 8) {
          R v[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) v[u] = row[k + u];
#pragma unroll
          for (int u = 0; u < 8; ++u) result = x_add(result, v[u]);
        }
        for (; k < end; ++k) result = x_add(result, row[k]);
      } else {
        constexpr int N = NP ? NP : 1;
        const bool gb = begin % N != 0, ge = end % N != 0;
        const int plb = gb ? begin / N + 1 : begin / N, ple = end / N;
        if (gb)
          for (int j = begin % N; j < N; ++j) result = x_add(result, row[(plb - 1) * N + j]);
        if (plb != ple) {
          R acc[N];
#pragma unroll
          for (int j = 0; j < N; ++j) acc[j] = (R)0;
          constexpr int KU = N >= 8 ? 1 : 8 / N;  // packs per unrolled step: at least eight loads in flight
          int k = plb;
          for (; k + KU <= ple; k += KU) {
            R v[KU * N];
#pragma unroll
            for (int u = 0; u < KU * N; ++u) v[u] = row[k * N + u];
#pragma unroll
            for (int u = 0; u < KU * N; ++u) acc[u % N] = x_add(acc[u % N], v[u]);
          }
          for (; k < ple; ++k)
#pragma unroll
            for (int j = 0; j < N; ++j) acc[j] = x_add(acc[j], row[k * N + j]);
          R sum = (R)0;
#pragma unroll
          for (int j = 0; j < N; ++j) sum = x_add(sum, acc[j]);
          result = x_add(result, sum);
        }
        if (ge)
          for (int j = 0; j < end % N; ++j) result = x_add(result, row[ple * N + j]);
      }
      out[c] = result;
    }
    __syncwarp();  // every lane is done with the buffer before it is staged again
  }
}

// Fallback for columns too long to stage: one thread per column straight from global memory (uncoalesced, correct).
template <typename R, bool PACKED>
__global__ void k_view_reduction_cm_direct(const R* __restrict__ in, R* __restrict__ out, int64_t ncol, int begin, int end,
                                           int64_t pitch, int N) {
  const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= ncol) return;
  const R* row = in + (size_t)c * pitch;
  out[c] = reduce_in_order<R, PACKED>([&](int k) { return row[k]; }, begin, end, N);
}

// ------------------------------------------------------------------------------------------- host side
template <typename R>
int view_reduction(const R* in, R* out, int64_t ncol, int begin, int end, int64_t pitch, int col_major, int packn,
                   cudaStream_t st) {
  if (ncol == 0) return EKAT_B200_OK;
  const bool packed = packn >= 1;
  if (!col_major) {
    const unsigned blocks = (unsigned)((ncol + 127) / 128);
    if (packed) k_view_reduction_lm<R, true><<<blocks, 128, 0, st>>>(in, out, ncol, begin, end, pitch, packn);
    else k_view_reduction_lm<R, false><<<blocks, 128, 0, st>>>(in, out, ncol, begin, end, pitch, 1);
    note_launch(sizeof(R) == 8 ? "view_reduction_lm<f64>" : "view_reduction_lm<f32>");
    return check_cuda(cudaGetLastError());
  }
  // scalar range the order reads: [begin, end), widened to whole packs where the packed order's first/last
  // partial pack reads beyond it (it does when begin and end share a pack, :113-119 / :135-141)
  int lo = begin, hi = end;
  if (packed) {
    if (begin % packn) hi = tmax(hi, (begin / packn + 1) * packn);
    if (end % packn) lo = tmin(lo, (end / packn) * packn);
  }
  const int64_t ntiles = (ncol + 31) / 32;
  {
    // TMA path: 16-byte aligned rows and an even-sized staged range (widened inside the row, which is pitch long)
    constexpr int E = 16 / sizeof(R);
    const int tlo = lo / E * E, thi = (int)tmin<int64_t>(pitch, (int64_t)(hi + E - 1) / E * E);
    const int tn = thi - tlo, TP = tn + E;
    const size_t tsmem = 16 + (size_t)2 * 32 * TP * sizeof(R);
    const bool ok = (reinterpret_cast<uintptr_t>(in) & 15) == 0 && pitch % E == 0 && tn % E == 0 && tn > 0 &&
                    tsmem <= kMaxSmemPerCta && (packn == 0 || packn == 1 || packn == 2 || packn == 4 || packn == 8 || packn == 16) &&
                    tuning().variant == kVarAuto;
    if (ok) {
      const int per_sm = (int)tmax<size_t>(1, tmin<size_t>(16, (kMaxSmemPerCta + 1024) / (tsmem + 1024)));
      const unsigned grid = (unsigned)tmin<int64_t>(ntiles, (int64_t)148 * per_sm);
      int rc = EKAT_B200_OK;
#define EKB_VR(NP_)                                                                                            \
  {                                                                                                            \
    auto k = k_view_reduction_cm_tma<R, NP_>;                                                                  \
    rc = allow_smem(k, tsmem);                                                                                 \
    if (!rc) k<<<grid, 32, tsmem, st>>>(in, out, ncol, begin, end, pitch, tlo, thi, TP, ntiles);               \
  }
      switch (packn) {
        case 0: EKB_VR(0) break;
        case 1: EKB_VR(1) break;
        case 2: EKB_VR(2) break;
        case 4: EKB_VR(4) break;
        case 8: EKB_VR(8) break;
        default: EKB_VR(16) break;
      }
#undef EKB_VR
      if (rc) return rc;
      note_launch(sizeof(R) == 8 ? "view_reduction_cm_tma<f64>" : "view_reduction_cm_tma<f32>");
      return check_cuda(cudaGetLastError());
    }
  }
  const int n = hi - lo;
  const int P = n | 1;
  const size_t smem = (size_t)2 * 32 * P * sizeof(R);
  const bool aligned = (reinterpret_cast<uintptr_t>(in) % sizeof(R)) == 0;
  if (n > 0 && smem <= kMaxSmemPerCta && aligned) {
    auto k = packed ? k_view_reduction_cm<R, true> : k_view_reduction_cm<R, false>;
    int rc = allow_smem(k, smem);
    if (rc) return rc;
    const int per_sm = (int)tmax<size_t>(1, tmin<size_t>(16, (kMaxSmemPerCta + 1024) / (smem + 1024)));
    const unsigned grid = (unsigned)tmin<int64_t>(ntiles, (int64_t)148 * per_sm);
    k<<<grid, 32, smem, st>>>(in, out, ncol, begin, end, pitch, packed ? packn : 1, lo, hi, P, ntiles);
    note_launch(sizeof(R) == 8 ? "view_reduction_cm<f64>" : "view_reduction_cm<f32>");
    return check_cuda(cudaGetLastError());
  }
  const unsigned blocks = (unsigned)((ncol + 127) / 128);
  if (packed) k_view_reduction_cm_direct<R, true><<<blocks, 128, 0, st>>>(in, out, ncol, begin, end, pitch, packn);
  else k_view_reduction_cm_direct<R, false><<<blocks, 128, 0, st>>>(in, out, ncol, begin, end, pitch, 1);
  note_launch(sizeof(R) == 8 ? "view_reduction_cm_direct<f64>" : "view_reduction_cm_direct<f32>");
  return check_cuda(cudaGetLastError());
}

template int view_reduction<double>(const double*, double*, int64_t, int, int, int64_t, int, int, cudaStream_t);
template int view_reduction<float>(const float*, float*, int64_t, int, int, int64_t, int, int, cudaStream_t);

}  // namespace ekb
