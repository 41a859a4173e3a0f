// util.cu -- synthetic-input fills, layout marshalling and verification reductions.
//
//  * fills: bit-identical to the reference's deterministic test fills
//    (tests/algorithm/tridiag_tests.hpp:11-28 fill_tridiag_matrix, :30-40 fill_data_matrix,
//    driven per problem as tests/algorithm/tridiag_tests_performance.cpp:190-196) and to the
//    counter-based LinInterp coordinates of SURVEY.md section 8d (ekat_b200/synth.py is the
//    numpy statement of the same generator).
//  * transposes: [ncol][n] <-> [n][pitch].
//  * checksum / residual: device-scalar reductions feeding one NCCL all-reduce (section 8e);
//    residual = the reference's acceptance norm (tests/algorithm/tridiag_tests.hpp:42-106).
#include "common.cuh"

namespace ekb {

// ----------------------------------------------------------------------------- fills
template <typename R>
__global__ void k_fill_tridiag(R* __restrict__ dl, R* __restrict__ d, R* __restrict__ du, R* __restrict__ X,
                               int64_t ncol, int64_t col0, int nlev, int nA, int nrhs, int64_t a_stride,
                               int64_t x_stride, int lev_major, int seed_mod) {
  // one thread per (level, column); loops over the nA matrices / nrhs right-hand sides
  const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int i = blockIdx.y;
  if (c >= ncol) return;
  const int seed = (int)(seed_mod > 0 ? (col0 + c) % seed_mod : (col0 + c));
  for (int p = 0; p < nA; ++p) {
    const int k = seed + p + i;
    // dl = (k%5==0 ? -1 : 1) * 1.3 * (0.1 + ((k*k) % 11)), evaluated in double, left to right
    const double vdl = x_mul(x_mul(k % 5 == 0 ? -1.0 : 1.0, 1.3), x_add(0.1, (double)((k * k) % 11)));
    const double vdu = x_mul(x_mul(k % 7 == 0 ? -1.0 : 1.0, 1.7), x_add(0.2, (double)((k * k) % 13)));
    const R rdl = (R)vdl, rdu = (R)vdu;
    // d = (k%3==0 ? -1 : 1) * (0.7 + |dl| + |du| + (k % 17)) with dl, du already rounded to R
    const double s = x_add(x_add(x_add(0.7, fabs((double)rdl)), fabs((double)rdu)), (double)(k % 17));
    const R rd = (R)x_mul(k % 3 == 0 ? -1.0 : 1.0, s);
    const size_t o = lev_major ? ((size_t)p * nlev + i) * a_stride + c : (size_t)c * a_stride + (size_t)i * nA + p;
    dl[o] = rdl;
    du[o] = rdu;
    d[o] = rd;
  }
  if (X) {
    const int seedx = nrhs;  // fill_data_matrix(X, in.nrhs)
    for (int j = 0; j < nrhs; ++j) {
      const int sg = (7 * i + 11 * j + 3 * i * j) % 3 == 0 ? -1 : 1;
      const int m = (17 * (i - 19) + 13 * (j - 11) + 5 * (i - 5) * (j - 7) + seedx) % 47;
      const size_t o = lev_major ? ((size_t)j * nlev + i) * x_stride + c : (size_t)c * x_stride + (size_t)i * nrhs + j;
      X[o] = (R)x_mul(x_mul((double)sg, 1.7), (double)m);
    }
  }
}

__host__ __device__ __forceinline__ uint64_t splitmix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
__device__ __forceinline__ double unit_rand(uint64_t key, unsigned stream, unsigned k) {
  const uint64_t v = splitmix64(key + ((uint64_t)stream << 32) + k);
  return (double)(v >> 11) * (1.0 / 9007199254740992.0);  // [0,1), 53 bits
}

// One thread per column (the cumulative sums are sequential by construction).
//   x1[0] = u, x1[k] = x1[k-1] + (0.1 + 0.9 u)      strictly increasing (lin_interp_test.cpp:182-216)
//   y1[k] = 100 u
//   x2[k] = lo + (hi - lo) * (c_k / W), c_k = sum_{m<=k} (0.05 + u_m), W = c_{km2-1} + (0.05 + u_extra)
//           lo/hi = x1[0] + delta/1000, x1[km1-1] - delta/1000     (lin_interp_test.cpp:386-390)
template <typename R>
__global__ void k_fill_lin_interp(R* __restrict__ x1, R* __restrict__ y1, R* __restrict__ x2, int64_t ncol,
                                  int64_t col0, int km1, int km2, int64_t pitch1, int64_t pitch2, int lev_major,
                                  uint64_t seed) {
  const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= ncol) return;
  const uint64_t key = splitmix64(seed ^ splitmix64((uint64_t)(col0 + c)));
  double acc = 0, first = 0;
  for (int k = 0; k < km1; ++k) {
    const double u = unit_rand(key, 0, k);
    acc = k == 0 ? u : x_add(acc, x_add(0.1, x_mul(0.9, u)));
    if (k == 0) first = acc;
    const size_t o = lev_major ? (size_t)k * pitch1 + c : (size_t)c * pitch1 + k;
    x1[o] = (R)acc;
    if (y1) y1[o] = (R)x_mul(100.0, unit_rand(key, 1, k));
  }
  if (!x2 || km2 <= 0) return;
  const double last = acc, delta = x_sub(last, first);
  const double lo = x_add(first, x_div(delta, 1000.0)), hi = x_sub(last, x_div(delta, 1000.0));
  double W = 0;
  for (int k = 0; k <= km2; ++k) W = x_add(W, x_add(0.05, unit_rand(key, 2, k)));
  double ck = 0;
  for (int k = 0; k < km2; ++k) {
    ck = x_add(ck, x_add(0.05, unit_rand(key, 2, k)));
    const size_t o = lev_major ? (size_t)k * pitch2 + c : (size_t)c * pitch2 + k;
    x2[o] = (R)x_add(lo, x_mul(x_sub(hi, lo), x_div(ck, W)));
  }
}

// ------------------------------------------------------------------------ transposes
// src[c*src_stride + k] -> dst[k*pitch + c], 32x32 tiles through padded shared memory.
template <typename R>
__global__ void k_col_to_lev(const R* __restrict__ src, R* __restrict__ dst, int64_t ncol, int n,
                             int64_t src_stride, int64_t pitch) {
  __shared__ R tile[32][33];
  const int64_t c0 = (int64_t)blockIdx.x * 32;
  const int k0 = blockIdx.y * 32;
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int64_t c = c0 + r;
    const int k = k0 + threadIdx.x;
    if (c < ncol && k < n) tile[r][threadIdx.x] = src[(size_t)c * src_stride + k];
  }
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int k = k0 + r;
    const int64_t c = c0 + threadIdx.x;
    if (c < ncol && k < n) dst[(size_t)k * pitch + c] = tile[threadIdx.x][r];
  }
}
template <typename R>
__global__ void k_lev_to_col(const R* __restrict__ src, R* __restrict__ dst, int64_t ncol, int n, int64_t pitch,
                             int64_t dst_stride) {
  __shared__ R tile[32][33];
  const int64_t c0 = (int64_t)blockIdx.x * 32;
  const int k0 = blockIdx.y * 32;
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int k = k0 + r;
    const int64_t c = c0 + threadIdx.x;
    if (c < ncol && k < n) tile[r][threadIdx.x] = src[(size_t)k * pitch + c];
  }
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int64_t c = c0 + r;
    const int k = k0 + threadIdx.x;
    if (c < ncol && k < n) dst[(size_t)c * dst_stride + k] = tile[threadIdx.x][r];
  }
}

// ekat::transpose<c2f | f2c> for rank-3 arrays (reference src/kokkos/ekat_math_utils.hpp:124-147), on the device and
// with pitches: the C side is [ni][nk][nj] (element (i,k,j) at c[i*c_stride + k*nj + j]; the reference's batch layout
// X[col][lev][rhs]), the Fortran side [nj][nk][ni] (element at f[(j*nk + k)*f_pitch + i]; the native level-major layout
// X[rhs][lev][col]). With q = k*nj + j the C side is a 2-D array [ni][nk*nj], so one 32 x 32 tile through padded
// shared memory serves: rows of q are contiguous on the C side, rows of i on the Fortran side (row (q % nj)*nk + q / nj).
// Roofline: HBM, 2 * word bytes per element.
template <typename R, bool C2F>
__global__ void k_transpose3(const R* __restrict__ src, R* __restrict__ dst, int64_t ni, int nk, int nj, int64_t c_stride,
                             int64_t f_pitch) {
  __shared__ R tile[32][33];
  const int64_t i0 = (int64_t)blockIdx.x * 32;
  const int q0 = blockIdx.y * 32, nq = nk * nj;
  if (C2F) {
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
      const int64_t i = i0 + r;
      const int q = q0 + threadIdx.x;
      if (i < ni && q < nq) tile[r][threadIdx.x] = ld_stream(src + (size_t)i * c_stride + q);
    }
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
      const int q = q0 + r;
      const int64_t i = i0 + threadIdx.x;
      if (i < ni && q < nq) st_stream(dst + ((size_t)(q % nj) * nk + q / nj) * f_pitch + i, tile[threadIdx.x][r]);
    }
  } else {
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
      const int q = q0 + r;
      const int64_t i = i0 + threadIdx.x;
      if (i < ni && q < nq) tile[r][threadIdx.x] = ld_stream(src + ((size_t)(q % nj) * nk + q / nj) * f_pitch + i);
    }
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
      const int64_t i = i0 + r;
      const int q = q0 + threadIdx.x;
      if (i < ni && q < nq) st_stream(dst + (size_t)i * c_stride + q, tile[threadIdx.x][r]);
    }
  }
}

// Row repacking of the host_to_device / device_to_host family (reference src/pack/ekat_pack_kokkos.hpp:332-1035): a
// rank-2 scalar array [nrows][n] (row stride src_stride) becomes the scalarized storage of a view of Pack<R,N> rows,
// [nrows][npack*N] (row stride dst_pitch >= n), the padding lanes zero as a freshly allocated Kokkos view has them
// (zero_pad), or left alone; with dst_pitch == n it is the reverse direction (device_to_host drops the padding).
template <typename R>
__global__ void __launch_bounds__(256)
k_repack(const R* __restrict__ src, R* __restrict__ dst, int64_t nrows, int n, int64_t src_stride, int64_t dst_pitch, int width) {
  const int64_t total = nrows * width;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (int64_t)gridDim.x * 256) {
    const int64_t r = i / width;
    const int k = (int)(i - r * width);
    st_stream(dst + (size_t)r * dst_pitch + k, k < n ? ld_stream(src + (size_t)r * src_stride + k) : (R)0);
  }
}

// ------------------------------------------------------------------------ reductions
template <typename R> struct Bits;
template <> struct Bits<double> { static __device__ uint64_t of(double v) { return (uint64_t)__double_as_longlong(v); } };
template <> struct Bits<float> { static __device__ uint64_t of(float v) { return (uint64_t)__float_as_uint(v); } };

template <typename R>
__global__ void k_checksum(const R* __restrict__ a, int64_t nrows, int64_t ncol, int64_t pitch,
                           unsigned long long* out) {
  uint64_t s = 0;
  const int64_t total = nrows * ncol;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = idx / ncol, c = idx - r * ncol;
    s += Bits<R>::of(a[(size_t)r * pitch + c]);
  }
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  __shared__ uint64_t ws[32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) ws[w] = s;
  __syncthreads();
  if (w == 0) {
    s = lane < (blockDim.x + 31) / 32 ? ws[lane] : 0;
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) atomicAdd(out, (unsigned long long)s);
  }
}

// thread per column; y = A x evaluated in R with unfused ops in the reference's order
// (tridiag_tests.hpp:62-85), |b - y| and |b| maxima in double (:96-105).
template <typename R>
__global__ void k_residual(const R* __restrict__ dl, const R* __restrict__ d, const R* __restrict__ du,
                           const R* __restrict__ B, const R* __restrict__ X, int64_t ncol, int nlev, int nA,
                           int nrhs, int64_t pitch, unsigned long long* out) {
  const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  double worst = 0;
  if (c < ncol) {
    double num = 0, den = 0;
    for (int j = 0; j < nrhs; ++j) {
      const size_t am = (size_t)(nA > 1 ? j : 0) * nlev, xm = (size_t)j * nlev;
      R xm1 = 0, x0 = X[xm * pitch + c];
      for (int i = 0; i < nlev; ++i) {
        const R xp1 = i + 1 < nlev ? X[(xm + i + 1) * pitch + c] : (R)0;
        const size_t a = (am + i) * pitch + c;
        R y;
        if (nlev == 1) y = x_mul(d[a], x0);
        else if (i == 0) y = x_add(x_mul(d[a], x0), x_mul(du[a], xp1));
        else if (i == nlev - 1) y = x_add(x_mul(dl[a], xm1), x_mul(d[a], x0));
        else y = x_add(x_add(x_mul(dl[a], xm1), x_mul(d[a], x0)), x_mul(du[a], xp1));
        const double bb = (double)B[(xm + i) * pitch + c];
        if (!isfinite((double)y)) { num = __longlong_as_double(0x7ff0000000000000LL); den = 1; }
        num = fmax(num, fabs(bb - (double)y));
        den = fmax(den, fabs(bb));
        xm1 = x0;
        x0 = xp1;
      }
    }
    worst = den > 0 ? num / den : num;
  }
  for (int o = 16; o > 0; o >>= 1) worst = fmax(worst, __shfl_xor_sync(0xffffffffu, worst, o));
  if ((threadIdx.x & 31) == 0) atomicMax(out, (unsigned long long)__double_as_longlong(worst));  // >= 0: bit order == value order
}

// ------------------------------------------------------------------------ self test
// x_div_pipelined (reciprocal hoisted, common.cuh) against the compiler's IEEE division on random operands:
// mode 0 = random bit patterns (every exponent, denormals, infinities, NaNs excluded from the count when both
// results are NaN), mode 1 = operands in the well-scaled range the solvers see. Counts bitwise mismatches.
// modes 2 / 3: the same two operand populations through the certified quotient (see below); 4 / 5: its flag rate.
template <typename R> struct DivRef;
template <> struct DivRef<double> {
  using U = unsigned long long;
  static __device__ double make(uint64_t b) { return __longlong_as_double((long long)b); }
  static __device__ U bits(double v) { return (U)__double_as_longlong(v); }
  static __device__ double ref(double a, double b) { return __ddiv_rn(a, b); }
};
template <> struct DivRef<float> {
  using U = unsigned;
  static __device__ float make(uint64_t b) { return __uint_as_float((unsigned)(b >> 16)); }
  static __device__ U bits(float v) { return __float_as_uint(v); }
  static __device__ float ref(float a, float b) { return __fdiv_rn(a, b); }
};
template <typename R>
__global__ void k_selftest_div(uint64_t n, uint64_t seed, int mode, unsigned long long* mismatches) {
  unsigned long long bad = 0;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t ra = splitmix64(seed + 2 * i), rb = splitmix64(seed + 2 * i + 1);
    R a, b;
    if ((mode & 1) == 0) {
      a = DivRef<R>::make(ra);
      b = DivRef<R>::make(rb);
    } else {
      a = (R)(((double)(ra >> 11) * (1.0 / 9007199254740992.0) - 0.5) * 200.0);
      b = (R)(((double)(rb >> 11) * (1.0 / 9007199254740992.0) + 0.01) * ((rb & 1) ? 60.0 : -60.0));
    }
    if (mode >= 2) {
      // modes 2, 3: the certified quotient of the bfb factorisation (common.cuh bfb_pivot_step) with a reciprocal taken
      // of a PERTURBED divisor (the predicted pivot: a few ulp to 2^-30 away) must be right whenever it is not flagged;
      // modes 4, 5: count the flags instead (how often the full-range redo would run)
      const uint64_t rp = splitmix64(seed ^ (0x9e3779b97f4a7c15ull * (i + 1)));
      const int sh = mode >= 6 ? 1 : (int)(rp & 31);  // perturbation 2^-53 .. 2^-22 relative (modes 6, 7: 2^-52, flag counts)
      const R bp = b * ((R)1 + (R)((rp >> 8 & 1) ? 1 : -1) * (R)ldexp(1.0, -(sizeof(R) == 8 ? 53 : 24) + (sizeof(R) == 8 ? sh : sh % 12)));
      for (int z = 0; z < 3; ++z) {
        const R an = z == 0 ? a : (z == 1 ? (R)0 : -(R)0);  // also +0 and -0 numerators (the sign of a zero quotient)
        const R r = rcp_seed(bp);
        const R q0 = x_mul(an, r);
        const R m = fma(r, fma(-b, q0, an), q0);
        bool flagged = false;
        certify_quotient(an, b, m, flagged);
        const R ws = DivRef<R>::ref(an, b);
        if (mode >= 4) bad += (flagged && z < 2) ? 1 : 0;  // (-0 numerators are always sent to the full-range division: not counted)
        else if (!flagged && DivRef<R>::bits(m) != DivRef<R>::bits(ws) && !(m != m && ws != ws)) ++bad;
      }
      continue;
    }
    const R q = x_div_pipelined(a, b), w = DivRef<R>::ref(a, b);
    if (DivRef<R>::bits(q) != DivRef<R>::bits(w) && !(q != q && w != w)) ++bad;
    // the speculative form must be right whenever it does not raise its flag; also probe zero numerators
    for (int z = 0; z < 2; ++z) {
      const R an = z ? (R)0 : a;
      bool flagged = false;
      const R qs = div_spec(an, b, rcp_refined(b), flagged);
      const R ws = DivRef<R>::ref(an, b);
      if (!flagged && DivRef<R>::bits(qs) != DivRef<R>::bits(ws) && !(qs != qs && ws != ws)) ++bad;
    }
  }
  if (bad) atomicAdd(mismatches, bad);
}
template <typename R>
int selftest_div(uint64_t n, uint64_t seed, int mode, uint64_t* mismatches, cudaStream_t st) {
  k_selftest_div<R><<<kNumSms * 8, 256, 0, st>>>(n, seed, mode, reinterpret_cast<unsigned long long*>(mismatches));
  note_launch("selftest_div");
  return check_cuda(cudaGetLastError());
}
template int selftest_div<double>(uint64_t, uint64_t, int, uint64_t*, cudaStream_t);
template int selftest_div<float>(uint64_t, uint64_t, int, uint64_t*, cudaStream_t);

// ------------------------------------------------------------------------ host side
template <typename R>
int fill_tridiag(R* dl, R* d, R* du, R* X, int64_t ncol, int64_t col0, int nlev, int nA, int nrhs,
                 int64_t a_stride, int64_t x_stride, int layout, int seed_mod, cudaStream_t st) {
  if (ncol == 0) return EKAT_B200_OK;
  dim3 grid((unsigned)((ncol + 127) / 128), (unsigned)nlev);
  k_fill_tridiag<R><<<grid, 128, 0, st>>>(dl, d, du, X, ncol, col0, nlev, nA, nrhs, a_stride, x_stride,
                                          layout == EKAT_B200_LAYOUT_LEV_MAJOR, seed_mod);
  note_launch("fill_tridiag");
  return check_cuda(cudaGetLastError());
}
template <typename R>
int fill_lin_interp(R* x1, R* y1, R* x2, int64_t ncol, int64_t col0, int km1, int km2, int64_t pitch1,
                    int64_t pitch2, int layout, uint64_t seed, cudaStream_t st) {
  if (ncol == 0) return EKAT_B200_OK;
  k_fill_lin_interp<R><<<(unsigned)((ncol + 127) / 128), 128, 0, st>>>(
      x1, y1, x2, ncol, col0, km1, km2, pitch1, pitch2, layout == EKAT_B200_LAYOUT_LEV_MAJOR, seed);
  note_launch("fill_lin_interp");
  return check_cuda(cudaGetLastError());
}
template <typename R>
int col_to_lev(const R* src, R* dst, int64_t ncol, int n, int64_t src_stride, int64_t pitch, cudaStream_t st) {
  if (ncol == 0 || n == 0) return EKAT_B200_OK;
  dim3 grid((unsigned)((ncol + 31) / 32), (unsigned)((n + 31) / 32)), block(32, 8);
  k_col_to_lev<R><<<grid, block, 0, st>>>(src, dst, ncol, n, src_stride, pitch);
  note_launch("col_to_lev");
  return check_cuda(cudaGetLastError());
}
template <typename R>
int lev_to_col(const R* src, R* dst, int64_t ncol, int n, int64_t pitch, int64_t dst_stride, cudaStream_t st) {
  if (ncol == 0 || n == 0) return EKAT_B200_OK;
  dim3 grid((unsigned)((ncol + 31) / 32), (unsigned)((n + 31) / 32)), block(32, 8);
  k_lev_to_col<R><<<grid, block, 0, st>>>(src, dst, ncol, n, pitch, dst_stride);
  note_launch("lev_to_col");
  return check_cuda(cudaGetLastError());
}
template <typename R>
int transpose3(const R* src, R* dst, int64_t ni, int nk, int nj, int64_t c_stride, int64_t f_pitch, int c2f, cudaStream_t st) {
  if (ni == 0 || nk == 0 || nj == 0) return EKAT_B200_OK;
  const int64_t nq = (int64_t)nk * nj;
  if (nq >= (int64_t)65535 * 32) return EKAT_B200_ERR_TOO_LARGE;
  dim3 grid((unsigned)((ni + 31) / 32), (unsigned)((nq + 31) / 32)), block(32, 8);
  if (c2f) k_transpose3<R, true><<<grid, block, 0, st>>>(src, dst, ni, nk, nj, c_stride, f_pitch);
  else k_transpose3<R, false><<<grid, block, 0, st>>>(src, dst, ni, nk, nj, c_stride, f_pitch);
  note_launch(c2f ? "transpose3_c2f" : "transpose3_f2c");
  return check_cuda(cudaGetLastError());
}
template <typename R>
int repack(const R* src, R* dst, int64_t nrows, int n, int64_t src_stride, int64_t dst_pitch, int zero_pad, cudaStream_t st) {
  if (nrows == 0 || n == 0) return EKAT_B200_OK;
  const int width = zero_pad ? (int)dst_pitch : n;
  const int64_t total = nrows * width;
  const unsigned blocks = (unsigned)tmin<int64_t>((total + 255) / 256, (int64_t)kNumSms * 8 * 16);
  k_repack<R><<<blocks, 256, 0, st>>>(src, dst, nrows, n, src_stride, dst_pitch, width);
  note_launch("repack");
  return check_cuda(cudaGetLastError());
}
template <typename R>
int checksum(const R* a, int64_t nrows, int64_t ncol, int64_t pitch, uint64_t* out, cudaStream_t st) {
  if (nrows * ncol == 0) return EKAT_B200_OK;
  const int64_t total = nrows * ncol;
  const unsigned blocks = (unsigned)tmin<int64_t>((total + 255) / 256, (int64_t)kNumSms * 16);
  k_checksum<R><<<blocks, 256, 0, st>>>(a, nrows, ncol, pitch, reinterpret_cast<unsigned long long*>(out));
  note_launch("checksum");
  return check_cuda(cudaGetLastError());
}
template <typename R>
int residual(const R* dl, const R* d, const R* du, const R* B, const R* X, int64_t ncol, int nlev, int nA,
             int nrhs, int64_t pitch, double* out, cudaStream_t st) {
  if (ncol == 0) return EKAT_B200_OK;
  k_residual<R><<<(unsigned)((ncol + 127) / 128), 128, 0, st>>>(dl, d, du, B, X, ncol, nlev, nA, nrhs, pitch,
                                                             reinterpret_cast<unsigned long long*>(out));
  note_launch("residual");
  return check_cuda(cudaGetLastError());
}

#define INST(R)                                                                                              \
  template int fill_tridiag<R>(R*, R*, R*, R*, int64_t, int64_t, int, int, int, int64_t, int64_t, int, int,   \
                               cudaStream_t);                                                                \
  template int fill_lin_interp<R>(R*, R*, R*, int64_t, int64_t, int, int, int64_t, int64_t, int, uint64_t,   \
                                  cudaStream_t);                                                             \
  template int col_to_lev<R>(const R*, R*, int64_t, int, int64_t, int64_t, cudaStream_t);                    \
  template int lev_to_col<R>(const R*, R*, int64_t, int, int64_t, int64_t, cudaStream_t);                    \
  template int transpose3<R>(const R*, R*, int64_t, int, int, int64_t, int64_t, int, cudaStream_t);          \
  template int repack<R>(const R*, R*, int64_t, int, int64_t, int64_t, int, cudaStream_t);                   \
  template int checksum<R>(const R*, int64_t, int64_t, int64_t, uint64_t*, cudaStream_t);                    \
  template int residual<R>(const R*, const R*, const R*, const R*, const R*, int64_t, int, int, int, int64_t, \
                           double*, cudaStream_t);
INST(double)
INST(float)

}  // namespace ekb
