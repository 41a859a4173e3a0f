// fused.cu -- "diffuse then remap" in one kernel (BASELINE config 4): shared-A, nrhs-RHS Thomas
// solve followed by LinInterp of every solution onto new levels, level-major operands.
//
// Fuses the two reference loops
//     thomas(team, dl, d, du, X)                      src/algorithm/ekat_tridiag.hpp:323-334
//     li.setup(...); li.lin_interp(...) per tracer     src/algorithm/ekat_lin_interp_impl.hpp:21-87
// so that the nrhs solutions never travel to HBM between them: the back-substituted columns
// stay in shared memory ([rhs][level][column]) and are the y1 operand of the interpolation, and the
// bracket search runs once per target level for all nrhs tracers (the reference's
// setup-once / interpolate-many contract, ekat_lin_interp.hpp:13-16).
// HBM traffic per column: (3 + nrhs) source level-arrays + x1 + x2 in, nrhs target arrays out
// ((3+10+2+10)*128*8 = 25,600 B at nrhs = 10).
//
// Tridiagonal arithmetic: the FMA/reciprocal Thomas of tridiag_levmajor.cu (tolerance-level).
// Interpolation arithmetic: the reference's 4 statements, unfused, in order (:153-156).
#include "tridiag_lm_common.cuh"

namespace ekb {

template <typename R>
__device__ __forceinline__ int bracket_strided(const R* __restrict__ x1, int stride, int km1, R v) {
  int lo = 0, n = km1;
  while (n > 0) {
    const int step = n >> 1, it = lo + step;
    const bool ge = v >= x1[(size_t)it * stride];
    lo = ge ? it + 1 : lo;
    n = ge ? n - step - 1 : step;
  }
  return lo > 0 ? lo - 1 : lo;
}

// Thread roles as in k_tridiag_lm_shared: warp 0 factorises, the others own (column group, rhs row).
template <typename R, int PF, int CPT>
__global__ void k_tridiag_lin_interp(const R* __restrict__ dl, const R* __restrict__ d, const R* __restrict__ du,
                                     const R* __restrict__ B, const R* __restrict__ x1, const R* __restrict__ x2,
                                     R* __restrict__ Y2, R* __restrict__ X, int64_t ncol, int nlev, int km2, int nrhs,
                                     int Tc, int64_t pitch) {
  using V = Vec<R, CPT>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int tid = threadIdx.x, nthr = blockDim.x;
  const bool role_a = tid < 32;
  const int idx = role_a ? tid : tid - 32;
  const int t = role_a ? (tid < Tc ? tid : Tc - 1) : idx % Tc;
  const int ty = role_a ? 0 : idx / Tc, Ty = (nthr - 32) / Tc;
  const size_t plane = (size_t)nlev * Tc;
  V* f0 = reinterpret_cast<V*>(smem_raw);  // m = dl*r, later the staged x1 tile
  V* f1 = f0 + plane;                      // r = 1/pivot
  V* f2 = f1 + plane;                      // c = du*r
  V* sx = f2 + plane;                      // [nrhs][nlev][Tc]: rhs -> y -> solution

  int64_t c0;
  int nlive;
  own_columns<CPT>(((int64_t)blockIdx.x * Tc + t) * CPT, ncol, c0, nlive);

  if (role_a) {
    if (tid < Tc)
      lm_factorize<R, false, false, PF, CPT>(const_cast<R*>(dl) + c0, const_cast<R*>(d) + c0, const_cast<R*>(du) + c0, pitch,
                                             nlev, f0, f1, f2, Tc, t, 0);
  } else if (ty < Ty) {
    for (int j = ty; j < nrhs; j += Ty)
      lm_issue_rhs<R, CPT>(B + (size_t)j * nlev * pitch + c0, pitch, nlev, sx + (size_t)j * plane, Tc, t);
    cp_async_wait<0>();
  }
  __syncthreads();
  if (!role_a && ty < Ty) {
    for (int j = ty; j < nrhs; j += Ty)
      lm_solve<R, false, CPT, true>(f0, f1, f2, sx + (size_t)j * plane, Tc, t, nlev,
                                    X ? X + (size_t)j * nlev * pitch + c0 : nullptr, pitch, nlive);
  }
  __syncthreads();
  if (km2 <= 0) return;

  // ---- remap: planes are plain [level][tcols] scalar arrays from here on ------------------------------
  const int tcols = Tc * CPT;
  const int64_t tile0 = (int64_t)blockIdx.x * tcols;
  R* sx1 = reinterpret_cast<R*>(f0);  // the multiplier plane is dead: stage the source coordinates over it
  for (int e = tid; e < nlev * tcols; e += nthr) {
    const int i = e / tcols, c = e - i * tcols;
    sx1[e] = ld_stream(x1 + (size_t)i * pitch + tmin<int64_t>(tile0 + c, ncol - 1));
  }
  __syncthreads();
  const int rows = nthr / tcols;
  if (tid < rows * tcols) {
    const int c = tid % tcols, r0 = tid / tcols;
    const int64_t col = tile0 + c;
    const bool live = col < ncol;
    const int64_t gcol = live ? col : ncol - 1;
    const R* x1c = sx1 + c;
    const R* ysol = reinterpret_cast<const R*>(sx) + c;
    const size_t splane = (size_t)nlev * tcols;
    for (int k2 = r0; k2 < km2; k2 += rows) {
      const R xt = ld_stream(x2 + (size_t)k2 * pitch + gcol);
      const int k1 = bracket_strided<R>(x1c, tcols, nlev, xt);
      const int k1ph = k1 == nlev - 1 ? k1 - 1 : k1 + 1;
      const R xa = x1c[(size_t)k1 * tcols], xb = x1c[(size_t)k1ph * tcols];
      const R dxt = x_sub(xt, xa), dx = x_sub(xb, xa);
      const R rdx = rcp_refined(dx);  // one reciprocal shared by the nrhs IEEE divisions of this target
      for (int j = 0; j < nrhs; ++j) {
        const R* y1c = ysol + (size_t)j * splane;
        const R ya = y1c[(size_t)k1 * tcols], yb = y1c[(size_t)k1ph * tcols];
        R v = x_sub(yb, ya);
        v = x_mul(v, dxt);
        v = div_given_rcp(v, dx, rdx);
        v = x_add(v, ya);
        if (live) st_stream(Y2 + ((size_t)j * km2 + k2) * pitch + col, v);
      }
    }
  }
}

template <typename R, int CPT>
static int launch_fused(const R* dl, const R* d, const R* du, const R* B, const R* x1, const R* x2, R* Y2, R* X,
                        int64_t ncol, int nlev, int km2, int nrhs, int64_t pitch, cudaStream_t stream) {
  const Tuning tu = tuning();
  int tcols = tu.cols_per_cta > 0 ? tu.cols_per_cta : 16;
  const size_t per_col = (size_t)(3 + nrhs) * nlev * sizeof(R);
  while (tcols > CPT && (size_t)tcols * per_col > kMaxSmemPerCta) tcols /= 2;
  if ((size_t)tcols * per_col > kMaxSmemPerCta || tcols < CPT) return EKAT_B200_ERR_TOO_LARGE;
  if (tcols > 32 * CPT) tcols = 32 * CPT;
  const int Tc = tcols / CPT;
  int Ty = nrhs;
  while (32 + Tc * Ty > 1024) Ty = (Ty + 1) / 2;
  const size_t smem = (size_t)tcols * per_col;
  auto k = k_tridiag_lin_interp<R, 16, CPT>;
  int rc = allow_smem(k, smem);
  if (rc) return rc;
  int nthr = 32 + Tc * Ty;
  if (nthr < 4 * tcols && 4 * tcols <= 1024) nthr = 4 * tcols;  // enough thread rows for the remap phase
  nthr = (nthr + 31) / 32 * 32;
  k<<<(unsigned)((ncol + tcols - 1) / tcols), nthr, smem, stream>>>(dl, d, du, B, x1, x2, Y2, X, ncol, nlev, km2, nrhs, Tc,
                                                                    pitch);
  note_launch(sizeof(R) == 8 ? "tridiag_lin_interp<f64>" : "tridiag_lin_interp<f32>");
  return check_cuda(cudaGetLastError());
}

template <typename R>
int tridiag_tma_shared_oop(bool exact, R* dl, R* d, R* du, const R* B, R* X, int64_t ncol, int nlev, int nrhs,
                           int64_t a_stride, int64_t x_stride, cudaStream_t stream);
template <typename R>
int lin_interp_lm(const R* x1, const R* x2, const R* y1, R* y2, int64_t ncol, int km1, int km2, int64_t pitch, int nfields,
                  cudaStream_t stream);

template <typename R>
int solve_remap_tma(const R* dl, const R* d, const R* du, const R* Bm, const R* x1, const R* x2, R* Y2, R* X, int64_t ncol,
                    int nlev, int km2, int nrhs, int64_t pitch, cudaStream_t stream);

template <typename R>
int tridiag_lin_interp(const R* dl, const R* d, const R* du, const R* B, const R* x1, const R* x2, R* Y2, R* X,
                       int64_t ncol, int nlev, int km2, int nrhs, int64_t pitch, cudaStream_t stream) {
  auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  // One kernel, every plane of a 16-column tile on chip (fused_tma.cu), whenever the tile fits: nothing but the inputs
  // and the remapped fields crosses HBM. Tuning variant kVarScalarCopies skips it (A/B against the two-pass path below).
  if (tuning().variant == kVarAuto) {
    const int rc = solve_remap_tma<R>(dl, d, du, B, x1, x2, Y2, X, ncol, nlev, km2, nrhs, pitch, stream);
    if (rc != EKAT_B200_ERR_UNSUPPORTED) return rc;
  }
  // Fast path: the persistent TMA solve (B -> X, out of place) followed by the level-major interpolation (X -> Y2),
  // both at their own HBM roofline. The solutions make one round trip through HBM/L2 between the two; when the
  // caller does not ask for them (X == nullptr) they are staged in Y2 itself, which has their shape when
  // km2 == nlev, and interpolated in place tile by tile. The one-kernel version below keeps them on chip but
  // holds only 16 columns per SM and runs at a fifth of this speed (profiles/).
  if ((tuning().variant == kVarAuto || tuning().variant == kVarScalarCopies) && km2 > 0 && (X != nullptr || km2 == nlev)) {
    R* sol = X ? X : Y2;
    int rc = tridiag_tma_shared_oop<R>(false, const_cast<R*>(dl), const_cast<R*>(d), const_cast<R*>(du), B, sol, ncol, nlev,
                                       nrhs, pitch, pitch, stream);
    if (rc == EKAT_B200_OK) rc = lin_interp_lm<R>(x1, x2, sol, Y2, ncol, nlev, km2, pitch, nrhs, stream);
    if (rc != EKAT_B200_ERR_UNSUPPORTED) return rc;
  }
  constexpr int VEC = 16 / sizeof(R);
  const bool vec = al(dl) && al(d) && al(du) && al(B) && (X == nullptr || al(X)) && (pitch * sizeof(R)) % 16 == 0;
  return vec ? launch_fused<R, VEC>(dl, d, du, B, x1, x2, Y2, X, ncol, nlev, km2, nrhs, pitch, stream)
             : launch_fused<R, 1>(dl, d, du, B, x1, x2, Y2, X, ncol, nlev, km2, nrhs, pitch, stream);
}

template int tridiag_lin_interp<double>(const double*, const double*, const double*, const double*, const double*,
                                        const double*, double*, double*, int64_t, int, int, int, int64_t, cudaStream_t);
template int tridiag_lin_interp<float>(const float*, const float*, const float*, const float*, const float*,
                                       const float*, float*, float*, int64_t, int, int, int, int64_t, cudaStream_t);

}  // namespace ekb
