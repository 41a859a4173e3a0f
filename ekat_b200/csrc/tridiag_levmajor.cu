// tridiag_levmajor.cu -- batched Thomas / bfb solvers on the native level-major layout.
//
// Replaces the per-column loops around ekat::tridiag::thomas and ekat::tridiag::bfb
// (reference: src/algorithm/ekat_tridiag.hpp:177-214 team Thomas, :216-290 serial Thomas,
// :292-320 + :632-678 bfb). One thread owns the recurrences of CPT = 16/sizeof(Real) adjacent
// columns; consecutive threads own consecutive column groups, so every level step of a warp is one
// coalesced 512-byte access on the [level][column] arrays.
//
// HBM traffic is the algorithmic minimum -- read dl,d,du,B once, write X once ((3+2*nrhs) words per
// column-level): the forward sweep's intermediates live in shared memory ([plane][level][thread]
// vectors, conflict-free) instead of being written back and re-read as the reference's in-place
// sweeps do. The kernels are bound by the serial divide chain per level (measured on B200: dependent
// DFMA 8 cycles, IEEE divide 131, fast reciprocal 63 -- tools/microbench.cu), not by issue slots, so
// only ~100 f64 columns of 128 levels fit an SM and each thread must keep many loads in flight:
// operands arrive by cp.async, PF levels ahead, straight into the planes they are updated in
// (tridiag_lm_common.cuh). A register prefetch ring does not work here: a warp has 6 scoreboard slots,
// so waiting for an old load also waits for the newest ones and the full HBM latency is exposed at
// every level (measured: 600 cycles per level, 16% of the HBM roofline).
//
//   EXACT (bfb):   q = dl[i]/d'[i-1]; d'[i] = d[i] - q*du[i-1]; x'[i] = x[i] - q*x'[i-1];
//                  x[n-1] = x'[n-1]/d'[n-1]; x[i-1] = (x'[i-1] - du[i-1]*x[i])/d'[i-1]
//                  every * - / separately rounded: bit-identical to the reference's bfb == thomas at
//                  -ffp-contract=off. The division is IEEE (same operations as __ddiv_rn) with its
//                  reciprocal part hoisted off the back-substitution chain (common.cuh).
//   FAST (thomas): reciprocal form with FMAs, r = 1/(d[i] - dl[i]*c[i-1]); c[i] = du[i]*r;
//                  y[i] = (x[i] - dl[i]*y[i-1])*r;  x[i] = y[i] - c[i]*x[i+1]
//                  -- 1 reciprocal per level instead of 2 divisions, one FMA per level in the
//                  back-substitution chain; agrees to ~1e-15 relative.
#include "tridiag_lm_common.cuh"

namespace ekb {

// TMA-fed warp-per-tile kernel (tridiag_tma.cu); EKAT_B200_ERR_UNSUPPORTED if the operands do not qualify.
template <typename R, bool CM>
int tridiag_bfb_stream(R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nmat, int64_t a_stride, int64_t x_stride,
                       cudaStream_t stream, int clob);  // tridiag_bfb_stream.cu
template <typename R>
int tridiag_bfb_shared(R* dl, R* d, R* du, R* X, const R* Xin, int64_t ncol, int nlev, int nrhs, int64_t a_stride, int64_t x_stride,
                       cudaStream_t stream);  // tridiag_bfb_shared.cu
template <typename R>
int tridiag_tma(bool exact, R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nmat, int64_t a_stride,
                int64_t x_stride, cudaStream_t stream);

// Persistent TMA-fed kernel for one matrix shared by nrhs right-hand sides (tridiag_tma_shared.cu).
template <typename R>
int tridiag_tma_shared(bool exact, R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nrhs, int64_t a_stride,
                       int64_t x_stride, cudaStream_t stream);

// One warp per tile, whole tile on chip, for short columns / small batches (tridiag_small.cu).
template <typename R>
int tridiag_small(R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nmat, int64_t a_stride, int64_t x_stride, bool force,
                  cudaStream_t stream);

// ---------------------------------------------------------------------------------------
// One matrix per rhs (nA == nrhs, incl. nrhs == 1): thread <-> (column group, blockIdx.y = matrix).
// Planes ([nlev][T] vectors):  EXACT: P0 = d -> d', P1 = x -> x', P2 = du;  ring R0 = dl
//                              FAST : P0 = du -> c, P1 = x -> y;            rings R0 = dl, R1 = d
// ---------------------------------------------------------------------------------------
template <typename R, bool EXACT, bool CLOBBER, int PF, int CPT>
__global__ void k_tridiag_lm(R* __restrict__ dl, R* __restrict__ d, R* __restrict__ du, R* __restrict__ X,
                             int64_t ncol, int nlev, int64_t a_stride, int64_t x_stride) {
  using V = Vec<R, CPT>;
  constexpr int B = sizeof(R) * CPT;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int T = blockDim.x, t = threadIdx.x;
  const size_t plane = (size_t)nlev * T;
  V* p0 = reinterpret_cast<V*>(smem_raw);
  V* p1 = p0 + plane;
  V* p2 = p1 + plane;                      // EXACT only
  V* r0 = EXACT ? p2 + plane : p1 + plane;  // [PF][T]
  V* r1 = r0 + (size_t)PF * T;              // FAST only

  int64_t c0;
  int nlive;
  own_columns<CPT>(((int64_t)blockIdx.x * T + t) * CPT, ncol, c0, nlive);
  const size_t mat = (size_t)blockIdx.y * nlev;
  R* pdl = dl + mat * a_stride + c0;
  R* pd = d + mat * a_stride + c0;
  R* pdu = du + mat * a_stride + c0;
  R* px = X + mat * x_stride + c0;

  auto issue = [&](int i, int slot) {
    if (i < nlev) {
      const size_t s = (size_t)i * T + t, ga = (size_t)i * a_stride;
      cp_async<B>(&r0[(size_t)slot * T + t], pdl + ga);
      if (EXACT) {
        cp_async<B>(&p0[s], pd + ga);
        cp_async<B>(&p2[s], pdu + ga);
      } else {
        cp_async<B>(&r1[(size_t)slot * T + t], pd + ga);
        cp_async<B>(&p0[s], pdu + ga);
      }
      cp_async<B>(&p1[s], px + (size_t)i * x_stride);
    }
    cp_async_commit();
  };
#pragma unroll
  for (int i = 0; i < PF; ++i) issue(i, i);
  cp_async_wait<PF - 1>();
  V a = r0[t], b = EXACT ? p0[t] : r1[t], c = EXACT ? p2[t] : p0[t], x = p1[t];
  V dprev, xprev, uprev;
  int slot = 0;
#pragma unroll 2
  for (int i = 0; i < nlev; ++i) {
    V an, bn, cn, xn;
    const int slot_n = slot + 1 == PF ? 0 : slot + 1;
    if (i + 1 < nlev) {
      cp_async_wait<PF - 2>();  // every group but the newest PF-2 has landed => level i+1 is in shared memory
      const size_t s = (size_t)(i + 1) * T + t, rs = (size_t)slot_n * T + t;
      an = r0[rs];
      bn = EXACT ? p0[s] : r1[rs];
      cn = EXACT ? p2[s] : p0[s];
      xn = p1[s];
    }
    issue(i + PF, slot);  // ring slot of level i is free: its operands are in registers
    const size_t s = (size_t)i * T + t;
    V o0, o1;
#pragma unroll
    for (int k = 0; k < CPT; ++k) {
      if (EXACT) {
        R dcur = b.v[k], xcur = x.v[k], q = 0;
        if (i > 0) {
          q = x_div_pipelined(a.v[k], dprev.v[k]);
          dcur = x_sub(b.v[k], x_mul(q, uprev.v[k]));
          xcur = x_sub(x.v[k], x_mul(q, xprev.v[k]));
        }
        o0.v[k] = dcur;
        o1.v[k] = xcur;
        dprev.v[k] = dcur;
        xprev.v[k] = xcur;
        uprev.v[k] = c.v[k];
        a.v[k] = q;  // kept for CLOBBER
      } else {
        const R den = i > 0 ? f_fma(-a.v[k], uprev.v[k], b.v[k]) : b.v[k];
        const R r = f_rcp(den);
        const R cc = c.v[k] * r;
        const R yy = (i > 0 ? f_fma(-a.v[k], xprev.v[k], x.v[k]) : x.v[k]) * r;
        o0.v[k] = cc;
        o1.v[k] = yy;
        uprev.v[k] = cc;
        xprev.v[k] = yy;
        a.v[k] = den;  // kept for CLOBBER
      }
    }
    p0[s] = o0;
    p1[s] = o1;
    if (CLOBBER && i > 0 && nlive > 0) {
      // reference side effect: multipliers in dl + pivots in d (bfb / team thomas, :299-302); FAST: pivots in d (:221)
      if (EXACT) {
        st_vec_stream<R, CPT>(pdl + (size_t)i * a_stride, a, nlive);
        st_vec_stream<R, CPT>(pd + (size_t)i * a_stride, o0, nlive);
      } else {
        st_vec_stream<R, CPT>(pd + (size_t)i * a_stride, a, nlive);
      }
    }
    a = an;
    b = bn;
    c = cn;
    x = xn;
    slot = slot_n;
  }
  cp_async_wait<0>();

  // back substitution straight out of shared memory, solution streamed to HBM
  size_t s = (size_t)(nlev - 1) * T + t;
  V xv = xprev;
  if (EXACT) {
#pragma unroll
    for (int k = 0; k < CPT; ++k) xv.v[k] = x_div_pipelined(xprev.v[k], dprev.v[k]);
  }
  if (nlive > 0) st_vec_stream<R, CPT>(px + (size_t)(nlev - 1) * x_stride, xv, nlive);
#pragma unroll 4
  for (int i = nlev - 2; i >= 0; --i) {
    s = (size_t)i * T + t;
    const V xi = p1[s], m0 = p0[s];
    if (EXACT) {
      const V ui = p2[s];
#pragma unroll
      for (int k = 0; k < CPT; ++k) {
        const R r = rcp_refined(m0.v[k]);  // off the x chain
        xv.v[k] = div_given_rcp(x_sub(xi.v[k], x_mul(ui.v[k], xv.v[k])), m0.v[k], r);
      }
    } else {
#pragma unroll
      for (int k = 0; k < CPT; ++k) xv.v[k] = f_fma(-m0.v[k], xv.v[k], xi.v[k]);
    }
    if (nlive > 0) st_vec_stream<R, CPT>(px + (size_t)i * x_stride, xv, nlive);
  }
}

// ---------------------------------------------------------------------------------------
// One matrix shared by nrhs right-hand sides (nA == 1, nrhs > 1).
// CTA = Tc column groups x (1 + Ty) thread rows: row 0 factorises each column's matrix once into
// shared memory while rows 1..Ty stream their right-hand sides in (one cp.async group per rhs) --
// the structure of the reference (factorize in Kokkos::single, barrier, per-rhs solves:
// ekat_tridiag.hpp:323-334, :632-648) with the barrier wait hidden behind the loads. After the
// barrier every (column group, rhs) recurrence runs in place out of shared memory.
// blockIdx.y = rhs group of G right-hand sides. smem: 3 factor planes + G rhs planes, [nlev][Tc] vectors.
// ---------------------------------------------------------------------------------------
template <typename R, bool EXACT, bool CLOBBER, int PF, int CPT>
__global__ void k_tridiag_lm_shared(R* __restrict__ dl, R* __restrict__ d, R* __restrict__ du,
                                    R* __restrict__ X, int64_t ncol, int nlev, int nrhs, int G, int Tc,
                                    int64_t a_stride, int64_t x_stride) {
  using V = Vec<R, CPT>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  // warp 0 = factorisers (lanes < Tc), threads 32.. = (column group, rhs row) solvers: roles never share a warp
  const int tid = threadIdx.x;
  const bool role_a = tid < 32;
  const int idx = role_a ? tid : tid - 32;
  const int t = role_a ? (tid < Tc ? tid : Tc - 1) : idx % Tc;
  const int ty = role_a ? 0 : idx / Tc, Ty = ((int)blockDim.x - 32) / Tc;
  const size_t plane = (size_t)nlev * Tc;
  V* f0 = reinterpret_cast<V*>(smem_raw);
  V* f1 = f0 + plane;
  V* f2 = f1 + plane;
  V* sx = f2 + plane;

  int64_t c0;
  int nlive;
  own_columns<CPT>(((int64_t)blockIdx.x * Tc + t) * CPT, ncol, c0, nlive);
  // CLOBBER writes the factorisation back into dl/d, so the matrix may be read from HBM only once per column: the
  // CTA then walks every right-hand-side group itself (gridDim.y == 1) instead of one group per blockIdx.y -- with
  // one CTA per group the later groups would factorise the multipliers/pivots the first one had already stored.
  const int ngroups = (nrhs + G - 1) / G;
  const int g_begin = CLOBBER ? 0 : (int)blockIdx.y, g_end = CLOBBER ? ngroups : (int)blockIdx.y + 1;
  for (int g = g_begin; g < g_end; ++g) {
    const int j0 = g * G;
    const int jn = min(G, nrhs - j0);
    if (role_a) {
      if (tid < Tc && g == g_begin)
        lm_factorize<R, EXACT, CLOBBER, PF, CPT>(dl + c0, d + c0, du + c0, a_stride, nlev, f0, f1, f2, Tc, t, nlive);
    } else if (ty < Ty) {
      for (int jj = ty; jj < jn; jj += Ty)
        lm_issue_rhs<R, CPT>(X + (size_t)(j0 + jj) * nlev * x_stride + c0, x_stride, nlev, sx + (size_t)jj * plane, Tc, t);
      cp_async_wait<0>();
    }
    __syncthreads();
    if (!role_a && ty < Ty) {
      for (int jj = ty; jj < jn; jj += Ty)
        lm_solve<R, EXACT, CPT, false>(f0, f1, f2, sx + (size_t)jj * plane, Tc, t, nlev,
                                       X + (size_t)(j0 + jj) * nlev * x_stride + c0, x_stride, nlive);
    }
    if (g + 1 < g_end) __syncthreads();  // the rhs planes are reused by the next group
  }
}

// ---------------------------------------------------------------------------------------
// host-side launch logic
// ---------------------------------------------------------------------------------------
static int pow2_floor(int v) {
  int p = 1;
  while (2 * p <= v) p *= 2;
  return p;
}

template <typename R>
static bool vec_ok(const void* a, const void* b, const void* c, const void* x, int64_t a_stride, int64_t x_stride) {
  auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  return al(a) && al(b) && al(c) && al(x) && (a_stride * sizeof(R)) % 16 == 0 && (x_stride * sizeof(R)) % 16 == 0;
}

template <typename R, bool EXACT, bool CLOBBER, int PF, int CPT>
static int launch_single(R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nmat, int64_t a_stride,
                         int64_t x_stride, int T, cudaStream_t stream, const char* name) {
  const size_t smem = ((size_t)(EXACT ? 3 : 2) * nlev + (size_t)(EXACT ? 1 : 2) * PF) * T * sizeof(R) * CPT;
  auto k = k_tridiag_lm<R, EXACT, CLOBBER, PF, CPT>;
  int rc = allow_smem(k, smem);
  if (rc) return rc;
  const int64_t tc = (int64_t)T * CPT;
  dim3 grid((unsigned)((ncol + tc - 1) / tc), (unsigned)nmat);
  k<<<grid, T, smem, stream>>>(dl, d, du, X, ncol, nlev, a_stride, x_stride);
  note_launch(name);
  return check_cuda(cudaGetLastError());
}

template <typename R, bool EXACT, bool CLOBBER, int CPT>
static int dispatch_pf_single(int pf, R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nmat, int64_t as,
                              int64_t xs, int T, cudaStream_t st, const char* name) {
  switch (pf) {
    case 4: return launch_single<R, EXACT, CLOBBER, 4, CPT>(dl, d, du, X, ncol, nlev, nmat, as, xs, T, st, name);
    case 8: return launch_single<R, EXACT, CLOBBER, 8, CPT>(dl, d, du, X, ncol, nlev, nmat, as, xs, T, st, name);
    case 24: return launch_single<R, EXACT, CLOBBER, 24, CPT>(dl, d, du, X, ncol, nlev, nmat, as, xs, T, st, name);
    case 32: return launch_single<R, EXACT, CLOBBER, 32, CPT>(dl, d, du, X, ncol, nlev, nmat, as, xs, T, st, name);
    default: return launch_single<R, EXACT, CLOBBER, 16, CPT>(dl, d, du, X, ncol, nlev, nmat, as, xs, T, st, name);
  }
}

template <typename R, bool EXACT, bool CLOBBER, int CPT>
static int launch_shared(R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nrhs, int64_t a_stride,
                         int64_t x_stride, int Tc, int G, int Ty, cudaStream_t stream, const char* name) {
  const size_t smem = (size_t)(3 + G) * nlev * Tc * sizeof(R) * CPT;
  auto k = k_tridiag_lm_shared<R, EXACT, CLOBBER, 16, CPT>;
  int rc = allow_smem(k, smem);
  if (rc) return rc;
  const int64_t tc = (int64_t)Tc * CPT;
  dim3 grid((unsigned)((ncol + tc - 1) / tc), CLOBBER ? 1u : (unsigned)((nrhs + G - 1) / G));
  k<<<grid, 32 + Tc * Ty, smem, stream>>>(dl, d, du, X, ncol, nlev, nrhs, G, Tc, a_stride, x_stride);
  note_launch(name);
  return check_cuda(cudaGetLastError());
}

template <typename R, bool EXACT, int CPT>
static int run_levmajor(R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nA, int nrhs, int64_t a_stride,
                        int64_t x_stride, bool clobber, cudaStream_t stream) {
  const Tuning tu = tuning();
  const bool f64 = sizeof(R) == 8;
  if (nA == nrhs) {
    // thread per (column group, matrix)
    int pf = tu.prefetch > 0 ? tu.prefetch : 16;
    if (pf != 4 && pf != 8 && pf != 16 && pf != 24 && pf != 32) pf = 16;
    const size_t per_col = ((size_t)(EXACT ? 3 : 2) * nlev + (size_t)(EXACT ? 1 : 2) * pf) * sizeof(R);
    // columns per CTA: 32 by default (a few CTAs per SM, staggered in phase); fewer only if one column is huge
    int tc = tu.cols_per_cta > 0 ? tu.cols_per_cta : 32;
    if ((size_t)tc * per_col > kMaxSmemPerCta) tc = pow2_floor((int)(kMaxSmemPerCta / per_col));
    if (tc < CPT) return EKAT_B200_ERR_TOO_LARGE;
    const int T = tc / CPT;
    const char* name = EXACT ? (f64 ? "tridiag_lm<f64,bfb>" : "tridiag_lm<f32,bfb>")
                             : (f64 ? "tridiag_lm<f64,thomas>" : "tridiag_lm<f32,thomas>");
    return clobber ? dispatch_pf_single<R, EXACT, true, CPT>(pf, dl, d, du, X, ncol, nlev, nrhs, a_stride, x_stride, T, stream, name)
                   : dispatch_pf_single<R, EXACT, false, CPT>(pf, dl, d, du, X, ncol, nlev, nrhs, a_stride, x_stride, T, stream, name);
  }
  // shared matrix, nrhs > 1
  int tcols = tu.cols_per_cta > 0 ? tu.cols_per_cta : 16;
  const size_t per_col_plane = (size_t)nlev * sizeof(R);
  int G = nrhs;
  // shrink the rhs group, then the column count, until the tile fits on chip
  while ((size_t)(3 + G) * tcols * per_col_plane > kMaxSmemPerCta && G > 1) G = (G + 1) / 2;
  while ((size_t)(3 + G) * tcols * per_col_plane > kMaxSmemPerCta && tcols > CPT) tcols /= 2;
  if ((size_t)(3 + G) * tcols * per_col_plane > kMaxSmemPerCta || tcols < CPT) return EKAT_B200_ERR_TOO_LARGE;
  if (tcols > 32 * CPT) tcols = 32 * CPT;  // the factorisers are one warp
  const int Tc = tcols / CPT;
  int Ty = G;
  while (32 + Tc * Ty > 1024) Ty = (Ty + 1) / 2;
  const char* name = EXACT ? (f64 ? "tridiag_lm_shared<f64,bfb>" : "tridiag_lm_shared<f32,bfb>")
                           : (f64 ? "tridiag_lm_shared<f64,thomas>" : "tridiag_lm_shared<f32,thomas>");
  return clobber ? launch_shared<R, EXACT, true, CPT>(dl, d, du, X, ncol, nlev, nrhs, a_stride, x_stride, Tc, G, Ty, stream, name)
                 : launch_shared<R, EXACT, false, CPT>(dl, d, du, X, ncol, nlev, nrhs, a_stride, x_stride, Tc, G, Ty, stream, name);
}

// Solve on LEV_MAJOR operands. exact: bfb arithmetic; else fast Thomas.
template <typename R>
int tridiag_levmajor(bool exact, R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nA, int nrhs,
                     int64_t a_stride, int64_t x_stride, unsigned flags, cudaStream_t stream) {
  const bool clobber = (flags & EKAT_B200_CLOBBER_DIAGS) != 0;
  const int variant = tuning().variant;
  // One matrix per column (shared by all its right-hand sides): the persistent warp-specialised kernel. With a single
  // right-hand side it is used for thomas (factor and solve chains on different warps: 90% of the HBM peak against 77%
  // for the one-warp kernel); bfb with a single right-hand side is bound by its division chain on either kernel and
  // keeps the one-warp kernel, which holds three tiles per SM instead of two (measured: 54% against 47%); so does f32
  // thomas (72% against 63%: half the bytes per level for the same number of barrier hand-offs).
  // Latency-bound batches of short columns (BASELINE config 1): the whole tile on chip, eliminated from both ends
  // (tridiag_small.cu decides by shape whether it applies; kVarSmall forces it for tests and A/B measurements).
  if (!exact && nA == nrhs && !clobber && (variant == kVarAuto || variant == kVarSmall)) {
    const int rc = tridiag_small<R>(dl, d, du, X, ncol, nlev, nrhs, a_stride, x_stride, variant == kVarSmall, stream);
    if (rc != EKAT_B200_ERR_UNSUPPORTED) return rc;
  }
  if (exact && nA == 1 && nrhs > 1 && !clobber && variant == kVarAuto) {
    // bfb, several right-hand sides sharing one matrix: forward sweeps and back-substitutions on different warps, so that
    // item i+1 goes down the columns while item i comes back up (tridiag_bfb_shared.cu)
    const int rc = tridiag_bfb_shared<R>(dl, d, du, X, X, ncol, nlev, nrhs, a_stride, x_stride, stream);
    if (rc != EKAT_B200_ERR_UNSUPPORTED) return rc;
  }
  const bool split_ok = nA == 1 && !clobber && var_allows_shared(variant);
  if (split_ok && (nrhs > 1 || (!exact && sizeof(R) == 8) || var_forces_shared(variant))) {
    const int rc = tridiag_tma_shared<R>(exact, dl, d, du, X, ncol, nlev, nrhs, a_stride, x_stride, stream);
    if (rc != EKAT_B200_ERR_UNSUPPORTED) return rc;
  }
  if (nA == nrhs && clobber && variant == kVarAuto) {
    // the in-place side effect (multipliers in dl(1..), pivots in d for bfb, :292-303; pivots in d for serial thomas,
    // :216-227, which then runs the bfb sweep -- within its tolerance by construction): the input-only ring kernel
    // stores the factorisation next to the solution (tridiag_bfb_stream.cu)
    const int rc = tridiag_bfb_stream<R, false>(dl, d, du, X, ncol, nlev, nrhs, a_stride, x_stride, stream, exact ? 3 : 2);
    if (rc != EKAT_B200_ERR_UNSUPPORTED) return rc;
  }
  if (nA == nrhs && !clobber && var_allows_warp_tile(variant)) {
    const int rc = tridiag_tma<R>(exact, dl, d, du, X, ncol, nlev, nrhs, a_stride, x_stride, stream);
    if (rc != EKAT_B200_ERR_UNSUPPORTED) return rc;
  }
  constexpr int VEC = 16 / sizeof(R);
  const bool vec = tuning().variant != kVarScalarCopies && vec_ok<R>(dl, d, du, X, a_stride, x_stride);
  if (vec) {
    return exact ? run_levmajor<R, true, VEC>(dl, d, du, X, ncol, nlev, nA, nrhs, a_stride, x_stride, clobber, stream)
                 : run_levmajor<R, false, VEC>(dl, d, du, X, ncol, nlev, nA, nrhs, a_stride, x_stride, clobber, stream);
  }
  return exact ? run_levmajor<R, true, 1>(dl, d, du, X, ncol, nlev, nA, nrhs, a_stride, x_stride, clobber, stream)
               : run_levmajor<R, false, 1>(dl, d, du, X, ncol, nlev, nA, nrhs, a_stride, x_stride, clobber, stream);
}

template int tridiag_levmajor<double>(bool, double*, double*, double*, double*, int64_t, int, int, int, int64_t,
                                      int64_t, unsigned, cudaStream_t);
template int tridiag_levmajor<float>(bool, float*, float*, float*, float*, int64_t, int, int, int, int64_t,
                                     int64_t, unsigned, cudaStream_t);

}  // namespace ekb
