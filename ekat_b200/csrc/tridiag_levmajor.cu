// tridiag_levmajor.cu -- batched Thomas / bfb solvers on the native level-major layout.
//
// Replaces the per-column loops around ekat::tridiag::thomas and ekat::tridiag::bfb
// (reference: src/algorithm/ekat_tridiag.hpp:177-214 team Thomas, :216-290 serial
// Thomas, :292-320 + :632-678 bfb). One thread owns one (column, rhs) recurrence;
// consecutive threads own consecutive columns, so every level step of a warp is one
// coalesced access ([level][column] arrays, column fastest).
//
// HBM traffic is the algorithmic minimum -- read dl,d,du,B once, write X once
// ((3+2*nrhs) words per column-level): the forward sweep's intermediates live in
// shared memory ([array][level][column-in-CTA], bank-conflict-free) instead of
// being written back and re-read as the reference's in-place sweeps do. Global
// loads are software-pipelined PF levels ahead through a register ring (all
// addresses are known up front), because the kernels are bound by the serial
// divide chain per level, not by issue slots: few threads per SM (shared memory
// holds ~100 f64 columns of 128 levels), each with deep memory-level parallelism.
//
//   EXACT (bfb):   q = dl[i]/d'[i-1]; d'[i] = d[i] - q*du[i-1]; x'[i] = x[i] - q*x'[i-1];
//                  x[n-1] = x'[n-1]/d'[n-1]; x[i-1] = (x'[i-1] - du[i-1]*x[i])/d'[i-1]
//                  every * - / separately rounded (common.cuh x_*): bit-identical to the
//                  reference's bfb == thomas at -ffp-contract=off.
//   FAST (thomas): reciprocal form with FMAs, r = 1/(d[i] - dl[i]*c[i-1]); c[i] = du[i]*r;
//                  y[i] = (x[i] - dl[i]*y[i-1])*r;  x[i] = y[i] - c[i]*x[i+1]
//                  -- 1 reciprocal per level instead of 2 divisions, no divide in the
//                  back-substitution chain; agrees to ~1e-15 relative.
#include "common.cuh"

namespace ekb {

// ---------------------------------------------------------------------------------------
// One matrix per rhs (nA == nrhs, incl. nrhs == 1): thread <-> (column, blockIdx.y = rhs).
// smem: s0 = d' (EXACT) / c (FAST), s1 = x' / y, s2 = du (EXACT only); each [nlev][T].
// ---------------------------------------------------------------------------------------
template <typename R, bool EXACT, bool CLOBBER, int PF>
__global__ void k_tridiag_lm(R* __restrict__ dl, R* __restrict__ d, R* __restrict__ du, R* __restrict__ X,
                             int64_t ncol, int nlev, int64_t a_stride, int64_t x_stride) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int T = blockDim.x, t = threadIdx.x;
  R* s0 = reinterpret_cast<R*>(smem_raw);
  R* s1 = s0 + (size_t)nlev * T;
  R* s2 = s1 + (size_t)nlev * T;

  int64_t col = (int64_t)blockIdx.x * T + t;
  const bool live = col < ncol;
  if (!live) col = ncol - 1;  // keep the warp convergent: dead lanes redo the last column, never store
  const size_t mat = (size_t)blockIdx.y * nlev;
  R* pdl = dl + mat * a_stride + col;
  R* pd = d + mat * a_stride + col;
  R* pdu = du + mat * a_stride + col;
  R* px = X + mat * x_stride + col;

  R rdl[PF], rd[PF], rdu[PF], rx[PF];
#pragma unroll
  for (int u = 0; u < PF; ++u) {
    if (u < nlev) {
      rdl[u] = ld_stream(pdl + (size_t)u * a_stride);
      rd[u] = ld_stream(pd + (size_t)u * a_stride);
      rdu[u] = ld_stream(pdu + (size_t)u * a_stride);
      rx[u] = ld_stream(px + (size_t)u * x_stride);
    }
  }

  R dprev = 0, xprev = 0, uprev = 0;
  for (int base = 0; base < nlev; base += PF) {
#pragma unroll
    for (int u = 0; u < PF; ++u) {
      const int i = base + u;
      if (i < nlev) {
        const R a = rdl[u], b = rd[u], c = rdu[u], x = rx[u];
        const int nx = i + PF;
        if (nx < nlev) {
          rdl[u] = ld_stream(pdl + (size_t)nx * a_stride);
          rd[u] = ld_stream(pd + (size_t)nx * a_stride);
          rdu[u] = ld_stream(pdu + (size_t)nx * a_stride);
          rx[u] = ld_stream(px + (size_t)nx * x_stride);
        }
        const size_t si = (size_t)i * T + t;
        if (EXACT) {
          R dcur = b, xcur = x;
          if (i > 0) {
            const R q = x_div(a, dprev);
            dcur = x_sub(b, x_mul(q, uprev));
            xcur = x_sub(x, x_mul(q, xprev));
            if (CLOBBER && live) {  // reference leaves multipliers in dl, pivots in d (:299-302)
              st_stream(pdl + (size_t)i * a_stride, q);
              st_stream(pd + (size_t)i * a_stride, dcur);
            }
          }
          s0[si] = dcur;
          s1[si] = xcur;
          s2[si] = c;
          dprev = dcur;
          xprev = xcur;
          uprev = c;
        } else {
          const R den = i > 0 ? f_fma(-a, uprev, b) : b;
          const R r = f_rcp(den);
          const R ccur = c * r;
          const R xcur = (i > 0 ? f_fma(-a, xprev, x) : x) * r;
          if (CLOBBER && live && i > 0) st_stream(pd + (size_t)i * a_stride, den);  // pivots, as serial thomas (:221)
          s0[si] = ccur;
          s1[si] = xcur;
          uprev = ccur;
          xprev = xcur;
        }
      }
    }
  }

  // back substitution straight out of shared memory, solution streamed to HBM
  if (EXACT) {
    size_t si = (size_t)(nlev - 1) * T + t;
    R xv = x_div(s1[si], s0[si]);
    if (live) st_stream(px + (size_t)(nlev - 1) * x_stride, xv);
#pragma unroll 4
    for (int i = nlev - 1; i > 0; --i) {
      si = (size_t)(i - 1) * T + t;
      xv = x_div(x_sub(s1[si], x_mul(s2[si], xv)), s0[si]);
      if (live) st_stream(px + (size_t)(i - 1) * x_stride, xv);
    }
  } else {
    R xv = s1[(size_t)(nlev - 1) * T + t];
    if (live) st_stream(px + (size_t)(nlev - 1) * x_stride, xv);
#pragma unroll 8
    for (int i = nlev - 2; i >= 0; --i) {
      const size_t si = (size_t)i * T + t;
      xv = f_fma(-s0[si], xv, s1[si]);
      if (live) st_stream(px + (size_t)i * x_stride, xv);
    }
  }
}

// ---------------------------------------------------------------------------------------
// One matrix shared by nrhs right-hand sides (nA == 1, nrhs > 1).
// CTA = Tc columns x G rhs (threadIdx.y strides over the CTA's rhs group, blockIdx.y =
// group). Phase 1: the ty == 0 threads factorise once per column into shared memory --
// the structure of the reference (factorize in Kokkos::single, barrier, per-rhs solves:
// ekat_tridiag.hpp:323-334, :632-648). Phase 2: every (column, rhs) thread sweeps.
// smem: sa0, sa1, sa2 : [nlev][Tc]   (EXACT: q, d', du   FAST: m = dl*r, r, c = du*r)
//       sx            : [G][nlev][Tc]
// ---------------------------------------------------------------------------------------
template <typename R, bool EXACT, bool CLOBBER, int PF>
__global__ void k_tridiag_lm_shared(R* __restrict__ dl, R* __restrict__ d, R* __restrict__ du,
                                    R* __restrict__ X, int64_t ncol, int nlev, int nrhs, int G,
                                    int64_t a_stride, int64_t x_stride) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int Tc = blockDim.x, t = threadIdx.x, ty = threadIdx.y;
  const size_t plane = (size_t)nlev * Tc;
  R* sa0 = reinterpret_cast<R*>(smem_raw);
  R* sa1 = sa0 + plane;
  R* sa2 = sa1 + plane;
  R* sx = sa2 + plane;

  int64_t col = (int64_t)blockIdx.x * Tc + t;
  const bool live = col < ncol;
  if (!live) col = ncol - 1;
  const int j0 = blockIdx.y * G;
  const int jn = min(G, nrhs - j0);  // rhs in this CTA's group

  if (ty == 0) {
    R* pdl = dl + col;
    R* pd = d + col;
    R* pdu = du + col;
    R rdl[PF], rd[PF], rdu[PF];
#pragma unroll
    for (int u = 0; u < PF; ++u) {
      if (u < nlev) {
        rdl[u] = ld_stream(pdl + (size_t)u * a_stride);
        rd[u] = ld_stream(pd + (size_t)u * a_stride);
        rdu[u] = ld_stream(pdu + (size_t)u * a_stride);
      }
    }
    R dprev = 0, uprev = 0;
    const bool wr = CLOBBER && live && blockIdx.y == 0;
    for (int base = 0; base < nlev; base += PF) {
#pragma unroll
      for (int u = 0; u < PF; ++u) {
        const int i = base + u;
        if (i < nlev) {
          const R a = rdl[u], b = rd[u], c = rdu[u];
          const int nx = i + PF;
          if (nx < nlev) {
            rdl[u] = ld_stream(pdl + (size_t)nx * a_stride);
            rd[u] = ld_stream(pd + (size_t)nx * a_stride);
            rdu[u] = ld_stream(pdu + (size_t)nx * a_stride);
          }
          const size_t si = (size_t)i * Tc + t;
          if (EXACT) {
            R q = 0, dcur = b;
            if (i > 0) {
              q = x_div(a, dprev);
              dcur = x_sub(b, x_mul(q, uprev));
              if (wr) {
                st_stream(pdl + (size_t)i * a_stride, q);
                st_stream(pd + (size_t)i * a_stride, dcur);
              }
            }
            sa0[si] = q;
            sa1[si] = dcur;
            sa2[si] = c;
            dprev = dcur;
            uprev = c;
          } else {
            const R den = i > 0 ? f_fma(-a, uprev, b) : b;
            const R r = f_rcp(den);
            const R ccur = c * r;
            if (wr && i > 0) st_stream(pd + (size_t)i * a_stride, den);
            sa0[si] = a * r;
            sa1[si] = r;
            sa2[si] = ccur;
            uprev = ccur;
          }
        }
      }
    }
  }
  __syncthreads();

  for (int jj = ty; jj < jn; jj += blockDim.y) {
    R* px = X + (size_t)(j0 + jj) * nlev * x_stride + col;
    R* sxj = sx + (size_t)jj * plane;
    R rx[PF];
#pragma unroll
    for (int u = 0; u < PF; ++u)
      if (u < nlev) rx[u] = ld_stream(px + (size_t)u * x_stride);
    R xprev = 0;
    for (int base = 0; base < nlev; base += PF) {
#pragma unroll
      for (int u = 0; u < PF; ++u) {
        const int i = base + u;
        if (i < nlev) {
          const R x = rx[u];
          const int nx = i + PF;
          if (nx < nlev) rx[u] = ld_stream(px + (size_t)nx * x_stride);
          const size_t si = (size_t)i * Tc + t;
          R xcur;
          if (EXACT) {
            xcur = i > 0 ? x_sub(x, x_mul(sa0[si], xprev)) : x;
          } else {
            const R xr = x * sa1[si];
            xcur = i > 0 ? f_fma(-sa0[si], xprev, xr) : xr;
          }
          sxj[si] = xcur;
          xprev = xcur;
        }
      }
    }
    if (EXACT) {
      size_t si = (size_t)(nlev - 1) * Tc + t;
      R xv = x_div(sxj[si], sa1[si]);
      if (live) st_stream(px + (size_t)(nlev - 1) * x_stride, xv);
#pragma unroll 4
      for (int i = nlev - 1; i > 0; --i) {
        si = (size_t)(i - 1) * Tc + t;
        xv = x_div(x_sub(sxj[si], x_mul(sa2[si], xv)), sa1[si]);
        if (live) st_stream(px + (size_t)(i - 1) * x_stride, xv);
      }
    } else {
      R xv = sxj[(size_t)(nlev - 1) * Tc + t];
      if (live) st_stream(px + (size_t)(nlev - 1) * x_stride, xv);
#pragma unroll 8
      for (int i = nlev - 2; i >= 0; --i) {
        const size_t si = (size_t)i * Tc + t;
        xv = f_fma(-sa2[si], xv, sxj[si]);
        if (live) st_stream(px + (size_t)i * x_stride, xv);
      }
    }
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

template <typename R, bool EXACT, bool CLOBBER, int PF>
static int launch_single(R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nmat, int64_t a_stride,
                         int64_t x_stride, int T, cudaStream_t stream, const char* name) {
  const size_t smem = (size_t)(EXACT ? 3 : 2) * nlev * T * sizeof(R);
  auto k = k_tridiag_lm<R, EXACT, CLOBBER, PF>;
  int rc = allow_smem(k, smem);
  if (rc) return rc;
  dim3 grid((unsigned)((ncol + T - 1) / T), (unsigned)nmat);
  k<<<grid, T, smem, stream>>>(dl, d, du, X, ncol, nlev, a_stride, x_stride);
  note_launch(name);
  return check_cuda(cudaGetLastError());
}

template <typename R, bool EXACT, bool CLOBBER, int PF>
static int launch_shared(R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nrhs, int64_t a_stride,
                         int64_t x_stride, int Tc, int G, int Ty, cudaStream_t stream, const char* name) {
  const size_t smem = (size_t)(3 + G) * nlev * Tc * sizeof(R);
  auto k = k_tridiag_lm_shared<R, EXACT, CLOBBER, PF>;
  int rc = allow_smem(k, smem);
  if (rc) return rc;
  dim3 grid((unsigned)((ncol + Tc - 1) / Tc), (unsigned)((nrhs + G - 1) / G));
  dim3 block(Tc, Ty);
  k<<<grid, block, smem, stream>>>(dl, d, du, X, ncol, nlev, nrhs, G, a_stride, x_stride);
  note_launch(name);
  return check_cuda(cudaGetLastError());
}

template <typename R, bool EXACT, bool CLOBBER>
static int dispatch_pf_single(int pf, R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nmat, int64_t as,
                              int64_t xs, int T, cudaStream_t st, const char* name) {
  switch (pf) {
    case 4: return launch_single<R, EXACT, CLOBBER, 4>(dl, d, du, X, ncol, nlev, nmat, as, xs, T, st, name);
    case 16: return launch_single<R, EXACT, CLOBBER, 16>(dl, d, du, X, ncol, nlev, nmat, as, xs, T, st, name);
    default: return launch_single<R, EXACT, CLOBBER, 8>(dl, d, du, X, ncol, nlev, nmat, as, xs, T, st, name);
  }
}

// Solve on LEV_MAJOR operands. exact: bfb arithmetic; else fast Thomas.
template <typename R>
int tridiag_levmajor(bool exact, R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nA, int nrhs,
                     int64_t a_stride, int64_t x_stride, unsigned flags, cudaStream_t stream) {
  const bool clobber = (flags & EKAT_B200_CLOBBER_DIAGS) != 0;
  const Tuning tu = tuning();
  const int pf = tu.prefetch > 0 ? tu.prefetch : 8;
  if (nA == nrhs) {
    // thread per (column, matrix). Columns per CTA: a warp unless a column is too long for that.
    const size_t per_col = (size_t)(exact ? 3 : 2) * nlev * sizeof(R);
    int T = tu.cols_per_cta > 0 ? tu.cols_per_cta : 32;
    if ((size_t)T * per_col > kMaxSmemPerCta) T = pow2_floor((int)(kMaxSmemPerCta / per_col));
    if (T < 1) return EKAT_B200_ERR_TOO_LARGE;
    const char* name = exact ? (sizeof(R) == 8 ? "tridiag_lm<f64,bfb>" : "tridiag_lm<f32,bfb>")
                             : (sizeof(R) == 8 ? "tridiag_lm<f64,thomas>" : "tridiag_lm<f32,thomas>");
    if (exact) {
      return clobber ? dispatch_pf_single<R, true, true>(pf, dl, d, du, X, ncol, nlev, nrhs, a_stride, x_stride, T, stream, name)
                     : dispatch_pf_single<R, true, false>(pf, dl, d, du, X, ncol, nlev, nrhs, a_stride, x_stride, T, stream, name);
    }
    return clobber ? dispatch_pf_single<R, false, true>(pf, dl, d, du, X, ncol, nlev, nrhs, a_stride, x_stride, T, stream, name)
                   : dispatch_pf_single<R, false, false>(pf, dl, d, du, X, ncol, nlev, nrhs, a_stride, x_stride, T, stream, name);
  }
  // shared matrix, nrhs > 1
  int Tc = tu.cols_per_cta > 0 ? tu.cols_per_cta : 16;
  const size_t per_col_plane = (size_t)nlev * sizeof(R);
  int G = nrhs;
  // shrink the rhs group, then the column count, until the tile fits on chip
  while ((size_t)(3 + G) * Tc * per_col_plane > kMaxSmemPerCta && G > 1) G = (G + 1) / 2;
  while ((size_t)(3 + G) * Tc * per_col_plane > kMaxSmemPerCta && Tc > 1) Tc /= 2;
  if ((size_t)(3 + G) * Tc * per_col_plane > kMaxSmemPerCta) return EKAT_B200_ERR_TOO_LARGE;
  int Ty = G;
  while (Tc * Ty > 1024) Ty = (Ty + 1) / 2;
  const char* name = exact ? (sizeof(R) == 8 ? "tridiag_lm_shared<f64,bfb>" : "tridiag_lm_shared<f32,bfb>")
                           : (sizeof(R) == 8 ? "tridiag_lm_shared<f64,thomas>" : "tridiag_lm_shared<f32,thomas>");
  if (exact) {
    return clobber ? launch_shared<R, true, true, 8>(dl, d, du, X, ncol, nlev, nrhs, a_stride, x_stride, Tc, G, Ty, stream, name)
                   : launch_shared<R, true, false, 8>(dl, d, du, X, ncol, nlev, nrhs, a_stride, x_stride, Tc, G, Ty, stream, name);
  }
  return clobber ? launch_shared<R, false, true, 8>(dl, d, du, X, ncol, nlev, nrhs, a_stride, x_stride, Tc, G, Ty, stream, name)
                 : launch_shared<R, false, false, 8>(dl, d, du, X, ncol, nlev, nrhs, a_stride, x_stride, Tc, G, Ty, stream, name);
}

template int tridiag_levmajor<double>(bool, double*, double*, double*, double*, int64_t, int, int, int, int64_t,
                                      int64_t, unsigned, cudaStream_t);
template int tridiag_levmajor<float>(bool, float*, float*, float*, float*, int64_t, int, int, int, int64_t,
                                     int64_t, unsigned, cudaStream_t);

}  // namespace ekb
