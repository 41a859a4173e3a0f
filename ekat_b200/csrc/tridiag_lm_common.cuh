// tridiag_lm_common.cuh -- device building blocks shared by the level-major tridiagonal kernels
// (tridiag_levmajor.cu) and the fused solve+remap kernel (fused.cu).
//
// Data movement: every operand is brought from HBM to shared memory with cp.async (16-byte LDGSTS when
// the pitch allows it: one thread owns CPT = 16/sizeof(Real) adjacent columns, so a warp's level step is
// one fully coalesced 512-byte access), each level in its own commit group, PF levels in flight per
// thread. Operands that the back-substitution needs again land directly in their final "plane"
// ([level][thread] vectors, overwritten in place by the forward sweep); operands that are dead after
// the forward sweep go through a small PF-deep ring. No thread ever reads shared memory written by
// another thread in these sweeps, so no block barrier is needed around the pipeline.
#pragma once
#include "common.cuh"

namespace ekb {

template <typename R, int CPT>
struct alignas(sizeof(R) * CPT) Vec {
  R v[CPT];
};

template <typename R, int CPT>
__device__ __forceinline__ void st_vec_stream(R* g, const Vec<R, CPT>& x, int nlive) {
  if (CPT > 1 && nlive == CPT) {
    if constexpr (sizeof(R) * CPT == 16) {
      __stcs(reinterpret_cast<uint4*>(g), *reinterpret_cast<const uint4*>(&x));
    } else if constexpr (sizeof(R) * CPT == 8) {
      __stcs(reinterpret_cast<uint2*>(g), *reinterpret_cast<const uint2*>(&x));
    }
  } else {
#pragma unroll
    for (int k = 0; k < CPT; ++k)
      if (k < nlive) __stcs(g + k, x.v[k]);
  }
}

// Column ownership of a thread: CPT adjacent columns starting at c0; nlive of them exist.
// Threads past the end re-read the last vector (keeps warps convergent) and never store.
template <int CPT>
__device__ __forceinline__ void own_columns(int64_t first, int64_t ncol, int64_t& c0, int& nlive) {
  c0 = first;
  const int64_t left = ncol - c0;
  nlive = left >= CPT ? CPT : (left > 0 ? (int)left : 0);
  if (nlive == 0) c0 = ((ncol - 1) / CPT) * CPT;
}

// ------------------------------------------------------------------------------------------------
// Factorisation sweep of one matrix per owned column, rolling cp.async window of PF levels.
//   EXACT: planes  q (multipliers, from dl), d' (pivots, from d), du          (bfb, ekat_tridiag.hpp:292-303)
//   FAST : planes  m = dl*r, r = 1/pivot, c = du*r                            (reciprocal-form Thomas)
// All three operands land in place in (pl0, pl1, pl2); `stride` = T (vectors between levels).
// ------------------------------------------------------------------------------------------------
template <typename R, bool EXACT, bool CLOBBER, int PF, int CPT>
__device__ __forceinline__ void lm_factorize(R* __restrict__ gdl, R* __restrict__ gd, R* __restrict__ gdu,
                                             int64_t a_stride, int nlev, Vec<R, CPT>* pl0, Vec<R, CPT>* pl1,
                                             Vec<R, CPT>* pl2, int stride, int t, int nlive) {
  using V = Vec<R, CPT>;
  constexpr int B = sizeof(R) * CPT;
  auto issue = [&](int i) {
    if (i < nlev) {
      const size_t s = (size_t)i * stride + t, g = (size_t)i * a_stride;
      cp_async<B>(&pl0[s], gdl + g);
      cp_async<B>(&pl1[s], gd + g);
      cp_async<B>(&pl2[s], gdu + g);
    }
    cp_async_commit();
  };
#pragma unroll
  for (int i = 0; i < PF; ++i) issue(i);
  cp_async_wait<PF - 1>();
  V a = pl0[t], b = pl1[t], c = pl2[t];
  V dprev, uprev;
#pragma unroll 2
  for (int i = 0; i < nlev; ++i) {
    V an, bn, cn;
    if (i + 1 < nlev) {
      cp_async_wait<PF - 2>();
      const size_t s = (size_t)(i + 1) * stride + t;
      an = pl0[s];
      bn = pl1[s];
      cn = pl2[s];
    }
    issue(i + PF);
    const size_t s = (size_t)i * stride + t;
    V o0, o1, o2;
#pragma unroll
    for (int k = 0; k < CPT; ++k) {
      if (EXACT) {
        R q = 0, dcur = b.v[k];
        if (i > 0) {
          q = x_div_pipelined(a.v[k], dprev.v[k]);
          dcur = x_sub(b.v[k], x_mul(q, uprev.v[k]));
        }
        o0.v[k] = q;
        o1.v[k] = dcur;
        o2.v[k] = c.v[k];
        dprev.v[k] = dcur;
        uprev.v[k] = c.v[k];
      } else {
        const R den = i > 0 ? f_fma(-a.v[k], uprev.v[k], b.v[k]) : b.v[k];
        const R r = f_rcp(den);
        const R cc = c.v[k] * r;
        o0.v[k] = a.v[k] * r;
        o1.v[k] = r;
        o2.v[k] = cc;
        uprev.v[k] = cc;
        if (CLOBBER) dprev.v[k] = den;
      }
    }
    pl0[s] = o0;
    pl1[s] = o1;
    if (!EXACT) pl2[s] = o2;
    if (CLOBBER && i > 0 && nlive > 0) {
      // reference side effect: bfb/team-thomas leave multipliers in dl and pivots in d (:299-302);
      // serial thomas leaves pivots in d (:221)
      if (EXACT) st_vec_stream<R, CPT>(gdl + (size_t)i * a_stride, o0, nlive);
      st_vec_stream<R, CPT>(gd + (size_t)i * a_stride, EXACT ? o1 : dprev, nlive);
    }
    a = an;
    b = bn;
    c = cn;
  }
  cp_async_wait<0>();
}

// ------------------------------------------------------------------------------------------------
// Solve sweeps of one right-hand side per owned column, in place in the plane `px` (already filled
// with B), against factor planes (pl0, pl1, pl2) as written by lm_factorize; solution streamed to gx
// (if non-null) and, if KEEP, left in px (the fused kernel interpolates from it).
// ------------------------------------------------------------------------------------------------
template <typename R, bool EXACT, int CPT, bool KEEP>
__device__ __forceinline__ void lm_solve(const Vec<R, CPT>* pl0, const Vec<R, CPT>* pl1, const Vec<R, CPT>* pl2,
                                         Vec<R, CPT>* px, int stride, int t, int nlev, R* __restrict__ gx,
                                         int64_t x_stride, int nlive) {
  using V = Vec<R, CPT>;
  V xp = px[t];
  if (!EXACT) {
    const V r0 = pl1[t];
#pragma unroll
    for (int k = 0; k < CPT; ++k) xp.v[k] = xp.v[k] * r0.v[k];
    px[t] = xp;
  }
#pragma unroll 4
  for (int i = 1; i < nlev; ++i) {
    const size_t s = (size_t)i * stride + t;
    const V x = px[s], m = pl0[s];
    V o;
    if (EXACT) {
#pragma unroll
      for (int k = 0; k < CPT; ++k) o.v[k] = x_sub(x.v[k], x_mul(m.v[k], xp.v[k]));  // :315-316
    } else {
      const V r = pl1[s];
#pragma unroll
      for (int k = 0; k < CPT; ++k) o.v[k] = f_fma(-m.v[k], xp.v[k], x.v[k] * r.v[k]);
    }
    px[s] = o;
    xp = o;
  }
  // back substitution
  size_t s = (size_t)(nlev - 1) * stride + t;
  V xv = xp;
  if (EXACT) {
    const V dn = pl1[s];
#pragma unroll
    for (int k = 0; k < CPT; ++k) xv.v[k] = x_div_pipelined(xp.v[k], dn.v[k]);  // :317
    if (KEEP) px[s] = xv;
  }
  if (gx && nlive > 0) st_vec_stream<R, CPT>(gx + (size_t)(nlev - 1) * x_stride, xv, nlive);
#pragma unroll 4
  for (int i = nlev - 2; i >= 0; --i) {
    s = (size_t)i * stride + t;
    const V x = px[s], u = pl2[s];
    if (EXACT) {
      const V dd = pl1[s];
#pragma unroll
      for (int k = 0; k < CPT; ++k) {
        const R r = rcp_refined(dd.v[k]);  // independent of the x chain: overlaps the previous level
        xv.v[k] = div_given_rcp(x_sub(x.v[k], x_mul(u.v[k], xv.v[k])), dd.v[k], r);  // :318-319
      }
    } else {
#pragma unroll
      for (int k = 0; k < CPT; ++k) xv.v[k] = f_fma(-u.v[k], xv.v[k], x.v[k]);
    }
    if (KEEP) px[s] = xv;
    if (gx && nlive > 0) st_vec_stream<R, CPT>(gx + (size_t)i * x_stride, xv, nlive);
  }
}

// Issue the cp.async of every level of one right-hand side into its plane (one commit group).
template <typename R, int CPT>
__device__ __forceinline__ void lm_issue_rhs(const R* __restrict__ gx, int64_t x_stride, int nlev, Vec<R, CPT>* px,
                                             int stride, int t) {
  constexpr int B = sizeof(R) * CPT;
  for (int i = 0; i < nlev; ++i) cp_async<B>(&px[(size_t)i * stride + t], gx + (size_t)i * x_stride);
  cp_async_commit();
}

}  // namespace ekb
