// thomas_two_sided.cuh -- Thomas elimination FROM BOTH ENDS ("burn at both ends"), shared by the kernels that hold whole
// columns on chip (fused_tma.cu, tridiag_cm_rhs.cu; tridiag_small.cu has its own form with deferred divisions).
//
// The column is split at level M = 8 * ceil(nch / 2): levels 0 .. M-1 are eliminated downwards (the classical forward
// sweep of ekat::tridiag::thomas, reference src/algorithm/ekat_tridiag.hpp:177-214), levels nlev-1 .. M upwards (the same
// recurrence on the reversed system: du plays dl's part and vice versa); the two halves meet in a 2x2 system for
// x(M-1), x(M), and the back-substitution runs outwards from there. Every thread runs BOTH halves interleaved: two
// independent chains where there was one, so the dependent latency per column halves at the same instruction count.
// Levels are taken 4 at a time per half (kHS), operands of both halves in registers before anything is stored.
// FULL: every chunk of 8 levels is whole (no per-level guards); otherwise levels past the end are identity rows.
// Pointers are the thread's column (element of level 0); `ls` = elements between levels.
// Agreement with the reference's one-sided Thomas: rounding level (diagonally dominant systems), as tridiag_small.cu.
#pragma once
#include "tridiag_tma.cuh"

namespace ekb {

constexpr int kHS = 4;  // levels per half-step

// level of slot s (0..7) of chunk k in processing order: ascending for the top half, descending for the bottom half
template <bool UP>
__device__ __forceinline__ int slot_level(int k, int s) { return k * kLB + (UP ? kLB - 1 - s : s); }

// Factor, one half-step (4 slots) of a half; state (q1, q2, cp, gp) per half: ratio form of tridiag_tma_shared.cu.
// In place: pm <- multipliers (coupling to the level processed BEFORE: dl*g above, du*g below), pr <- 1/pivot,
// pc <- coupling to the level processed AFTER times 1/pivot (c' above, a~' below).
template <typename R>
struct FacState {
  R q1, q2, cp, gp;
};
template <typename R, bool UP, bool FULL>
__device__ __forceinline__ void factor_load(const R* pm, const R* pr, const R* pc, int ls, int k, int s0, int nlev, R (&a)[kHS],
                                            R (&b)[kHS], R (&c)[kHS]) {
#pragma unroll
  for (int u = 0; u < kHS; ++u) {
    const int lev = slot_level<UP>(k, s0 + u);
    const bool ok = FULL || lev < nlev;
    const int o = (ok ? lev : 0) * ls;
    const R vl = pm[o], vd = pr[o], vu = pc[o];
    a[u] = UP ? vu : vl;
    c[u] = UP ? vl : vu;
    b[u] = vd;
    if (lev == 0) (UP ? c[u] : a[u]) = (R)0;         // dl(0) is never used (ekat_tridiag.hpp:24-27)
    if (lev == nlev - 1) (UP ? a[u] : c[u]) = (R)0;  // du(nrow-1) is never used
    if (!ok) {
      a[u] = (R)0;
      b[u] = (R)1;
      c[u] = (R)0;
    }
  }
}
template <typename R, bool UP, bool FULL>
__device__ __forceinline__ void factor_run(R* pm, R* pr, R* pc, int ls, int k, int s0, int nlev, const R (&a)[kHS], const R (&b)[kHS],
                                           const R (&c)[kHS], FacState<R>& st) {
  constexpr int RN = TmaCfg<R>::RN;
#pragma unroll
  for (int u = 0; u < kHS; ++u) {
    const int s = s0 + u;
    R q;
    if (s % RN == 0) q = f_fma(-(a[u] * st.cp), st.q2, b[u]);  // renormalised state: q[prev] == 1
    else q = f_fma(b[u], st.q1, -((a[u] * st.cp) * st.q2));
    const R rq = t_rcp(q);
    const R g = (s % RN == 0) ? rq : st.q1 * rq;  // q[prev]/q[this] = 1/pivot
    const int lev = slot_level<UP>(k, s);
    if (FULL || lev < nlev) {
      const int o = lev * ls;
      pm[o] = a[u] * st.gp;
      pr[o] = g;
      pc[o] = c[u] * g;
    }
    if ((s + 1) % RN == 0) {
      st.q2 = g;
      st.q1 = (R)1;
    } else {
      st.q2 = (s % RN == 0) ? (R)1 : st.q1;
      st.q1 = q;
    }
    st.cp = c[u];
    st.gp = g;
  }
}

// Solver thread, way in (elimination of its right-hand side towards the middle): z = x - m*z_prev, stored as z * (1/pivot).
template <typename R, bool UP, bool FULL>
__device__ __forceinline__ void sweep_in_load(const R* pm, const R* pr, int la, const R* px, int lx, int k, int s0, int nlev,
                                              R (&m)[kHS], R (&g)[kHS], R (&x)[kHS]) {
#pragma unroll
  for (int u = 0; u < kHS; ++u) {
    const int lev = slot_level<UP>(k, s0 + u);
    const bool ok = FULL || lev < nlev;
    const int l = ok ? lev : 0;
    m[u] = ok ? pm[l * la] : (R)0;
    g[u] = ok ? pr[l * la] : (R)1;
    x[u] = ok ? px[l * lx] : (R)0;
  }
}
template <typename R, bool UP, bool FULL>
__device__ __forceinline__ void sweep_in_run(R* px, int lx, int k, int s0, int nlev, const R (&m)[kHS], const R (&g)[kHS],
                                             const R (&x)[kHS], R& z) {
#pragma unroll
  for (int u = 0; u < kHS; ++u) {
    z = f_fma(-m[u], z, x[u]);
    const int lev = slot_level<UP>(k, s0 + u);
    if (FULL || lev < nlev) px[lev * lx] = z * g[u];
  }
}
// The two halves meet: x(M-1) + c'(M-1) x(M) = y(M-1),  a~'(M) x(M-1) + x(M) = y~(M). The solved pair is written over
// y(M-1), y~(M): the way out starts from x = 0 with the coupling of these two levels switched off and finds them in place.
template <typename R>
__device__ __forceinline__ void sweep_meet(const R* pc, int la, R* px, int lx, int M) {
  const R ya = px[(M - 1) * lx], yb = px[M * lx], ca = pc[(M - 1) * la], ab = pc[M * la];
  const R rdet = f_rcp(f_fma(-ca, ab, (R)1));
  px[(M - 1) * lx] = f_fma(-ca, yb, ya) * rdet;
  px[M * lx] = f_fma(-ab, ya, yb) * rdet;
}
// Way out (back-substitution from the middle): x = y - c*x_prev, slots of a chunk in REVERSE processing order.
template <typename R, bool UP, bool FULL>
__device__ __forceinline__ void sweep_out_load(const R* pc, int la, const R* px, int lx, int k, int s0, int nlev, R (&c)[kHS],
                                               R (&y)[kHS]) {
#pragma unroll
  for (int u = 0; u < kHS; ++u) {
    const int lev = slot_level<UP>(k, s0 + kHS - 1 - u);
    const bool ok = FULL || lev < nlev;
    const int l = ok ? lev : 0;
    c[u] = ok ? pc[l * la] : (R)0;
    y[u] = ok ? px[l * lx] : (R)0;
  }
}
template <typename R, bool UP, bool FULL>
__device__ __forceinline__ void sweep_out_run(R* px, int lx, int k, int s0, int nlev, const R (&c)[kHS], const R (&y)[kHS], R& xv) {
#pragma unroll
  for (int u = 0; u < kHS; ++u) {
    const int lev = slot_level<UP>(k, s0 + kHS - 1 - u);
    if (FULL || lev < nlev) {
      xv = f_fma(-c[u], xv, y[u]);
      px[lev * lx] = xv;
    }
  }
}

// Both sweeps of one right-hand side. wait_step(t): chunk t of the top half and chunk nch-1-t of the bottom half are
// factorised. pm, pr, pc: the column's factor planes (level stride la); px: its right-hand side (level stride lx).
template <typename R, bool FULL, class Wait>
__device__ __forceinline__ void two_sided_sweeps(const R* pm, const R* pr, const R* pc, int la, R* px, int lx, int nlev, Wait wait_step) {
  const int nch = (nlev + kLB - 1) / kLB, kt = (nch + 1) / 2, M = tmin(kt * kLB, nlev);
  R zt = (R)0, zb = (R)0;
  for (int t = 0; t < kt; ++t) {  // way in, both halves interleaved
    wait_step(t);
    const int kb = nch - 1 - t;
    const bool bot = kb >= kt;
#pragma unroll
    for (int s0 = 0; s0 < kLB; s0 += kHS) {
      R m0[kHS], g0[kHS], x0[kHS], m1[kHS], g1[kHS], x1[kHS];
      sweep_in_load<R, false, FULL>(pm, pr, la, px, lx, t, s0, nlev, m0, g0, x0);
      if (bot) sweep_in_load<R, true, FULL>(pm, pr, la, px, lx, kb, s0, nlev, m1, g1, x1);
      sweep_in_run<R, false, FULL>(px, lx, t, s0, nlev, m0, g0, x0, zt);
      if (bot) sweep_in_run<R, true, FULL>(px, lx, kb, s0, nlev, m1, g1, x1, zb);
    }
  }
  if (M < nlev) sweep_meet<R>(pc, la, px, lx, M);
  R xt = (R)0, xb = (R)0;
  for (int t = kt - 1; t >= 0; --t) {  // way out
    const int kb = nch - 1 - t;
    const bool bot = kb >= kt;
#pragma unroll
    for (int s0 = kLB - kHS; s0 >= 0; s0 -= kHS) {
      R c0[kHS], y0[kHS], c1[kHS], y1[kHS];
      sweep_out_load<R, false, FULL>(pc, la, px, lx, t, s0, nlev, c0, y0);
      if (bot) sweep_out_load<R, true, FULL>(pc, la, px, lx, kb, s0, nlev, c1, y1);
      if (s0 == kLB - kHS && M < nlev) {  // the interface levels are solved: no coupling beyond them
        if (t == kt - 1) c0[0] = (R)0;
        if (kb == kt) c1[0] = (R)0;
      }
      sweep_out_run<R, false, FULL>(px, lx, t, s0, nlev, c0, y0, xt);
      if (bot) sweep_out_run<R, true, FULL>(px, lx, kb, s0, nlev, c1, y1, xb);
    }
  }
}

// The factorisation of one column from both ends; done_step(t) after the chunks of step t are in place.
template <typename R, bool FULL, class Done>
__device__ __forceinline__ void two_sided_factor(R* pm, R* pr, R* pc, int la, int nlev, Done done_step) {
  const int nch = (nlev + kLB - 1) / kLB, kt = (nch + 1) / 2;
  FacState<R> ft = {(R)1, (R)0, (R)0, (R)0}, fb = {(R)1, (R)0, (R)0, (R)0};
  for (int t = 0; t < kt; ++t) {
    const int kb = nch - 1 - t;
    const bool bot = kb >= kt;
#pragma unroll
    for (int s0 = 0; s0 < kLB; s0 += kHS) {
      R a0[kHS], b0[kHS], c0[kHS], a1[kHS], b1[kHS], c1[kHS];
      factor_load<R, false, FULL>(pm, pr, pc, la, t, s0, nlev, a0, b0, c0);
      if (bot) factor_load<R, true, FULL>(pm, pr, pc, la, kb, s0, nlev, a1, b1, c1);
      factor_run<R, false, FULL>(pm, pr, pc, la, t, s0, nlev, a0, b0, c0, ft);
      if (bot) factor_run<R, true, FULL>(pm, pr, pc, la, kb, s0, nlev, a1, b1, c1, fb);
    }
    done_step(t);
  }
}

}  // namespace ekb
