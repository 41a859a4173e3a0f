// tridiag_cr_warp.cu -- cyclic reduction with one WARP per (column, group of right-hand sides): a column's rows live in
// the registers of the warp's lanes (RPT consecutive rows per lane), neighbours across lanes travel by warp shuffles,
// and there is no shared memory and no barrier anywhere. One matrix per column (nA == 1), up to 256 levels.
//
// Replaces the per-column loops around ekat::tridiag::cr(team, dl, d, du, X) with X rank 1 or 2 (reference
// src/algorithm/ekat_tridiag.hpp:385-458, :460-544) for the shapes cyclic reduction exists for -- few columns, where a
// thread per column cannot fill the machine (guidance :71-74) -- and for the reference layout
// (tests/algorithm/tridiag_tests_performance.cpp:33-37: cr on 2,048 problems x 128 rows x 43 right-hand sides is the
// reference perf driver's default). Every row update is the reference's, operation for operation and separately
// rounded, with its clamped-neighbour trick (:408-418), so the results are bit-identical to the reference's CR at
// -ffp-contract=off.
//
// Why this shape. The reference's levels are separated by team barriers because its rows live in memory. Here lane l
// owns rows RPT*l .. RPT*l+RPT-1:
//   * levels with offset os < RPT stay inside a lane, except that local row 0 needs its left neighbour from lane l-1
//     (one shuffle of a row per level);
//   * levels with os = RPT*s, s = 1, 2, 4, 8, 16 involve local row 0 of the lanes that are multiples of 2s; their
//     neighbours are lanes l-s and l+s: two shuffles of a row per level. A warp executes in lock step, so "the level
//     is complete" needs no barrier;
//   * the bottom 2x2 solve (:427-437) is computed by every lane from the two broadcast rows; the way up mirrors the
//     way down with only solution values travelling.
// In the reference layout a column's rows are contiguous, so the loads and stores of a warp are coalesced as they
// stand; level-major operands are read with a stride (acceptable where this kernel is used for them: few columns,
// L2-resident). The matrix work is repeated by every right-hand-side group of a column (as the reference's
// sub-teams would not: :474-503), which is what keeps the groups independent.
#include "common.cuh"

namespace ekb {

struct CrwArgs {
  void *dl, *d, *du, *X;
  int64_t ncol, a_stride, x_stride;
  int nlev, nrhs, ngroups;
};

template <typename R, int G>
struct CrwRow {
  R l, d, u, x[G];
};

template <typename R, int G>
__device__ __forceinline__ CrwRow<R, G> crw_shfl_up(const CrwRow<R, G>& r, int s) {
  CrwRow<R, G> o;
  o.l = __shfl_up_sync(0xffffffffu, r.l, s);
  o.d = __shfl_up_sync(0xffffffffu, r.d, s);
  o.u = __shfl_up_sync(0xffffffffu, r.u, s);
#pragma unroll
  for (int g = 0; g < G; ++g) o.x[g] = __shfl_up_sync(0xffffffffu, r.x[g], s);
  return o;
}
template <typename R, int G>
__device__ __forceinline__ CrwRow<R, G> crw_shfl_down(const CrwRow<R, G>& r, int s) {
  CrwRow<R, G> o;
  o.l = __shfl_down_sync(0xffffffffu, r.l, s);
  o.d = __shfl_down_sync(0xffffffffu, r.d, s);
  o.u = __shfl_down_sync(0xffffffffu, r.u, s);
#pragma unroll
  for (int g = 0; g < G; ++g) o.x[g] = __shfl_down_sync(0xffffffffu, r.x[g], s);
  return o;
}
template <typename R, int G>
__device__ __forceinline__ CrwRow<R, G> crw_bcast(const CrwRow<R, G>& r, int src) {
  CrwRow<R, G> o;
  o.l = __shfl_sync(0xffffffffu, r.l, src);
  o.d = __shfl_sync(0xffffffffu, r.d, src);
  o.u = __shfl_sync(0xffffffffu, r.u, src);
#pragma unroll
  for (int g = 0; g < G; ++g) o.x[g] = __shfl_sync(0xffffffffu, r.x[g], src);
  return o;
}

// One row of a down-sweep level (:404-419 / :484-503). A missing neighbour is the row itself times a zero factor, as
// the reference's clamped index makes it -- including that d(i) then sees the NEW du(i) / dl(i).
template <typename R, int G>
__device__ __forceinline__ CrwRow<R, G> crw_down(const CrwRow<R, G>& me, const CrwRow<R, G>& m, const CrwRow<R, G>& p, bool im_ok,
                                                 bool ip_ok) {
  const R f1 = im_ok ? x_div(-me.l, m.d) : (R)0;  // :408-409
  const R f2 = ip_ok ? x_div(-me.u, p.d) : (R)0;
  CrwRow<R, G> o;
#pragma unroll
  for (int g = 0; g < G; ++g)
    o.x[g] = x_add(me.x[g], x_add(x_mul(f1, im_ok ? m.x[g] : me.x[g]), x_mul(f2, ip_ok ? p.x[g] : me.x[g])));  // :418 / :495
  o.l = x_mul(f1, im_ok ? m.l : me.l);  // :415 / :500
  o.u = x_mul(f2, ip_ok ? p.u : me.u);
  o.d = x_add(me.d, x_add(x_mul(f1, im_ok ? m.u : o.u), x_mul(f2, ip_ok ? p.l : o.l)));
  return o;
}
// One row of an up-sweep level (:446-452 / :532-540); the left neighbour always exists.
template <typename R, int G>
__device__ __forceinline__ void crw_up(CrwRow<R, G>& me, const R (&xm)[G], const R (&xp)[G], bool ip_ok) {
  const R rd = rcp_refined(me.d);  // the part of the IEEE quotients by d(i) that all right-hand sides share (common.cuh)
#pragma unroll
  for (int g = 0; g < G; ++g) {
    R f = 0;
    f = x_add(f, x_mul(me.l, xm[g]));
    f = x_add(f, ip_ok ? x_mul(me.u, xp[g]) : (R)0);
    me.x[g] = div_given_rcp(x_sub(me.x[g], f), me.d, rd);
  }
}

// LM: level-major operands dl,d,du [nlev][a_stride], X [nrhs][nlev][x_stride]; else the reference layout
// dl,d,du [ncol][a_stride], X [ncol][x_stride] with X(lev, rhs) at lev*nrhs + rhs.
template <typename R, int RPT, int G, bool LM>
__global__ void __launch_bounds__(128)
k_tridiag_cr_warp(const CrwArgs a) {
  const int lane = threadIdx.x & 31;
  const int64_t gw = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t col = gw / a.ngroups;
  if (col >= a.ncol) return;  // whole warps only
  const int j0 = (int)(gw - col * a.ngroups) * G;
  const int nrow = a.nlev, nrhs = a.nrhs;
  const R* gl = static_cast<const R*>(a.dl);
  const R* gd = static_cast<const R*>(a.d);
  const R* gu = static_cast<const R*>(a.du);
  R* gx = static_cast<R*>(a.X);
  auto aoff = [&](int i) { return LM ? (size_t)i * a.a_stride + col : (size_t)col * a.a_stride + i; };
  auto xoff = [&](int i, int j) {
    return LM ? ((size_t)j * nrow + i) * a.x_stride + col : (size_t)col * a.x_stride + (size_t)i * nrhs + j;
  };

  CrwRow<R, G> row[RPT];
#pragma unroll
  for (int k = 0; k < RPT; ++k) {
    const int i = RPT * lane + k;
    const bool ok = i < nrow;  // rows beyond the column: never active, never a valid neighbour
    row[k].l = ok ? __ldg(gl + aoff(i)) : (R)0;
    row[k].d = ok ? __ldg(gd + aoff(i)) : (R)1;
    row[k].u = ok ? __ldg(gu + aoff(i)) : (R)0;
#pragma unroll
    for (int g = 0; g < G; ++g) row[k].x[g] = (ok && j0 + g < nrhs) ? gx[xoff(i, j0 + g)] : (R)0;
  }

  // ------------------------------------------------------------------ way down :401-423
  int os = 1;
#pragma unroll
  for (int ls = 1; ls < RPT; ls <<= 1) {  // offsets inside a lane
    if (2 * ls < nrow) {
      const CrwRow<R, G> hl = crw_shfl_up(row[RPT - ls], 1);  // row RPT*lane - ls
#pragma unroll
      for (int k = 0; k < RPT; k += 2 * ls) {
        const int i = RPT * lane + k;
        const bool im_ok = i - ls >= 0, ip_ok = i + ls < nrow;
        const CrwRow<R, G> n = crw_down<R, G>(row[k], k == 0 ? hl : row[k >= ls ? k - ls : 0], row[k + ls], im_ok, ip_ok);
        if (i < nrow) row[k] = n;
      }
      os = 2 * ls;
    }
  }
#pragma unroll
  for (int s = 1; s < 32; s <<= 1) {  // offsets of s lanes: local row 0 of the lanes that are multiples of 2s
    if (2 * RPT * s < nrow) {
      const CrwRow<R, G> m = crw_shfl_up(row[0], s), p = crw_shfl_down(row[0], s);
      const int i = RPT * lane;
      const bool im_ok = lane >= s, ip_ok = lane + s < 32 && i + RPT * s < nrow;
      const CrwRow<R, G> n = crw_down<R, G>(row[0], m, p, im_ok, ip_ok);
      if ((lane & (2 * s - 1)) == 0 && i < nrow) row[0] = n;
      os = 2 * RPT * s;
    }
  }
  // `os` is now the first offset that is not a level: rows 0 and os are what is left (:427-437 / :508-521)
  {
    const CrwRow<R, G> r0 = crw_bcast(row[0], 0);
    if (os >= nrow) {
      if (lane == 0) {
#pragma unroll
        for (int g = 0; g < G; ++g) row[0].x[g] = x_div(r0.x[g], r0.d);
      }
    } else {
      // row `os`: local row os of lane 0 when os < RPT, else local row 0 of lane os / RPT
      CrwRow<R, G> mine = row[0];
#pragma unroll
      for (int k = 1; k < RPT; ++k)
        if (k == os) mine = row[k];
      const int src = os < RPT ? 0 : os / RPT;
      const CrwRow<R, G> rs = crw_bcast(mine, src);
      const R det = x_sub(x_mul(r0.d, rs.d), x_mul(r0.u, rs.l));
      const R rdet = rcp_refined(det);
      R xa[G], xb[G];
#pragma unroll
      for (int g = 0; g < G; ++g) {
        xa[g] = div_given_rcp(x_sub(x_mul(rs.d, r0.x[g]), x_mul(r0.u, rs.x[g])), det, rdet);
        xb[g] = div_given_rcp(x_sub(x_mul(r0.d, rs.x[g]), x_mul(rs.l, r0.x[g])), det, rdet);
      }
      if (lane == 0) {
#pragma unroll
        for (int g = 0; g < G; ++g) row[0].x[g] = xa[g];
      }
      if (lane == src) {
#pragma unroll
        for (int k = 0; k < RPT; ++k) {
          if (k == (os < RPT ? os : 0)) {
#pragma unroll
            for (int g = 0; g < G; ++g) row[k].x[g] = xb[g];
          }
        }
      }
    }
  }
  // ------------------------------------------------------------------ way up :442-457
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) {
    if (2 * RPT * s <= os) {  // this offset was a level on the way down
      R xm[G], xp[G];
#pragma unroll
      for (int g = 0; g < G; ++g) {
        xm[g] = __shfl_up_sync(0xffffffffu, row[0].x[g], s);
        xp[g] = __shfl_down_sync(0xffffffffu, row[0].x[g], s);
      }
      const int i = RPT * lane;
      const bool ip_ok = lane + s < 32 && i + RPT * s < nrow;
      CrwRow<R, G> n = row[0];
      crw_up<R, G>(n, xm, xp, ip_ok);
      if ((lane & (2 * s - 1)) == s && i < nrow) row[0] = n;
    }
  }
#pragma unroll
  for (int ls = RPT / 2; ls >= 1; ls >>= 1) {
    if (2 * ls <= os) {
      R xr[G];  // local row 0 of the next lane: the right neighbour of the lane's last row of this level
#pragma unroll
      for (int g = 0; g < G; ++g) xr[g] = __shfl_down_sync(0xffffffffu, row[0].x[g], 1);
#pragma unroll
      for (int k = ls; k < RPT; k += 2 * ls) {
        const int i = RPT * lane + k;
        const bool ip_ok = i + ls < nrow && (k + ls < RPT || lane < 31);
        R xp[G];
#pragma unroll
        for (int g = 0; g < G; ++g) xp[g] = k + ls < RPT ? row[k + ls < RPT ? k + ls : 0].x[g] : xr[g];
        CrwRow<R, G> n = row[k];
        crw_up<R, G>(n, row[k - ls].x, xp, ip_ok);
        if (i < nrow) row[k] = n;
      }
    }
  }
#pragma unroll
  for (int k = 0; k < RPT; ++k) {
    const int i = RPT * lane + k;
#pragma unroll
    for (int g = 0; g < G; ++g)
      if (i < nrow && j0 + g < nrhs) gx[xoff(i, j0 + g)] = row[k].x[g];
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Columns longer than one warp holds (257 .. 2,048 levels, one right-hand side): the same scheme with NW = 2, 4 or 8
// warps per column -- "few columns, many levels", the shape cyclic reduction exists for (reference guidance
// src/algorithm/ekat_tridiag.hpp:71-74; the reference's own test runs nrow = 2048,
// tests/algorithm/tridiag_tests_correctness.cpp:410-414). Thread t of the CTA owns rows 8t .. 8t+7 in registers. A level
// needs at most ONE row from another warp: the left neighbour of a warp's lane 0 (on the way down) or the right
// neighbour of its last active lane (on the way up); levels whose offset is a multiple of 256 rows involve lane 0 of
// the warps only. That one row per warp and level goes through a small shared-memory array (two copies, alternating by
// level, so one __syncthreads per level is enough); everything else stays in registers and travels by shuffle.
constexpr int kCtaRPT = 8;

template <typename R, bool LM>
__global__ void __launch_bounds__(256)
k_tridiag_cr_cta(const CrwArgs a) {
  constexpr int RPT = kCtaRPT;
  using Row = CrwRow<R, 1>;
  __shared__ Row edge[2][8];
  __shared__ R xedge[2][8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, t = threadIdx.x, NW = blockDim.x >> 5;
  const int64_t col = blockIdx.x;
  const int nrow = a.nlev;
  const R* gl = static_cast<const R*>(a.dl);
  const R* gd = static_cast<const R*>(a.d);
  const R* gu = static_cast<const R*>(a.du);
  R* gx = static_cast<R*>(a.X);
  auto off = [&](int i, int64_t stride) { return LM ? (size_t)i * stride + col : (size_t)col * stride + i; };

  Row row[RPT];
#pragma unroll
  for (int k = 0; k < RPT; ++k) {
    const int i = RPT * t + k;
    const bool ok = i < nrow;
    row[k].l = ok ? __ldg(gl + off(i, a.a_stride)) : (R)0;
    row[k].d = ok ? __ldg(gd + off(i, a.a_stride)) : (R)1;
    row[k].u = ok ? __ldg(gu + off(i, a.a_stride)) : (R)0;
    row[k].x[0] = ok ? gx[off(i, a.x_stride)] : (R)0;
  }
  int lv = 0, os = 1;
  // ------------------------------------------------------------------ way down :401-423
#pragma unroll
  for (int ls = 1; ls < RPT; ls <<= 1) {  // offsets inside a thread; local row 0 takes its left neighbour from thread t-1
    if (2 * ls < nrow) {
      if (lane == 31) edge[lv & 1][warp] = row[RPT - ls];
      __syncthreads();
      Row hl = crw_shfl_up(row[RPT - ls], 1);
      if (lane == 0 && warp > 0) hl = edge[lv & 1][warp - 1];
#pragma unroll
      for (int k = 0; k < RPT; k += 2 * ls) {
        const int i = RPT * t + k;
        const bool im_ok = i - ls >= 0, ip_ok = i + ls < nrow;
        const Row n = crw_down<R, 1>(row[k], k == 0 ? hl : row[k >= ls ? k - ls : 0], row[k + ls], im_ok, ip_ok);
        if (i < nrow) row[k] = n;
      }
      os = 2 * ls;
      ++lv;
    }
  }
#pragma unroll
  for (int s = 1; s < 32; s <<= 1) {  // offsets of s threads inside a warp
    if (2 * RPT * s < nrow) {
      if (lane == 32 - s) edge[lv & 1][warp] = row[0];
      __syncthreads();
      Row m = crw_shfl_up(row[0], s);
      const Row p = crw_shfl_down(row[0], s);
      if (lane == 0 && warp > 0) m = edge[lv & 1][warp - 1];
      const int i = RPT * t;
      const bool im_ok = t >= s, ip_ok = lane + s < 32 && i + RPT * s < nrow;
      const Row n = crw_down<R, 1>(row[0], m, p, im_ok, ip_ok);
      if ((lane & (2 * s - 1)) == 0 && i < nrow) row[0] = n;
      os = 2 * RPT * s;
      ++lv;
    }
  }
  for (int sw = 1; sw < NW; sw <<= 1) {  // offsets of sw warps: lane 0 of the warps that are multiples of 2 sw
    if (2 * RPT * 32 * sw < nrow) {
      if (lane == 0) edge[lv & 1][warp] = row[0];
      __syncthreads();
      const int i = RPT * t;
      if (lane == 0 && (warp & (2 * sw - 1)) == 0 && i < nrow) {
        const bool im_ok = warp >= sw, ip_ok = warp + sw < NW && i + RPT * 32 * sw < nrow;
        const Row m = edge[lv & 1][im_ok ? warp - sw : warp], p = edge[lv & 1][ip_ok ? warp + sw : warp];
        row[0] = crw_down<R, 1>(row[0], m, p, im_ok, ip_ok);
      }
      os = 2 * RPT * 32 * sw;
      ++lv;
    }
  }
  // ------------------------------------------------------------------ rows 0 and os are what is left (:427-437)
  {
    const int ts = os / RPT, ks = os % RPT;  // owner of row os
    if (t == 0) edge[lv & 1][0] = row[0];
    if (os < nrow && t == ts) {
      Row mine = row[0];
#pragma unroll
      for (int k = 1; k < RPT; ++k)
        if (k == ks) mine = row[k];
      edge[lv & 1][1] = mine;
    }
    __syncthreads();
    if (t == 0 || (os < nrow && t == ts)) {
      const Row r0 = edge[lv & 1][0];
      if (os >= nrow) {
        row[0].x[0] = x_div(r0.x[0], r0.d);
      } else {
        const Row rs = edge[lv & 1][1];
        const R det = x_sub(x_mul(r0.d, rs.d), x_mul(r0.u, rs.l));
        const R xa = x_div(x_sub(x_mul(rs.d, r0.x[0]), x_mul(r0.u, rs.x[0])), det);
        const R xb = x_div(x_sub(x_mul(r0.d, rs.x[0]), x_mul(rs.l, r0.x[0])), det);
        if (t == 0) row[0].x[0] = xa;
        if (t == ts) {
#pragma unroll
          for (int k = 0; k < RPT; ++k)
            if (k == ks) row[k].x[0] = xb;
        }
      }
    }
    ++lv;
  }
  // ------------------------------------------------------------------ way up :442-457 (only solution values travel)
  for (int sw = NW >> 1; sw >= 1; sw >>= 1) {
    if (2 * RPT * 32 * sw <= os) {
      if (lane == 0) xedge[lv & 1][warp] = row[0].x[0];
      __syncthreads();
      const int i = RPT * t;
      if (lane == 0 && (warp & (2 * sw - 1)) == sw && i < nrow) {
        const bool ip_ok = warp + sw < NW && i + RPT * 32 * sw < nrow;
        const R xm[1] = {xedge[lv & 1][warp - sw]}, xp[1] = {xedge[lv & 1][ip_ok ? warp + sw : warp]};
        crw_up<R, 1>(row[0], xm, xp, ip_ok);
      }
      ++lv;
    }
  }
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) {
    if (2 * RPT * s <= os) {
      if (lane == 0) xedge[lv & 1][warp] = row[0].x[0];  // the right neighbour of a warp's lane 32-s is the next warp's lane 0
      __syncthreads();
      R xm[1], xp[1];
      xm[0] = __shfl_up_sync(0xffffffffu, row[0].x[0], s);
      xp[0] = __shfl_down_sync(0xffffffffu, row[0].x[0], s);
      if (lane + s >= 32) xp[0] = xedge[lv & 1][warp + 1 < NW ? warp + 1 : warp];
      const int i = RPT * t;
      const bool ip_ok = i + RPT * s < nrow;
      Row n = row[0];
      crw_up<R, 1>(n, xm, xp, ip_ok);
      if ((lane & (2 * s - 1)) == s && i < nrow) row[0] = n;
      ++lv;
    }
  }
#pragma unroll
  for (int ls = RPT / 2; ls >= 1; ls >>= 1) {
    if (2 * ls <= os) {
      if (lane == 0) xedge[lv & 1][warp] = row[0].x[0];
      __syncthreads();
      R xr[1];  // local row 0 of the next thread: the right neighbour of the thread's last row of this level
      xr[0] = __shfl_down_sync(0xffffffffu, row[0].x[0], 1);
      if (lane == 31) xr[0] = xedge[lv & 1][warp + 1 < NW ? warp + 1 : warp];
#pragma unroll
      for (int k = ls; k < RPT; k += 2 * ls) {
        const int i = RPT * t + k;
        const bool ip_ok = i + ls < nrow;
        R xp[1];
        xp[0] = k + ls < RPT ? row[k + ls < RPT ? k + ls : 0].x[0] : xr[0];
        Row n = row[k];
        crw_up<R, 1>(n, row[k - ls].x, xp, ip_ok);
        if (i < nrow) row[k] = n;
      }
      ++lv;
    }
  }
#pragma unroll
  for (int k = 0; k < RPT; ++k) {
    const int i = RPT * t + k;
    if (i < nrow) gx[off(i, a.x_stride)] = row[k].x[0];
  }
}

template <typename R, int RPT, int G, bool LM>
static int launch_cr_warp(const CrwArgs& a, cudaStream_t stream) {
  const int64_t nwarps = a.ncol * a.ngroups;
  const int64_t grid = (nwarps + 3) / 4;
  if (grid >= (int64_t)1 << 31) return EKAT_B200_ERR_UNSUPPORTED;
  k_tridiag_cr_warp<R, RPT, G, LM><<<(unsigned)grid, 128, 0, stream>>>(a);
  note_launch(sizeof(R) == 8 ? (LM ? "tridiag_cr_warp<f64>" : "tridiag_cr_warp_cm<f64>") : (LM ? "tridiag_cr_warp<f32>" : "tridiag_cr_warp_cm<f32>"));
  return check_cuda(cudaGetLastError());
}

// Up to 256 levels: one warp per (column, rhs group); up to 2,048 levels and one right-hand side: 2 .. 8 warps per
// column. Returns EKAT_B200_ERR_UNSUPPORTED beyond that.
template <typename R>
int tridiag_cr_warp(R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nrhs, int64_t a_stride, int64_t x_stride,
                    int layout, cudaStream_t stream) {
  const bool lm = layout == EKAT_B200_LAYOUT_LEV_MAJOR;
  if (nlev > 256) {  // several warps per column
    if (nlev > 256 * kCtaRPT || nrhs != 1 || ncol >= (int64_t)1 << 31) return EKAT_B200_ERR_UNSUPPORTED;
    int nw = 2;
    while (nw * 32 * kCtaRPT < nlev) nw *= 2;
    CrwArgs a;
    a.dl = dl; a.d = d; a.du = du; a.X = X;
    a.ncol = ncol; a.a_stride = a_stride; a.x_stride = x_stride;
    a.nlev = nlev; a.nrhs = 1; a.ngroups = 1;
    if (lm) k_tridiag_cr_cta<R, true><<<(unsigned)ncol, 32 * nw, 0, stream>>>(a);
    else k_tridiag_cr_cta<R, false><<<(unsigned)ncol, 32 * nw, 0, stream>>>(a);
    note_launch(sizeof(R) == 8 ? (lm ? "tridiag_cr_cta<f64>" : "tridiag_cr_cta_cm<f64>") : (lm ? "tridiag_cr_cta<f32>" : "tridiag_cr_cta_cm<f32>"));
    return check_cuda(cudaGetLastError());
  }
  // level-major operands are read with a stride here: with one right-hand side that only pays for small batches
  // (B200, f64, 128 levels: 2,048 columns 11 us against the tile kernel's 15; 4,736 columns 24 against 16)
  if (lm && nrhs == 1 && (ncol > 2048 || nlev > 128)) return EKAT_B200_ERR_UNSUPPORTED;
  const int rpt = nlev <= 64 ? 2 : (nlev <= 128 ? 4 : 8);
  const int gmax = rpt == 8 ? 4 : 8, G = nrhs == 1 ? 1 : gmax;
  CrwArgs a;
  a.dl = dl; a.d = d; a.du = du; a.X = X;
  a.ncol = ncol; a.a_stride = a_stride; a.x_stride = x_stride;
  a.nlev = nlev; a.nrhs = nrhs; a.ngroups = (nrhs + G - 1) / G;
#define EKB_CRW(RPT_, G_) (lm ? launch_cr_warp<R, RPT_, G_, true>(a, stream) : launch_cr_warp<R, RPT_, G_, false>(a, stream))
  if (rpt == 2) return G == 1 ? EKB_CRW(2, 1) : EKB_CRW(2, 8);
  if (rpt == 4) return G == 1 ? EKB_CRW(4, 1) : EKB_CRW(4, 8);
  return G == 1 ? EKB_CRW(8, 1) : EKB_CRW(8, 4);
#undef EKB_CRW
}

template int tridiag_cr_warp<double>(double*, double*, double*, double*, int64_t, int, int, int64_t, int64_t, int, cudaStream_t);
template int tridiag_cr_warp<float>(float*, float*, float*, float*, int64_t, int, int, int64_t, int64_t, int, cudaStream_t);

}  // namespace ekb
