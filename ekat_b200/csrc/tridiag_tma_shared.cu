// tridiag_tma_shared.cu -- batched Thomas / bfb with ONE matrix per column shared by all its right-hand sides
// (nA == 1; nrhs >= 1), TMA-fed, persistent warp-specialised CTAs. Native level-major operands; with one
// right-hand side also the reference's column-major operands (CM).
//
// Replaces the per-column loops around ekat::tridiag::thomas(team, dl, d, du, X[nrow][nrhs])
// (reference src/algorithm/ekat_tridiag.hpp:177-214, :323-334: factorize once in Kokkos::single, barrier, one
// solve per right-hand side) and ekat::tridiag::bfb with rank-1 A (:292-320, :632-648) -- BASELINE config 4
// (10 tracers diffused by one SHOC-style matrix per column) and, with nrhs == 1, the thomas of configs 1, 3, 5.
//
// Why this shape. Per column-level the kernel moves 3 + 2*nrhs words and does ~5*nrhs flops, so it is HBM-bound
// as long as the forward sweep's intermediates never leave the chip: a tile of 32 columns x 128 levels needs one
// plane (32 KB, f64) per right-hand side plus one for the factor, i.e. five right-hand sides fill an SM. With one
// or two CTAs per SM nothing hides a tile's load latency unless the CTA itself overlaps tiles, hence:
//  * persistent CTAs walk work items (column tile, group of rh right-hand sides);
//  * warp roles: rh solver warps (one right-hand side each, lane = column), one factor warp, two producer warps
//    (one elected lane each) that only issue TMA copies, and -- block size permitting -- one io warp per solver
//    warp that issues its tensor stores and refills (a TMA issue costs the issuing thread ~130 cycles);
//    every wait in the kernel is an mbarrier wait;
//  * the factor warp turns (dl, d) into (m, r) [thomas] or (dl) into (m) [bfb] in place in a ring stage that the
//    TMA filled and hands it to the solver warps through a second mbarrier; du becomes c' = du*r [thomas] / d
//    becomes the pivots d' [bfb] in place in a plane that stays until back-substitution. With factor and solve
//    chains on different warps a single right-hand side already runs at 90% of the HBM peak (thomas, f64);
//  * every solver warp keeps its right-hand side in a plane: B lands there by TMA, the forward sweep overwrites it
//    with y*r [thomas] / x' [bfb], the back-substitution overwrites it with the solution, which leaves by TMA
//    tensor stores 8 levels at a time. As soon as a chunk's store has been read out, the load of the NEXT work
//    item's chunk goes into the same slot; the slot order is mirrored from one item to the next (see below), so
//    that every chunk is requested a whole sweep before it is needed;
//  * thomas: the factor warp runs the continuant ("ratio") recurrence (no division on its dependent chain, see
//    tridiag_tma.cu); solver warps have one FMA per level on their chains in both sweeps.
//  * bfb: the reference's operation order with separately rounded * - / (bit-identical); du for the
//    back-substitution is read again from L2 through the ring instead of occupying a plane, and the ring's other
//    half carries the refined reciprocals of the pivots, computed once by the factor warp for all solver warps.
// HBM traffic is the algorithmic minimum (ncu: 8.09 GB moved for 8.14 GB algorithmic at 345,600 x 128 x 10); when
// the right-hand sides do not fit one CTA, the nrhs/rh groups of a tile run on neighbouring CTAs at the same time
// and share dl, d, du through L2. The hand-off protocol is checked by tools/sim_shared_protocol.py.
#include "tridiag_tma.cuh"

namespace ekb {

// bfb factor sweep over one chunk held in registers (ekat_tridiag.hpp:292-303): m = dl(i)/d'(i-1); d'(i) = d(i) - m*du(i-1).
template <typename R, int CPT, int LB, bool SLOW>
__device__ __forceinline__ void bfb_factor_chunk(const Vec<R, CPT> (&a)[LB], const Vec<R, CPT> (&b)[LB], const Vec<R, CPT> (&c)[LB],
                                                 Vec<R, CPT> (&om)[LB], Vec<R, CPT> (&od)[LB], R (&dprev)[CPT], R (&uprev)[CPT],
                                                 bool first, bool& bad) {
  R rp[CPT];  // ~ 1 / d'(i-1): reciprocal of the PREDICTED pivot, off the chain (common.cuh bfb_pivot_step)
  if (!SLOW) {
#pragma unroll
    for (int j = 0; j < CPT; ++j) rp[j] = rcp_seed(dprev[j]);
  }
#pragma unroll
  for (int u = 0; u < LB; ++u) {
#pragma unroll
    for (int j = 0; j < CPT; ++j) {
      const R av = (u == 0 && first) ? (R)0 : a[u].v[j];  // dl(0) is never used (:24-27)
      R m, dc;
      if (SLOW) {
        m = (u == 0 && first) ? (R)0 : x_div(av, dprev[j]);
        dc = x_sub(b[u].v[j], x_mul(m, uprev[j]));        // d(i) -= dl(i)*du(i-1)   :301
      } else {
        dc = dprev[j];
        m = bfb_factor_level(av, b[u].v[j], uprev[j], dc, rp[j], bad);
      }
      om[u].v[j] = m;
      od[u].v[j] = dc;
      dprev[j] = dc;
      uprev[j] = c[u].v[j];
    }
  }
}

// Phase trace for tuning (variant kVarTrace = 7 of ekat_b200_set_tuning selects the instrumented instantiation): per traced CTA
// (blockIdx 0 and gridDim/2) and work item < kTraceItems, 16 counters in clock64 cycles -- see trace_put call sites.
constexpr int kTraceItems = 64, kTraceWords = 16;
__device__ long long g_trace[2 * kTraceItems * kTraceWords];
__device__ __forceinline__ void trace_put(int it, int idx, long long v) {
  const int cs = blockIdx.x == 0 ? 0 : (blockIdx.x == gridDim.x / 2 ? 1 : -1);
  if (cs >= 0 && it < kTraceItems) g_trace[(cs * kTraceItems + it) * kTraceWords + idx] = v;
}

// grid = persistent CTAs; block = 32 * (rh + 3): warps [0, rh) solve, warp rh factorises, warp rh+1 feeds the ring,
// warp rh+2 feeds the factor plane.
// Shared memory: barriers | PC [nslot][LB][32] | PX [rh][nslot][LB][32] | RING [NS][2][LB][32]   (units of V)
//   Chunk k of a plane lives in slot k for even work items and in slot nch-1-k for odd ones: the back-substitution
//   frees slots top chunk first, which is exactly the order in which the next item's forward sweep needs them, so
//   every chunk of the next item is requested a whole sweep before it is used and no slot is double-buffered.
template <typename R, bool EXACT, int CPT, int RN, bool DBG, bool CM>
__global__ void __launch_bounds__(512, 1)
k_tridiag_tma_shared(const __grid_constant__ CUtensorMap mdl, const __grid_constant__ CUtensorMap md,
                     const __grid_constant__ CUtensorMap mdu, const __grid_constant__ CUtensorMap mx,
                     const __grid_constant__ CUtensorMap mxin, R* __restrict__ X, int64_t ncol, int nlev, int64_t x_stride, int rh,
                     int ngroups, int nwork, int NS, int io) {
  using V = Vec<R, CPT>;
  constexpr int LB = kLB, W = 32 * CPT, CH = LB * 32;
  constexpr unsigned CHUNK_BYTES = LB * W * sizeof(R);
  static_assert(LB % RN == 0, "renormalisation interval must divide the chunk");
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nfull = nlev / LB, ntail = nlev - nfull * LB, nch = nfull + (ntail ? 1 : 0), nslot = nch;
  const int UPT = EXACT ? 2 * nch : nch;  // ring uses per work item

  uint64_t* raw = reinterpret_cast<uint64_t*>(smem_raw);  // [NS]  TMA landed in ring stage
  uint64_t* ready = raw + NS;                             // [NS]  factor warp transformed the stage
  uint64_t* freeb = ready + NS;                           // [NS]  all solver warps have read the stage
  uint64_t* plc = freeb + NS;                             // [nslot] TMA landed in factor-plane slot
  uint64_t* cfree = plc + nslot;                          // [nslot] all solver warps have read the slot (back-substitution)
  uint64_t* plx = cfree + nslot;                          // [rh][nslot] TMA landed in a right-hand-side slot
  uint64_t* xdone = plx + rh * nslot;                     // [rh][nslot] solver warp finished a slot: solution ready to be stored
  constexpr size_t AL = CM ? 1024 : 128;  // column-major chunks are swizzled: they sit on 1 KB boundaries (see t_ld)
  const size_t off = ((size_t)(3 * NS + (2 + 2 * rh) * nslot) * 8 + AL - 1) & ~(AL - 1);
  V* PC = reinterpret_cast<V*>(smem_raw + off);
  V* PX = PC + (size_t)nslot * CH;
  V* RING = PX + (size_t)rh * nslot * CH;

  if (threadIdx.x == 0) {
    for (int s = 0; s < NS; ++s) {
      mbar_init(&raw[s], 1);
      mbar_init(&ready[s], 1);
      mbar_init(&freeb[s], rh);
    }
    for (int k = 0; k < nslot; ++k) {
      mbar_init(&plc[k], 1);
      mbar_init(&cfree[k], rh);
    }
    for (int k = 0; k < 2 * rh * nslot; ++k) mbar_init(&plx[k], 1);  // plx and xdone
    fence_mbar_init();
  }
  __syncthreads();

  const int nit = (nwork - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;  // work items of this CTA
  auto slot = [&](int k, int it) { return (it & 1) ? nch - 1 - k : k; };
  auto par = [&](int, int it) { return (unsigned)(it & 1); };  // every slot is filled once per work item
  // work item -> (first column of its tile, first right-hand side of its group); nwork < 2^30, so 32-bit arithmetic
  auto item_c0 = [&](int it) { return (int)(((unsigned)blockIdx.x + (unsigned)it * gridDim.x) / (unsigned)ngroups) * W; };
  auto item_j0 = [&](int it) { return (int)(((unsigned)blockIdx.x + (unsigned)it * gridDim.x) % (unsigned)ngroups) * rh; };
  // ring position of a role: stage and phase parity of the use it is at, advanced use by use (no division per chunk)
  int rs = 0;
  unsigned rph = 0;
  auto ring_advance = [&]() {
    if (++rs == NS) {
      rs = 0;
      rph ^= 1u;
    }
  };
  const uint64_t pol_stream = l2_policy_evict_first();
  // CM: the reference's column-major operands (one right-hand side): boxes are 32 columns x LB levels, coordinates
  // (level, column) instead of (column, level); element access through t_ld / t_st (tridiag_tma.cuh)
  auto ld = [&](const V* chunk, int u) { return t_ld<R, CPT, LB, CM>(chunk, u, lane); };
  auto stv = [&](V* chunk, int u, const V& v) { t_st<R, CPT, LB, CM>(chunk, u, lane, v); };
  auto tload = [&](void* dst, const CUtensorMap* m, int col, int row, uint64_t* bar) {
    if (CM) tma_load_2d(dst, m, row, col, bar);
    else tma_load_2d(dst, m, col, row, bar);
  };
  auto tload_hint = [&](void* dst, const CUtensorMap* m, int col, int row, uint64_t* bar, uint64_t pol) {
    if (CM) tma_load_2d_hint(dst, m, row, col, bar, pol);
    else tma_load_2d_hint(dst, m, col, row, bar, pol);
  };
  auto tstore = [&](const CUtensorMap* m, int col, int row, const void* src) {
    if (CM) tma_store_2d(m, row, col, src);
    else tma_store_2d(m, col, row, src);
  };

  // right-hand side jw of the CTA: its plane, its barriers, and the copy of chunk k of work item it2 into its slot.
  // The copies and the tensor stores of a right-hand side are issued by one thread: lane 0 of its io warp
  // (warp rh+3+jw) when the CTA has io warps, else lane 0 of the solver warp itself. An issue costs that thread
  // ~130 cycles (measured), which is why it is taken off the solver warps whenever the block size allows.
  const int jw = warp < rh ? warp : (warp >= rh + 3 ? warp - rh - 3 : 0);
  V* PXj = PX + (size_t)jw * nslot * CH;
  uint64_t* plxj = plx + (size_t)jw * nslot;
  uint64_t* xdonej = xdone + (size_t)jw * nslot;
  auto load_x = [&](int it2, int k, int c0_2, int xrow_2) {
    const int sl = slot(k, it2);
    mbar_expect_tx(&plxj[sl], CHUNK_BYTES);
    tload_hint(PXj + (size_t)sl * CH, &mxin, c0_2, xrow_2 + k * LB, &plxj[sl], pol_stream);  // B (mxin) in, X (mx) out
  };
  constexpr int kSD = 3;  // a slot is refilled kSD steps after its tensor store was committed (wait_group.read<kSD>)
  const int jb0 = ntail ? 1 : 0;  // the partial top chunk (back-substitution step 0) leaves by plain stores

  if (warp < rh) {
    // ================================================================ solver warp: one right-hand side
    if (lane == 0 && !io) {
      for (int k = 0; k < nch; ++k) load_x(0, k, item_c0(0), (item_j0(0) + jw) * nlev);
    }
    for (int it = 0; it < nit; ++it) {
      const int c0 = item_c0(it), xrow0 = (item_j0(it) + jw) * nlev;
      const bool has1 = it + 1 < nit;
      const int c0_n1 = item_c0(it + 1), xrow_n1 = (item_j0(it + 1) + jw) * nlev;
      auto refill = [&](int jb) {  // lane 0: the slot of this item's chunk nch-1-jb is free again: it is the next item's chunk jb
        if (has1) load_x(it + 1, jb, c0_n1, xrow_n1);
      };
      // ------------------------------------------------------------ forward sweep
      R yprev[CPT];
#pragma unroll
      for (int j = 0; j < CPT; ++j) yprev[j] = (R)0;
      long long t0 = 0, tl = 0, acc[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
      if (DBG) t0 = tl = clock64();
#define EKB_TICK(i)                  \
  if (DBG) {                         \
    const long long n_ = clock64();  \
    acc[i] += n_ - tl;               \
    tl = n_;                         \
  }
      for (int k = 0; k < nch; ++k) {
        const int sl = slot(k, it), s = rs;
        mbar_wait(&plxj[sl], par(k, it));
        EKB_TICK(0)
        mbar_wait(&ready[s], rph);
        EKB_TICK(1)
        ring_advance();
        const V* rm = RING + (size_t)s * 2 * CH;
        V* px = PXj + (size_t)sl * CH;
        V m[LB], r[LB], x[LB];
#pragma unroll
        for (int u = 0; u < LB; ++u) {
          m[u] = ld(rm, u);
          if (!EXACT) r[u] = ld(rm + CH, u);
          x[u] = ld(px, u);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&freeb[s]);
        EKB_TICK(2)
#pragma unroll
        for (int u = 0; u < LB; ++u) {
#pragma unroll
          for (int j = 0; j < CPT; ++j) {
            if (EXACT) {
              yprev[j] = x_sub(x[u].v[j], x_mul(m[u].v[j], yprev[j]));  // X(i) -= dl(i)*X(i-1)   :315-316
              x[u].v[j] = yprev[j];
            } else {
              yprev[j] = f_fma(-m[u].v[j], yprev[j], x[u].v[j]);
              x[u].v[j] = yprev[j] * r[u].v[j];
            }
          }
        }
#pragma unroll
        for (int u = 0; u < LB; ++u) stv(px, u, x[u]);
        EKB_TICK(3)
      }
      // ------------------------------------------------------------ back substitution
      R xv[CPT];
#pragma unroll
      for (int j = 0; j < CPT; ++j) xv[j] = (R)0;
      for (int jb = 0; jb < nch; ++jb) {
        const int k = nch - 1 - jb, sl = slot(k, it);
        int s = 0;
        if (EXACT) {
          s = rs;
          mbar_wait(&ready[s], rph);  // passed on by the factor warp: every ring use completes raw, ready, free in turn
          mbar_wait(&raw[s], rph);    // (already complete; direct acquire of the TMA's writes)
          ring_advance();
        }
        EKB_TICK(4)
        const V* bu = RING + (size_t)s * 2 * CH;  // EXACT: du of this chunk, re-read through L2
        const V* pc = PC + (size_t)sl * CH;
        V* px = PXj + (size_t)sl * CH;
        V y[LB], c[LB], ui[LB], rr[LB];
#pragma unroll
        for (int u = 0; u < LB; ++u) {
          y[u] = ld(px, u);
          c[u] = ld(pc, u);
          if (EXACT) {
            ui[u] = ld(bu, u);
            rr[u] = ld(bu + CH, u);  // rcp_refined(d') from the factor warp
          }
        }
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(&cfree[sl]);
          if (EXACT) mbar_arrive(&freeb[s]);
        }
        EKB_TICK(5)
        if (k == nfull) {
          // partial top chunk (ntail levels): a tensor store would spill into the next right-hand side -> plain stores
          const int64_t col = (int64_t)c0 + lane * CPT;
#pragma unroll
          for (int u = LB - 1; u >= 0; --u) {
            if (u < ntail) {
              V o;
#pragma unroll
              for (int j = 0; j < CPT; ++j) {
                if (EXACT) {
                  const R uu = (u == ntail - 1) ? (R)0 : ui[u].v[j];  // du(nrow-1) is never used (:24-27)
                  xv[j] = x_div(x_sub(y[u].v[j], x_mul(uu, xv[j])), c[u].v[j]);  // :317-319
                } else {
                  xv[j] = f_fma(-c[u].v[j], xv[j], y[u].v[j]);
                }
                o.v[j] = xv[j];
              }
              if (CM) {  // column-major: a column's levels are contiguous (only this ragged chunk goes uncoalesced)
#pragma unroll
                for (int j = 0; j < CPT; ++j) {
                  const int64_t cj = (int64_t)c0 + j * 32 + lane;
                  if (cj < ncol) __stcs(X + (size_t)cj * x_stride + (k * LB + u), o.v[j]);
                }
              } else {
                R* g = X + (size_t)(xrow0 + k * LB + u) * x_stride + col;
#pragma unroll
                for (int j = 0; j < CPT; ++j)
                  if (col + j < ncol) __stcs(g + j, o.v[j]);
              }
            }
          }
          if (lane == 0) {  // every lane read its part of the slot before the __syncwarp above
            if (io) mbar_arrive(&xdonej[sl]);
            else refill(jb);
          }
        } else {
          if (EXACT) {
            if (jb == 0) {  // this chunk ends with the last level
#pragma unroll
              for (int j = 0; j < CPT; ++j) ui[LB - 1].v[j] = (R)0;
            }
            R sxv[CPT];
            V ys[LB];
#pragma unroll
            for (int j = 0; j < CPT; ++j) sxv[j] = xv[j];
#pragma unroll
            for (int u = 0; u < LB; ++u) ys[u] = y[u];
            bool bad = false;
#pragma unroll
            for (int u = LB - 1; u >= 0; --u) {
#pragma unroll
              for (int j = 0; j < CPT; ++j) {
                const R num = x_sub(y[u].v[j], x_mul(ui[u].v[j], xv[j]));  // X(i-1) - du(i-1)*X(i)   :318-319
                xv[j] = div_spec(num, c[u].v[j], rr[u].v[j], bad);
                y[u].v[j] = xv[j];
              }
            }
            if (__any_sync(0xffffffffu, bad)) {
#pragma unroll
              for (int j = 0; j < CPT; ++j) xv[j] = sxv[j];
#pragma unroll
              for (int u = 0; u < LB; ++u) y[u] = ys[u];
              bfb_bwd_chunk<R, CPT, LB, true>(y, c, ui, xv, bad);
            }
          } else {
#pragma unroll
            for (int u = LB - 1; u >= 0; --u) {
#pragma unroll
              for (int j = 0; j < CPT; ++j) {
                xv[j] = f_fma(-c[u].v[j], xv[j], y[u].v[j]);
                y[u].v[j] = xv[j];
              }
            }
          }
#pragma unroll
          for (int u = 0; u < LB; ++u) stv(px, u, y[u]);
          EKB_TICK(6)
          fence_proxy_async();  // generic-proxy writes of this chunk -> visible to the TMA store
          __syncwarp();
          EKB_TICK(7)
          if (lane == 0 && io) mbar_arrive(&xdonej[sl]);  // the io warp stores the slot and refills it
          if (lane == 0 && !io) {
            tstore(&mx, c0, xrow0 + k * LB, PXj + (size_t)sl * CH);
            EKB_TICK(8)
            tma_store_commit();
            EKB_TICK(9)
            if (jb - jb0 >= kSD) {
              tma_store_wait_read<kSD>();  // the store committed kSD steps ago has read its slot
              EKB_TICK(10)
              refill(jb - kSD);
            }
            EKB_TICK(11)
          }
        }
      }
      if (lane == 0 && !io) {
        tma_store_wait_read<0>();
        for (int jb = tmax(jb0, nch - kSD); jb < nch; ++jb) refill(jb);
      }
      __syncwarp();
      if (DBG && jw == 0 && lane == 0) {
        // forward: 0 wait for B chunk, 1 wait for factor warp, 2 shared loads + release, 3 recurrence + shared stores
        // backward: 4 wait for du (bfb), 5 shared loads + release, 6 recurrence + shared stores, 7 proxy fence + warp
        // sync, 8 lane 0: tensor store, wait for the previous store's read, refill
        // (lane 0 only) 8 tensor store issue, 9 commit, 10 wait_group.read, 11 refill (expect_tx + tensor load issue)
        for (int i = 0; i < 12; ++i) trace_put(it, i, acc[i]);
        trace_put(it, 12, t0);  // start time stamp
      }
#undef EKB_TICK
    }
  } else if (warp == rh) {
    // ================================================================ factor warp
    for (int it = 0; it < nit; ++it) {
      // FAST: q1 = q[i-1], q2 = q[i-2] (ratio form), cp = du[i-1], gp = 1/pivot[i-1].  EXACT: q1 = d'[i-1], cp = du[i-1].
      R q1[CPT], q2[CPT], cp[CPT], gp[CPT];
#pragma unroll
      for (int j = 0; j < CPT; ++j) {
        q1[j] = (R)1;
        q2[j] = (R)0;
        cp[j] = (R)0;
        gp[j] = (R)0;
      }
      long long f0 = 0, w_raw = 0, w_plc = 0;
      if (DBG) f0 = clock64();
      for (int k = 0; k < nch; ++k) {
        const int sl = slot(k, it), s = rs;
        long long ta = 0, tb = 0;
        if (DBG) ta = clock64();
        mbar_wait(&raw[s], rph);
        ring_advance();
        if (DBG) tb = clock64();
        mbar_wait(&plc[sl], par(k, it));
        if (DBG) {
          w_raw += tb - ta;
          w_plc += clock64() - tb;
        }
        V* r0 = RING + (size_t)s * 2 * CH;
        V* r1 = r0 + CH;
        V* pc = PC + (size_t)sl * CH;
        V a[LB], b[LB], c[LB], om[LB], o1[LB], oc[LB];
#pragma unroll
        for (int u = 0; u < LB; ++u) {
          a[u] = ld(r0, u);
          b[u] = EXACT ? ld(pc, u) : ld(r1, u);
          c[u] = EXACT ? ld(r1, u) : ld(pc, u);
        }
        if (EXACT) {
          R sd[CPT], su[CPT];
#pragma unroll
          for (int j = 0; j < CPT; ++j) {
            sd[j] = q1[j];
            su[j] = cp[j];
          }
          bool bad = false;
          bfb_factor_chunk<R, CPT, LB, false>(a, b, c, om, oc, q1, cp, k == 0, bad);
          if (__any_sync(0xffffffffu, bad)) {  // rare operands somewhere in the warp: redo with the full-range division
#pragma unroll
            for (int j = 0; j < CPT; ++j) {
              q1[j] = sd[j];
              cp[j] = su[j];
            }
            bfb_factor_chunk<R, CPT, LB, true>(a, b, c, om, oc, q1, cp, k == 0, bad);
          }
        } else {
          if (k == 0) {  // dl(0) is never used (ekat_tridiag.hpp:24-27) and may hold anything
#pragma unroll
            for (int j = 0; j < CPT; ++j) a[0].v[j] = (R)0;
          }
#pragma unroll
          for (int u = 0; u < LB; ++u) {
#pragma unroll
            for (int j = 0; j < CPT; ++j) {
              R q;
              if (u % RN == 0) q = f_fma(-(a[u].v[j] * cp[j]), q2[j], b[u].v[j]);  // renormalised state: q[i-1] == 1
              else q = f_fma(b[u].v[j], q1[j], -((a[u].v[j] * cp[j]) * q2[j]));
              const R rq = t_rcp(q);
              const R g = (u % RN == 0) ? rq : q1[j] * rq;  // q[i-1]/q[i] = 1/pivot[i]
              om[u].v[j] = a[u].v[j] * gp[j];               // multiplier dl[i]/pivot[i-1]
              o1[u].v[j] = g;
              oc[u].v[j] = (k * LB + u == nlev - 1) ? (R)0 : c[u].v[j] * g;  // du(nrow-1) is never used (:24-27)
              if ((u + 1) % RN == 0) {
                q2[j] = g;
                q1[j] = (R)1;
              } else {
                q2[j] = (u % RN == 0) ? (R)1 : q1[j];
                q1[j] = q;
              }
              cp[j] = c[u].v[j];
              gp[j] = g;
            }
          }
        }
#pragma unroll
        for (int u = 0; u < LB; ++u) {
          stv(r0, u, om[u]);
          if (!EXACT) stv(r1, u, o1[u]);
          stv(pc, u, oc[u]);
        }
        fence_proxy_async();  // these slots are later overwritten by TMA copies
        __syncwarp();
        if (lane == 0) mbar_arrive(&ready[s]);
      }
      if (DBG && lane == 0) {
        trace_put(it, 13, clock64() - f0);  // factor warp: forward sweep of this item, total
        trace_put(it, 14, w_raw);           //   of which waiting for dl, d
        trace_put(it, 15, w_plc);           //   of which waiting for du / d plane chunks
      }
      if (EXACT) {
        // the back-substitution's du chunks go through the same ring: hand each one on untouched, so that every
        // barrier of a stage sees every use (a waiter must never fall two phases behind)
        // ... and carry, in their second half, the refined reciprocals of this chunk's pivots (the part of the IEEE
        // division that does not depend on the right-hand side, common.cuh), computed once here instead of once per
        // solver warp
        for (int jb = 0; jb < nch; ++jb) {
          const int s = rs, sl = slot(nch - 1 - jb, it);
          const V* pc = PC + (size_t)sl * CH;
          V* r1 = RING + (size_t)s * 2 * CH + CH;
          V dd[LB];
#pragma unroll
          for (int u = 0; u < LB; ++u) dd[u] = ld(pc, u);
#pragma unroll
          for (int u = 0; u < LB; ++u) {
#pragma unroll
            for (int j = 0; j < CPT; ++j) dd[u].v[j] = rcp_refined(dd[u].v[j]);
          }
          mbar_wait(&raw[s], rph);  // du has landed in the first half; the stage is ours until `ready`
          ring_advance();
#pragma unroll
          for (int u = 0; u < LB; ++u) stv(r1, u, dd[u]);
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) mbar_arrive(&ready[s]);
        }
      }
    }
  } else if (warp == rh + 1) {
    // ================================================================ ring producer
    if (lane == 0) {
      bool wrapped = false;
      for (int it = 0; it < nit; ++it) {
        const int c0 = item_c0(it);
        for (int k = 0; k < UPT; ++k) {
          const int s = rs;
          if (wrapped) mbar_wait(&freeb[s], rph ^ 1u);  // the use one lap earlier has been read by every solver warp
          ring_advance();
          if (rs == 0) wrapped = true;
          V* st = RING + (size_t)s * 2 * CH;
          if (k < nch) {
            mbar_expect_tx(&raw[s], 2u * CHUNK_BYTES);
            tload(st, &mdl, c0, k * LB, &raw[s]);
            tload(st + CH, EXACT ? &mdu : &md, c0, k * LB, &raw[s]);
          } else {  // EXACT: du of chunk 2*nch-1-k for the back-substitution (an L2 hit)
            mbar_expect_tx(&raw[s], CHUNK_BYTES);
            tload(st, &mdu, c0, (2 * nch - 1 - k) * LB, &raw[s]);
          }
        }
      }
    }
  } else if (warp == rh + 2) {
    // ================================================================ factor-plane producer
    if (lane == 0) {
      auto load_c = [&](int it2, int k) {
        const int sl = slot(k, it2);
        mbar_expect_tx(&plc[sl], CHUNK_BYTES);
        tload(PC + (size_t)sl * CH, EXACT ? &md : &mdu, item_c0(it2), k * LB, &plc[sl]);
      };
      for (int k = 0; k < nch; ++k) load_c(0, k);
      for (int it = 0; it + 1 < nit; ++it)
        for (int jb = 0; jb < nch; ++jb) {  // the slot of this item's chunk nch-1-jb becomes the next item's chunk jb
          mbar_wait(&cfree[slot(nch - 1 - jb, it)], par(0, it));
          load_c(it + 1, jb);
        }
    }
  } else if (lane == 0) {
    // ================================================================ io warp of right-hand side jw (only if io)
    for (int k = 0; k < nch; ++k) load_x(0, k, item_c0(0), (item_j0(0) + jw) * nlev);
    for (int it = 0; it < nit; ++it) {
      const int c0 = item_c0(it), xrow0 = (item_j0(it) + jw) * nlev;
      const bool has1 = it + 1 < nit;
      const int c0_n1 = item_c0(it + 1), xrow_n1 = (item_j0(it + 1) + jw) * nlev;
      for (int jb = 0; jb < nch; ++jb) {
        const int k = nch - 1 - jb, sl = slot(k, it);
        mbar_wait(&xdonej[sl], par(0, it));
        if (k == nfull) {  // partial top chunk: already stored by the solver warp
          if (has1) load_x(it + 1, jb, c0_n1, xrow_n1);
        } else {
          tstore(&mx, c0, xrow0 + k * LB, PXj + (size_t)sl * CH);
          tma_store_commit();
          if (jb - jb0 >= kSD) {
            tma_store_wait_read<kSD>();
            if (has1) load_x(it + 1, jb - kSD, c0_n1, xrow_n1);
          }
        }
      }
      tma_store_wait_read<0>();
      if (has1)
        for (int jb = tmax(jb0, nch - kSD); jb < nch; ++jb) load_x(it + 1, jb, c0_n1, xrow_n1);
    }
  }
}

// ------------------------------------------------------------------------------------------- host side
static inline int shared_threads(int rh, int io) { return 32 * (rh + 3 + (io ? rh : 0)); }

template <typename R, bool EXACT, bool DBG, bool CM>
static int launch_shared_tma(R* dl, R* d, R* du, R* X, const R* Xin, int64_t ncol, int nlev, int nrhs, int64_t a_stride, int64_t x_stride,
                             int rh, int ns, int io, int per_sm, size_t smem, cudaStream_t stream, const char* name) {
  constexpr int CPT = TmaCfg<R>::CPT, RN = EXACT ? 1 : TmaCfg<R>::RN, W = 32 * CPT;
  CUtensorMap mdl, md, mdu, mx, mxin;
  if (CM) {  // [ncol][stride] rows whose first nlev entries are the column; 64-byte rows, swizzled (f64 only)
    if (!make_tensor_map_2d<R>(&mdl, dl, nlev, ncol, a_stride, kLB, W, 1) || !make_tensor_map_2d<R>(&md, d, nlev, ncol, a_stride, kLB, W, 1) ||
        !make_tensor_map_2d<R>(&mdu, du, nlev, ncol, a_stride, kLB, W, 1) || !make_tensor_map_2d<R>(&mx, X, nlev, ncol, x_stride, kLB, W, 1) ||
        !make_tensor_map_2d<R>(&mxin, Xin, nlev, ncol, x_stride, kLB, W, 1))
      return EKAT_B200_ERR_UNSUPPORTED;
  } else
  if (!make_tensor_map_2d<R>(&mdl, dl, ncol, nlev, a_stride, W, kLB, 0) || !make_tensor_map_2d<R>(&md, d, ncol, nlev, a_stride, W, kLB, 0) ||
      !make_tensor_map_2d<R>(&mdu, du, ncol, nlev, a_stride, W, kLB, 0) ||
      !make_tensor_map_2d<R>(&mx, X, ncol, (int64_t)nrhs * nlev, x_stride, W, kLB, 0) ||
      !make_tensor_map_2d<R>(&mxin, Xin, ncol, (int64_t)nrhs * nlev, x_stride, W, kLB, 0))
    return EKAT_B200_ERR_UNSUPPORTED;
  auto k = k_tridiag_tma_shared<R, EXACT, CPT, RN, DBG, CM>;
  int rc = allow_smem(k, smem);
  if (rc) return rc;
  const int ngroups = nrhs / rh;
  const int64_t ntiles = (ncol + W - 1) / W, nwork = ntiles * ngroups;
  if (nwork >= (int64_t)1 << 30) return EKAT_B200_ERR_UNSUPPORTED;
  const int64_t grid = tmin<int64_t>(nwork, (int64_t)kNumSms * tmax(1, per_sm));
  k<<<(unsigned)grid, shared_threads(rh, io), smem, stream>>>(mdl, md, mdu, mx, mxin, X, ncol, nlev, x_stride, rh, ngroups,
                                                            (int)nwork, ns, io);
  note_launch(name);
  return check_cuda(cudaGetLastError());
}

// Resident CTAs per SM of an instantiation at a block size and dynamic shared-memory size (registers and threads
// bind as well as shared memory). The occupancy query is a driver call, so its answers are remembered per thread.
template <typename R, bool EXACT, bool CM>
static int shared_ctas_per_sm(int threads, size_t smem) {
  struct Memo {
    int threads;
    size_t smem;
    int n;
  };
  static thread_local Memo memo[16] = {};
  static thread_local int used = 0;
  for (int i = 0; i < used; ++i)
    if (memo[i].threads == threads && memo[i].smem == smem) return memo[i].n;
  constexpr int CPT = TmaCfg<R>::CPT, RN = EXACT ? 1 : TmaCfg<R>::RN;
  auto k = k_tridiag_tma_shared<R, EXACT, CPT, RN, false, CM>;
  int n = 0;
  if (allow_smem(k, smem) != EKAT_B200_OK) return 1;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k, threads, smem) != cudaSuccess || n < 1) n = 1;
  if (used < 16) memo[used++] = {threads, smem, n};
  return n;
}

// Returns EKAT_B200_ERR_UNSUPPORTED when the operands do not meet TMA's requirements or a tile does not fit on
// chip; the caller then uses the cp.async kernels.
template <typename R, bool CM>
static int run_shared(bool exact, R* dl, R* d, R* du, R* X, const R* Xin, int64_t ncol, int nlev, int nrhs, int64_t a_stride,
                      int64_t x_stride, cudaStream_t stream) {
  if (CM && (nrhs != 1 || sizeof(R) != 8)) return EKAT_B200_ERR_UNSUPPORTED;  // right-hand sides interleave within a column there
  auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  if (!al(dl) || !al(d) || !al(du) || !al(X) || !al(Xin) || (a_stride * sizeof(R)) % 16 || (x_stride * sizeof(R)) % 16)
    return EKAT_B200_ERR_UNSUPPORTED;
  if (ncol >= (int64_t)1 << 31 || (int64_t)nrhs * nlev >= (int64_t)1 << 31) return EKAT_B200_ERR_UNSUPPORTED;
  constexpr int CPT = TmaCfg<R>::CPT, W = 32 * CPT;
  const Tuning tu = tuning();
  const int nch = (nlev + kLB - 1) / kLB, nslot = nch;
  const size_t chunk = (size_t)kLB * W * sizeof(R);
  auto smem_for = [&](int rh, int ns) {
    const size_t al_b = CM ? 1024 : 128;
    const size_t bars = ((size_t)(3 * ns + (2 + 2 * rh) * nslot) * 8 + al_b - 1) & ~(al_b - 1);
    return bars + (size_t)(1 + rh) * nslot * chunk + (size_t)ns * 2 * chunk;
  };
  // right-hand sides per CTA: the largest divisor of nrhs whose planes fit next to a ring of >= 4 stages (>= 2 if none does)
  int rh = 0, ns = 0;
  for (int want = 4; want >= 2 && !rh; want -= 2)
    for (int cand = tmin(nrhs, 13); cand >= 1; --cand) {
      if (nrhs % cand) continue;
      if (tu.cols_per_cta > 0 && cand != tu.cols_per_cta) continue;  // tuning override: right-hand sides per CTA
      int n = tu.prefetch > 8 ? tmin(tu.prefetch, 16) : 8;
      while (n >= want && smem_for(cand, n) > kMaxSmemPerCta) --n;
      if (n >= want) {
        rh = cand;
        ns = n;
        break;
      }
    }
  if (!rh) return EKAT_B200_ERR_UNSUPPORTED;
  if (tu.prefetch >= 2 && tu.prefetch <= ns) ns = tu.prefetch;  // tuning override: ring stages
  // one io warp per solver warp (TMA issue off the solver's critical path) while the block stays within 512 threads
  int io = (2 * rh + 3) * 32 <= 512 && tu.variant != kVarNoIoWarps ? 1 : 0;
  size_t smem = smem_for(rh, ns);
  auto ctas = [&](int cio, size_t sm) {
    return exact ? shared_ctas_per_sm<R, true, CM>(shared_threads(rh, cio), sm) : shared_ctas_per_sm<R, false, CM>(shared_threads(rh, cio), sm);
  };
  int per_sm = ctas(io, smem);
  // A deeper ring where it costs no residency: the factor warp's lead over the copies is what the ring buys, and at the
  // bench's shape 10 stages against 8 are worth 2 % (tools/sweep.py ring: 2 / 4 / 6 / 8 / 10 / 12 stages = 0.72 / 0.79 /
  // 0.86 / 0.897 / 0.915 / 0.903 of the HBM peak; 14 stages lose the second CTA of an SM: 0.60)
  if (tu.prefetch < 2 && ns == 8) {
    for (int n2 = 10; n2 > ns; --n2)
      if (smem_for(rh, n2) <= kMaxSmemPerCta && ctas(io, smem_for(rh, n2)) >= per_sm) {
        ns = n2;
        smem = smem_for(rh, ns);
        break;
      }
  }
  // Small batches (BASELINE config 1: 21,600 columns x 72 levels = 675 tiles for 148 SMs) are latency-bound: what a
  // launch costs is the number of WAVES of tiles, each a full forward + backward dependent chain, not bytes. There
  // the deep ring and the io warps are traded for residency: fewer ring stages (less shared memory) and no io warps
  // (fewer threads, so registers admit another CTA) until every tile is resident at once, if that is possible.
  const int64_t nwork = ((ncol + W - 1) / W) * (nrhs / rh);
  auto waves = [&](int n) { return (nwork + (int64_t)kNumSms * n - 1) / ((int64_t)kNumSms * n); };
  if (tu.prefetch < 2 && tu.variant == kVarAuto && nwork <= (int64_t)kNumSms * 16 && waves(per_sm) > 1) {
    int best_ns = ns, best_io = io, best_n = per_sm;
    for (int cio = io; cio >= 0; --cio)
      for (int cns = ns; cns >= 2; --cns) {
        const int n = ctas(cio, smem_for(rh, cns));
        if (waves(n) < waves(best_n)) {
          best_ns = cns;
          best_io = cio;
          best_n = n;
        }
      }
    ns = best_ns;
    io = best_io;
    per_sm = best_n;
    smem = smem_for(rh, ns);
  }
  const bool f64 = sizeof(R) == 8;
  const char* name = CM ? (exact ? "tridiag_tma_shared_cm<f64,bfb>" : "tridiag_tma_shared_cm<f64,thomas>")
                        : exact ? (f64 ? "tridiag_tma_shared<f64,bfb>" : "tridiag_tma_shared<f32,bfb>")
                                : (f64 ? "tridiag_tma_shared<f64,thomas>" : "tridiag_tma_shared<f32,thomas>");
  if (tu.variant == kVarTrace && !CM)  // instrumented instantiation (phase trace)
    return exact ? launch_shared_tma<R, true, true, false>(dl, d, du, X, Xin, ncol, nlev, nrhs, a_stride, x_stride, rh, ns, io, per_sm, smem, stream, name)
                 : launch_shared_tma<R, false, true, false>(dl, d, du, X, Xin, ncol, nlev, nrhs, a_stride, x_stride, rh, ns, io, per_sm, smem, stream, name);
  return exact ? launch_shared_tma<R, true, false, CM>(dl, d, du, X, Xin, ncol, nlev, nrhs, a_stride, x_stride, rh, ns, io, per_sm, smem, stream, name)
               : launch_shared_tma<R, false, false, CM>(dl, d, du, X, Xin, ncol, nlev, nrhs, a_stride, x_stride, rh, ns, io, per_sm, smem, stream, name);
}

// level-major operands: dl,d,du [nlev][a_stride], X [nrhs][nlev][x_stride]
template <typename R>
int tridiag_tma_shared(bool exact, R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nrhs, int64_t a_stride,
                       int64_t x_stride, cudaStream_t stream) {
  return run_shared<R, false>(exact, dl, d, du, X, X, ncol, nlev, nrhs, a_stride, x_stride, stream);
}
// same, out of place: right-hand sides read from B, solutions written to X (same shape and pitch; B is not modified)
template <typename R>
int tridiag_tma_shared_oop(bool exact, R* dl, R* d, R* du, const R* B, R* X, int64_t ncol, int nlev, int nrhs,
                           int64_t a_stride, int64_t x_stride, cudaStream_t stream) {
  return run_shared<R, false>(exact, dl, d, du, X, B, ncol, nlev, nrhs, a_stride, x_stride, stream);
}
// the reference's column-major operands, one right-hand side (f64): dl,d,du [ncol][a_stride], X [ncol][x_stride]
template <typename R>
int tridiag_tma_shared_cm(bool exact, R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int64_t a_stride, int64_t x_stride,
                          cudaStream_t stream) {
  return run_shared<R, true>(exact, dl, d, du, X, X, ncol, nlev, 1, a_stride, x_stride, stream);
}

int debug_trace(long long* out, int n) {
  const int total = 2 * kTraceItems * kTraceWords;
  if (n > total) n = total;
  return check_cuda(cudaMemcpyFromSymbol(out, g_trace, (size_t)n * sizeof(long long)));
}

template int tridiag_tma_shared<double>(bool, double*, double*, double*, double*, int64_t, int, int, int64_t, int64_t, cudaStream_t);
template int tridiag_tma_shared<float>(bool, float*, float*, float*, float*, int64_t, int, int, int64_t, int64_t, cudaStream_t);
template int tridiag_tma_shared_oop<double>(bool, double*, double*, double*, const double*, double*, int64_t, int, int, int64_t, int64_t,
                                            cudaStream_t);
template int tridiag_tma_shared_oop<float>(bool, float*, float*, float*, const float*, float*, int64_t, int, int, int64_t, int64_t,
                                           cudaStream_t);
template int tridiag_tma_shared_cm<double>(bool, double*, double*, double*, double*, int64_t, int, int64_t, int64_t, cudaStream_t);
template int tridiag_tma_shared_cm<float>(bool, float*, float*, float*, float*, int64_t, int, int64_t, int64_t, cudaStream_t);

}  // namespace ekb
