// tridiag_bfb_shared.cu -- ekat::tridiag::bfb with ONE matrix per column shared by several right-hand sides
// (BASELINE config 4's shape, bit-exact flavour), level-major operands, TMA-fed, persistent CTAs whose forward
// sweeps and back-substitutions run on DIFFERENT warps.
//
// Replaces the per-column loops around ekat::tridiag::bfb(team, dl, d, du, X[nrow][nrhs]) with rank-1 A (reference
// src/algorithm/ekat_tridiag.hpp:292-320, :632-648). Arithmetic: the reference's operation order with separately
// rounded * - / -- bit-identical.
//
// Why this shape. k_tridiag_tma_shared<bfb> gave every right-hand side ONE warp that swept forward behind the factor
// warp and then back. A clock64 trace of it (tools/trace_shared.py, profiles/r02_cr_blocked_negative.txt) showed a
// work item costing 36.7 K cycles = the solver warps' forward sweep (13.7 K, mostly waiting for the factor warp's
// division chain: 147 cycles per level) + their way back (22.3 K: a division per level, and du / reciprocals coming
// through the same 8-stage ring the forward operands use). The two halves cannot overlap on one warp, and with one
// ring the factor warp cannot lead by more than the ring's depth while the solver warps are on their way back. Here:
//   * rh FORWARD warps and rh BACKWARD warps: while the backward warps take item i up the columns, the factor warp
//     and the forward warps take item i+1 down, chunk k of item i+1 landing in the plane slot that the way back of
//     item i has just stored and released (the slot order is mirrored from one item to the next as in
//     tridiag_tma_shared.cu, so this is exactly the order both sweeps want);
//   * the way back has its own small ring for du (an L2 hit: the tile's du went down the forward ring a sweep ago)
//     and refines the reciprocals of the pivots itself, off the chain, so it never waits for the factor warp;
//   * forward -> backward hand-over of a right-hand side: one mbarrier per right-hand side (arrive = release of the
//     forward warp's shared-memory writes and, through its own waits on the factor warp, of the pivots);
//   * three producer threads (forward ring, du ring, pivot plane), each blocking only on its own consumers; the
//     backward warps issue their own tensor stores and refill the slot with the next item's right-hand side.
// Per work item the CTA is now bound by max(factor chain, way back) instead of their sum.
// HBM traffic is the algorithmic minimum; groups of a tile on neighbouring CTAs share dl, d, du through L2.
#include <type_traits>

#include "tridiag_tma.cuh"

namespace ekb {

// Phase trace for tuning, compiled in only with -DEKB_TRACE_BFB (tools/trace_bfb_shared.py builds that library for one run):
// per traced CTA (blockIdx 0 and gridDim/2) and work item < 64, 16 counters in clock64 cycles, written by lane 0 of the
// factor warp, forward warp 0, backward warp 0 and the three producers.
#ifdef EKB_TRACE_BFB
__device__ long long g_bfb_trace[2 * 64 * 16];
__device__ __forceinline__ void bfb_trace_put(int it, int idx, long long v) {
  const int cs = blockIdx.x == 0 ? 0 : (blockIdx.x == gridDim.x / 2 ? 1 : -1);
  if (cs >= 0 && it < 64 && (threadIdx.x & 31) == 0) g_bfb_trace[(cs * 64 + it) * 16 + idx] = v;
}
#define TR_DECL(name) long long name = 0
#define TR_WAIT(bar, par, acc)      \
  {                                 \
    const long long t_ = clock64(); \
    mbar_wait(bar, par);            \
    acc += clock64() - t_;          \
  }
#define TR_PUT(it, idx, v) bfb_trace_put(it, idx, v)
#define TR_NOW() clock64()
#else
#define TR_DECL(name)
#define TR_WAIT(bar, par, acc) mbar_wait(bar, par)
#define TR_PUT(it, idx, v)
#define TR_NOW() 0
#endif

// RAGGED: the level count is not a multiple of the chunk (the top chunk then leaves by plain stores); the instantiation
// without it carries none of that in the way-back loop, the warp whose instruction stream bounds the kernel
template <typename R, int CPT, bool RAGGED>
__global__ void __launch_bounds__(512, 1)
k_tridiag_bfb_shared(const __grid_constant__ CUtensorMap mdl, const __grid_constant__ CUtensorMap md,
                     const __grid_constant__ CUtensorMap mdu, const __grid_constant__ CUtensorMap mx,
                     const __grid_constant__ CUtensorMap mxin, R* __restrict__ X, int64_t ncol, int nlev, int64_t x_stride, int rh,
                     int ngroups, int nwork, int NS, int NU) {
  using V = Vec<R, CPT>;
  constexpr int LB = kLB, W = 32 * CPT, CH = LB * 32;
  constexpr unsigned CHUNK_BYTES = LB * W * sizeof(R);
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nfull = nlev / LB, ntail = RAGGED ? nlev - nfull * LB : 0, nch = nfull + (RAGGED ? 1 : 0), nslot = nch;

  uint64_t* raw = reinterpret_cast<uint64_t*>(smem_raw);  // [NS]  dl | du of a chunk landed in a forward-ring stage
  uint64_t* ready = raw + NS;                             // [NS]  the factor warp turned dl into multipliers there
  uint64_t* freeb = ready + NS;                           // [NS]  every forward warp has read the stage
  uint64_t* uraw = freeb + NS;                            // [NU]  du of a chunk landed in a stage of the way-back ring
  uint64_t* ufree = uraw + NU;                            // [NU]  every backward warp has read it
  uint64_t* plc = ufree + NU;                             // [nslot] d of a chunk landed in its pivot-plane slot
  uint64_t* cfree = plc + nslot;                          // [nslot] every backward warp has read the pivots of the slot
  uint64_t* fdone = cfree + nslot;                        // [rh]  forward sweep of a right-hand side complete
  uint64_t* plx = fdone + rh;                             // [rh][nslot] right-hand-side chunk landed in its slot
  const size_t off = ((size_t)(3 * NS + 2 * NU + 2 * nslot + rh + rh * nslot) * 8 + 127) & ~(size_t)127;
  V* PD = reinterpret_cast<V*>(smem_raw + off);  // d -> pivots d'
  V* PX = PD + (size_t)nslot * CH;               // [rh] B -> x' -> x
  V* RING = PX + (size_t)rh * nslot * CH;        // [NS][dl -> m | du]
  V* URING = RING + (size_t)NS * 2 * CH;         // [NU][du]

  if (threadIdx.x == 0) {
    for (int s = 0; s < NS; ++s) {
      mbar_init(&raw[s], 1);
      mbar_init(&ready[s], 1);
      mbar_init(&freeb[s], rh);
    }
    for (int s = 0; s < NU; ++s) {
      mbar_init(&uraw[s], 1);
      mbar_init(&ufree[s], rh);
    }
    for (int k = 0; k < nslot; ++k) {
      mbar_init(&plc[k], 1);
      mbar_init(&cfree[k], rh);
    }
    for (int k = 0; k < rh + rh * nslot; ++k) mbar_init(&fdone[k], 1);  // fdone and plx
    fence_mbar_init();
  }
  __syncthreads();

  const int nit = (nwork - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;  // work items of this CTA
  auto slot = [&](int k, int it) { return (it & 1) ? nch - 1 - k : k; };  // every slot is filled once per item: parity it & 1
  auto item_c0 = [&](int it) { return (int)(((unsigned)blockIdx.x + (unsigned)it * gridDim.x) / (unsigned)ngroups) * W; };
  auto item_j0 = [&](int it) { return (int)(((unsigned)blockIdx.x + (unsigned)it * gridDim.x) % (unsigned)ngroups) * rh; };
  // ring position of a role: stage and phase parity of the use it is at
  int rs = 0;
  unsigned rph = 0;
  auto ring_advance = [&](int n) {
    if (++rs == n) {
      rs = 0;
      rph ^= 1u;
    }
  };
  const uint64_t pol_stream = l2_policy_evict_first();
  // a slot is refilled kSD steps after its tensor store was committed (wait_group.read<kSD>); 1 / 2 / 3: 4.15 / 4.09 / 4.08 TB/s
  constexpr int kSD = 1;
  constexpr int jb0 = RAGGED ? 1 : 0;  // the partial top chunk (way-back step 0) leaves by plain stores

  if (warp < rh) {
    // ================================================================ forward warp of right-hand side `warp`
    const int jw = warp;
    V* PXj = PX + (size_t)jw * nslot * CH;
    uint64_t* plxj = plx + (size_t)jw * nslot;
    for (int it = 0; it < nit; ++it) {
      R yprev[CPT];
#pragma unroll
      for (int j = 0; j < CPT; ++j) yprev[j] = (R)0;
      TR_DECL(w_plx);
      TR_DECL(w_ready);
      if (jw == 0) TR_PUT(it, 4, TR_NOW());
      for (int k = 0; k < nch; ++k) {
        const int sl = slot(k, it), s = rs;
        TR_WAIT(&plxj[sl], (unsigned)(it & 1), w_plx);
        TR_WAIT(&ready[s], rph, w_ready);
        ring_advance(NS);
        const V* rm = RING + (size_t)s * 2 * CH;
        V* px = PXj + (size_t)sl * CH;
        V m[LB], x[LB];
#pragma unroll
        for (int u = 0; u < LB; ++u) {
          m[u] = rm[u * 32 + lane];
          x[u] = px[u * 32 + lane];
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&freeb[s]);
#pragma unroll
        for (int u = 0; u < LB; ++u) {
#pragma unroll
          for (int j = 0; j < CPT; ++j) {
            yprev[j] = x_sub(x[u].v[j], x_mul(m[u].v[j], yprev[j]));  // X(i) -= dl(i)*X(i-1)   :315-316
            x[u].v[j] = yprev[j];
          }
        }
#pragma unroll
        for (int u = 0; u < LB; ++u) px[u * 32 + lane] = x[u];
      }
      __syncwarp();
      if (jw == 0) {
        TR_PUT(it, 5, TR_NOW());
        TR_PUT(it, 6, w_plx);
        TR_PUT(it, 7, w_ready);
      }
      if (lane == 0) mbar_arrive(&fdone[jw]);  // x' of this right-hand side (and, through `ready`, the pivots) are in place
    }
  } else if (warp < 2 * rh) {
    // ================================================================ backward warp of right-hand side `warp - rh`
    // (it issues its own tensor stores and refills: separate io warps as in tridiag_tma_shared.cu were measured --
    // 19 warps leave 96 registers per thread, and the spills cost more than the ~130 cycles per issue they save:
    // 3.81 against 4.06 TB/s at config 4)
    const int jw = warp - rh;
    V* PXj = PX + (size_t)jw * nslot * CH;
    uint64_t* plxj = plx + (size_t)jw * nslot;
    auto load_x = [&](int it2, int k, int c0_2, int xrow_2) {  // lane 0
      const int sl = slot(k, it2);
      mbar_expect_tx(&plxj[sl], CHUNK_BYTES);
      tma_load_2d_hint(PXj + (size_t)sl * CH, &mxin, c0_2, xrow_2 + k * LB, &plxj[sl], pol_stream);  // B (mxin) in, X (mx) out
    };
    if (lane == 0)
      for (int k = 0; k < nch; ++k) load_x(0, k, item_c0(0), (item_j0(0) + jw) * nlev);
    for (int it = 0; it < nit; ++it) {
      const int c0 = item_c0(it), xrow0 = (item_j0(it) + jw) * nlev;
      const bool has1 = it + 1 < nit;
      const int c0_n1 = item_c0(it + 1), xrow_n1 = (item_j0(it + 1) + jw) * nlev;
      auto refill = [&](int jb) {  // lane 0: the slot of this item's chunk nch-1-jb is free again: it is the next item's chunk jb
        if (has1) load_x(it + 1, jb, c0_n1, xrow_n1);
      };
      TR_DECL(w_fdone);
      TR_DECL(w_uraw);
      TR_DECL(w_store);
      TR_DECL(w_sec_a);
      TR_DECL(w_sec_b);
      TR_DECL(w_sec_c);
      TR_WAIT(&fdone[jw], (unsigned)(it & 1), w_fdone);
      if (jw == 0) TR_PUT(it, 8, TR_NOW());
      R xv[CPT];
#pragma unroll
      for (int j = 0; j < CPT; ++j) xv[j] = (R)0;
      for (int jb = 0; jb < nch; ++jb) {
        const int k = nch - 1 - jb, sl = slot(k, it), s = rs;
        TR_WAIT(&uraw[s], rph, w_uraw);
#ifdef EKB_TRACE_BFB
        const long long ta_ = clock64();
#endif
        ring_advance(NU);
        const V* bu = URING + (size_t)s * CH;
        const V* pc = PD + (size_t)sl * CH;
        V* px = PXj + (size_t)sl * CH;
        V y[LB], c[LB], ui[LB];
#pragma unroll
        for (int u = 0; u < LB; ++u) {
          y[u] = px[u * 32 + lane];
          c[u] = pc[u * 32 + lane];
          ui[u] = bu[u * 32 + lane];
        }
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(&cfree[sl]);
          mbar_arrive(&ufree[s]);
        }
#ifdef EKB_TRACE_BFB
        const long long tb_ = clock64();
        w_sec_a += tb_ - ta_;
#endif
        if (RAGGED && k == nfull) {
          // partial top chunk (ntail levels): a tensor store would spill into the next right-hand side -> plain stores
          const int64_t col = (int64_t)c0 + lane * CPT;
#pragma unroll
          for (int u = LB - 1; u >= 0; --u) {
            if (u < ntail) {
              V o;
#pragma unroll
              for (int j = 0; j < CPT; ++j) {
                const R uu = (u == ntail - 1) ? (R)0 : ui[u].v[j];  // du(nrow-1) is never used (:24-27)
                xv[j] = x_div(x_sub(y[u].v[j], x_mul(uu, xv[j])), c[u].v[j]);  // :317-319
                o.v[j] = xv[j];
              }
              R* g = X + (size_t)(xrow0 + k * LB + u) * x_stride + col;
#pragma unroll
              for (int j = 0; j < CPT; ++j)
                if (col + j < ncol) __stcs(g + j, o.v[j]);
            }
          }
          if (lane == 0) refill(jb);  // every lane read its part of the slot before the __syncwarp above
        } else {
          if (jb == 0) {  // this chunk ends with the last level
#pragma unroll
            for (int j = 0; j < CPT; ++j) ui[LB - 1].v[j] = (R)0;
          }
          V rr[LB];  // the part of the quotients by the pivots that does not depend on the right-hand side
#pragma unroll
          for (int u = 0; u < LB; ++u) {
#pragma unroll
            for (int j = 0; j < CPT; ++j) rr[u].v[j] = rcp_refined(c[u].v[j]);
          }
          R sxv[CPT];
          V ys[LB];
#pragma unroll
          for (int j = 0; j < CPT; ++j) sxv[j] = xv[j];
#pragma unroll
          for (int u = 0; u < LB; ++u) ys[u] = y[u];
          unsigned badu = 0;  // the pivots are validated once per chunk, off the chain; a level checks numerator and quotient only
#pragma unroll
          for (int u = 0; u < LB; ++u) {
#pragma unroll
            for (int j = 0; j < CPT; ++j) badu |= divisor_in_range(c[u].v[j]) ? 0u : 1u;
          }
#pragma unroll
          for (int u = LB - 1; u >= 0; --u) {
#pragma unroll
            for (int j = 0; j < CPT; ++j) {
              const R num = x_sub(y[u].v[j], x_mul(ui[u].v[j], xv[j]));  // X(i-1) - du(i-1)*X(i)   :318-319
              xv[j] = div_spec_b(num, c[u].v[j], rr[u].v[j], badu);
              y[u].v[j] = xv[j];
            }
          }
          bool bad = badu != 0;
          if (__any_sync(0xffffffffu, bad)) {  // rare operands somewhere in the warp: redo with the full-range division
#pragma unroll
            for (int j = 0; j < CPT; ++j) xv[j] = sxv[j];
#pragma unroll
            for (int u = 0; u < LB; ++u) y[u] = ys[u];
            bfb_bwd_chunk<R, CPT, LB, true>(y, c, ui, xv, bad);
          }
#ifdef EKB_TRACE_BFB
          const long long tc_ = clock64();
          w_sec_b += tc_ - tb_;
#endif
#pragma unroll
          for (int u = 0; u < LB; ++u) px[u * 32 + lane] = y[u];
          fence_proxy_async();  // generic-proxy writes of this chunk -> visible to the TMA store
          __syncwarp();
#ifdef EKB_TRACE_BFB
          w_sec_a += clock64() - tc_;  // (trace: STS + fence + syncwarp are booked on the load section's counter)
#endif
          if (lane == 0) {
            tma_store_2d(&mx, c0, xrow0 + k * LB, PXj + (size_t)sl * CH);
            tma_store_commit();
            if (jb - jb0 >= kSD) {
#ifdef EKB_TRACE_BFB
              const long long t_ = clock64();
#endif
              tma_store_wait_read<kSD>();  // the store committed kSD steps ago has read its slot
#ifdef EKB_TRACE_BFB
              w_store += clock64() - t_;
#endif
              refill(jb - kSD);
            }
          }
#ifdef EKB_TRACE_BFB
          __syncwarp();
          w_sec_c += clock64() - tc_;
#endif
        }
      }
      if (lane == 0) {
        tma_store_wait_read<0>();
        for (int jb = tmax(jb0, nch - kSD); jb < nch; ++jb) refill(jb);
      }
      __syncwarp();
      if (jw == 0) {
        TR_PUT(it, 9, TR_NOW());
        TR_PUT(it, 10, w_fdone);
        TR_PUT(it, 11, w_uraw);
        TR_PUT(it, 12, w_store);
        TR_PUT(it, 13, w_sec_a);
        TR_PUT(it, 14, w_sec_b);
        TR_PUT(it, 15, w_sec_c);
      }
    }
  } else if (warp == 2 * rh) {
    // ================================================================ factor warp (:292-303)
    for (int it = 0; it < nit; ++it) {
      R dprev[CPT], uprev[CPT];
#pragma unroll
      for (int j = 0; j < CPT; ++j) {
        dprev[j] = (R)1;
        uprev[j] = (R)0;
      }
      TR_DECL(w_raw);
      TR_DECL(w_plc);
      TR_PUT(it, 0, TR_NOW());
      for (int k = 0; k < nch; ++k) {
        const int sl = slot(k, it), s = rs;
        TR_WAIT(&raw[s], rph, w_raw);
        ring_advance(NS);
        TR_WAIT(&plc[sl], (unsigned)(it & 1), w_plc);
        V* r0 = RING + (size_t)s * 2 * CH;
        V* pc = PD + (size_t)sl * CH;
        V a[LB], b[LB], c[LB], om[LB], od[LB];
#pragma unroll
        for (int u = 0; u < LB; ++u) {
          a[u] = r0[u * 32 + lane];
          b[u] = pc[u * 32 + lane];
          c[u] = r0[CH + u * 32 + lane];
        }
        R sd[CPT], su[CPT];
#pragma unroll
        for (int j = 0; j < CPT; ++j) {
          sd[j] = dprev[j];
          su[j] = uprev[j];
        }
        bool bad = false;
        auto sweep = [&](auto slow) {
          R rp[CPT];  // ~ 1 / d'(i-1): reciprocal of the PREDICTED pivot, off the chain (common.cuh bfb_pivot_step)
          if constexpr (!decltype(slow)::value) {
#pragma unroll
            for (int j = 0; j < CPT; ++j) rp[j] = rcp_seed(dprev[j]);
          }
#pragma unroll
          for (int u = 0; u < LB; ++u) {
#pragma unroll
            for (int j = 0; j < CPT; ++j) {
              const R av = (u == 0 && k == 0) ? (R)0 : a[u].v[j];  // dl(0) is never used (:24-27)
              R m, dc;
              if constexpr (decltype(slow)::value) {
                m = (u == 0 && k == 0) ? (R)0 : x_div(av, dprev[j]);
                dc = x_sub(b[u].v[j], x_mul(m, uprev[j]));         // d(i) -= dl(i)*du(i-1)   :301
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
        };
        sweep(std::false_type{});
        if (__any_sync(0xffffffffu, bad)) {  // rare operands somewhere in the warp: redo with the full-range division
#pragma unroll
          for (int j = 0; j < CPT; ++j) {
            dprev[j] = sd[j];
            uprev[j] = su[j];
          }
          sweep(std::true_type{});
        }
#pragma unroll
        for (int u = 0; u < LB; ++u) {
          r0[u * 32 + lane] = om[u];
          pc[u * 32 + lane] = od[u];
        }
        fence_proxy_async();  // these slots are later overwritten by TMA copies
        __syncwarp();
        if (lane == 0) mbar_arrive(&ready[s]);
      }
      TR_PUT(it, 1, TR_NOW());
      TR_PUT(it, 2, w_raw);
      TR_PUT(it, 3, w_plc);
    }
  } else if (warp == 2 * rh + 1) {
    // ================================================================ forward-ring producer
    if (lane == 0) {
      bool wrapped = false;
      for (int it = 0; it < nit; ++it) {
        const int c0 = item_c0(it);
        TR_DECL(w_freeb);
        for (int k = 0; k < nch; ++k) {
          const int s = rs;
          if (wrapped) TR_WAIT(&freeb[s], rph ^ 1u, w_freeb);  // the use one lap earlier has been read by every forward warp
          ring_advance(NS);
          if (rs == 0) wrapped = true;
          V* st = RING + (size_t)s * 2 * CH;
          mbar_expect_tx(&raw[s], 2u * CHUNK_BYTES);
          tma_load_2d(st, &mdl, c0, k * LB, &raw[s]);
          tma_load_2d(st + CH, &mdu, c0, k * LB, &raw[s]);
        }
      }
    }
  } else if (warp == 2 * rh + 2) {
    // ================================================================ way-back ring producer: du, top chunk first
    if (lane == 0) {
      bool wrapped = false;
      for (int it = 0; it < nit; ++it) {
        const int c0 = item_c0(it);
        for (int jb = 0; jb < nch; ++jb) {
          const int s = rs;
          if (wrapped) mbar_wait(&ufree[s], rph ^ 1u);
          ring_advance(NU);
          if (rs == 0) wrapped = true;
          mbar_expect_tx(&uraw[s], CHUNK_BYTES);
          tma_load_2d(URING + (size_t)s * CH, &mdu, c0, (nch - 1 - jb) * LB, &uraw[s]);
        }
      }
    }
  } else if (warp == 2 * rh + 3) {
    // ================================================================ pivot-plane producer
    if (lane == 0) {
      auto load_c = [&](int it2, int k) {
        const int sl = slot(k, it2);
        mbar_expect_tx(&plc[sl], CHUNK_BYTES);
        tma_load_2d(PD + (size_t)sl * CH, &md, item_c0(it2), k * LB, &plc[sl]);
      };
      for (int k = 0; k < nch; ++k) load_c(0, k);
      for (int it = 0; it + 1 < nit; ++it) {
        TR_DECL(w_cfree);
        for (int jb = 0; jb < nch; ++jb) {  // the slot of this item's chunk nch-1-jb becomes the next item's chunk jb
          TR_WAIT(&cfree[slot(nch - 1 - jb, it)], (unsigned)(it & 1), w_cfree);
          load_c(it + 1, jb);
        }
      }
    }
  }
}

// Returns EKAT_B200_ERR_UNSUPPORTED when the operands do not meet TMA's requirements or a tile does not fit on
// chip; the caller then uses k_tridiag_tma_shared<bfb> / the cp.async kernels.
template <typename R>
int tridiag_bfb_shared(R* dl, R* d, R* du, R* X, const R* Xin, int64_t ncol, int nlev, int nrhs, int64_t a_stride, int64_t x_stride,
                       cudaStream_t stream) {
  auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  if (!al(dl) || !al(d) || !al(du) || !al(X) || !al(Xin) || (a_stride * sizeof(R)) % 16 || (x_stride * sizeof(R)) % 16)
    return EKAT_B200_ERR_UNSUPPORTED;
  if (ncol >= (int64_t)1 << 31 || (int64_t)nrhs * nlev >= (int64_t)1 << 31) return EKAT_B200_ERR_UNSUPPORTED;
  constexpr int CPT = TmaCfg<R>::CPT, W = 32 * CPT;
  const Tuning tu = tuning();
  if (tu.variant != kVarAuto) return EKAT_B200_ERR_UNSUPPORTED;
  const int nch = (nlev + kLB - 1) / kLB, nslot = nch;
  const size_t chunk = (size_t)kLB * W * sizeof(R);
  auto smem_for = [&](int rh, int ns, int nu) {
    const size_t bars = ((size_t)(3 * ns + 2 * nu + 2 * nslot + rh + rh * nslot) * 8 + 127) & ~(size_t)127;
    return bars + (size_t)(1 + rh) * nslot * chunk + (size_t)ns * 2 * chunk + (size_t)nu * chunk;
  };
  // right-hand sides per CTA: the largest divisor of nrhs (at most 6: 2*rh + 4 warps) whose planes fit next to a forward
  // ring of >= 4 stages and a way-back ring of >= 3
  int rh = 0, ns = 0, nu = 0;
  for (int cand = tmin(nrhs, 6); cand >= 1 && !rh; --cand) {
    if (nrhs % cand) continue;
    if (tu.cols_per_cta > 0 && cand != tu.cols_per_cta) continue;  // tuning override: right-hand sides per CTA
    for (int n = 8; n >= 4 && !rh; --n)
      for (int u = 4; u >= 3 && !rh; --u)
        if (smem_for(cand, n, u) <= kMaxSmemPerCta) {
          rh = cand;
          ns = n;
          nu = u;
        }
  }
  if (!rh) return EKAT_B200_ERR_UNSUPPORTED;
  // Short columns: k_tridiag_tma_shared<bfb> holds all right-hand sides of a tile in one CTA (one factorisation per tile,
  // not one per group) where this kernel's 2*rh + 4 warps cannot -- measured at 72 levels x 10 rhs: 5.0 against 3.7 TB/s
  if (rh < nrhs && nrhs <= 13 && (size_t)(1 + nrhs) * nslot * chunk + 4 * 2 * chunk + 2048 <= kMaxSmemPerCta && tu.cols_per_cta <= 0)
    return EKAT_B200_ERR_UNSUPPORTED;
  if (tu.prefetch >= 2 && tu.prefetch <= ns) ns = tu.prefetch;
  const size_t smem = smem_for(rh, ns, nu);
  CUtensorMap mdl, md, mdu, mx, mxin;
  if (!make_tensor_map_2d<R>(&mdl, dl, ncol, nlev, a_stride, W, kLB, 0) || !make_tensor_map_2d<R>(&md, d, ncol, nlev, a_stride, W, kLB, 0) ||
      !make_tensor_map_2d<R>(&mdu, du, ncol, nlev, a_stride, W, kLB, 0) ||
      !make_tensor_map_2d<R>(&mx, X, ncol, (int64_t)nrhs * nlev, x_stride, W, kLB, 0) ||
      !make_tensor_map_2d<R>(&mxin, Xin, ncol, (int64_t)nrhs * nlev, x_stride, W, kLB, 0))
    return EKAT_B200_ERR_UNSUPPORTED;
  const int ngroups = nrhs / rh;
  const int64_t ntiles = (ncol + W - 1) / W, nwork = ntiles * ngroups;
  if (nwork >= (int64_t)1 << 30) return EKAT_B200_ERR_UNSUPPORTED;
  const int64_t grid = tmin<int64_t>(nwork, (int64_t)kNumSms);
  auto k = (nlev % kLB) ? k_tridiag_bfb_shared<R, CPT, true> : k_tridiag_bfb_shared<R, CPT, false>;
  const int rc = allow_smem(k, smem);
  if (rc) return rc;
  k<<<(unsigned)grid, 32 * (2 * rh + 4), smem, stream>>>(mdl, md, mdu, mx, mxin, X, ncol, nlev, x_stride, rh, ngroups, (int)nwork, ns, nu);
  note_launch(sizeof(R) == 8 ? "tridiag_bfb_shared<f64>" : "tridiag_bfb_shared<f32>");
  return check_cuda(cudaGetLastError());
}

#ifdef EKB_TRACE_BFB
extern "C" int ekat_b200_debug_trace_bfb(long long* out, int n) {
  if (n > 2 * 64 * 16) n = 2 * 64 * 16;
  return check_cuda(cudaMemcpyFromSymbol(out, g_bfb_trace, (size_t)n * sizeof(long long)));
}
#endif

template int tridiag_bfb_shared<double>(double*, double*, double*, double*, const double*, int64_t, int, int, int64_t, int64_t, cudaStream_t);
template int tridiag_bfb_shared<float>(float*, float*, float*, float*, const float*, int64_t, int, int, int64_t, int64_t, cudaStream_t);

}  // namespace ekb
