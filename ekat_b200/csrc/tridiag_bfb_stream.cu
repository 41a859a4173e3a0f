// tridiag_bfb_stream.cu -- ekat::tridiag::bfb, one matrix and one right-hand side per column, TMA-fed, with the
// copy ring decoupled from the results.
//
// Replaces the per-column loop around ekat::tridiag::bfb(team, dl, d, du, X) (reference
// src/algorithm/ekat_tridiag.hpp:292-320 bfb_thomas_factorize / bfb_thomas_solve, :632-678) for BASELINE configs
// 1, 3, 5. Arithmetic: the reference's operation order with separately rounded * - / (tridiag_tma.cuh
// bfb_fwd_chunk / bfb_bwd_chunk) -- bit-identical to the reference.
//
// Why this shape. bfb cannot leave its division chain (~110 cycles per level forward, ~45 backward), so an SM is
// only as fast as the number of columns it holds, and that is bounded by shared memory: the pivots d' of a tile
// (32 columns x nlev) must stay on chip between the sweeps. Everything else is therefore kept as small AND as busy
// as possible:
//   * x' goes back to X in place through L2 (evict-last) and is read again on the way back, as in
//     k_tridiag_bfb_spill (tridiag_tma.cu), so only one plane is resident;
//   * ring stages [dl | du | x] are INPUT ONLY: the sweeping warp moves a chunk's operands into registers and
//     releases the stage at once (ebar), long before the chunk's chain has run; results leave through two small
//     output buffers (OB, one per chunk parity). In k_tridiag_bfb_spill a stage was tied up from the copy's issue
//     until the tensor store of its results had read it -- ncu showed the sweeping warps waiting for copies;
//   * one thread of an io warp issues every copy (a tensor copy costs its issuing thread ~130 cycles);
//   * two sweeping warps take the chunks of a tile in turn (PP): while one runs the chain of chunk k the other
//     stores the results of chunk k-1 and stages the operands of chunk k+1 in registers; the chain state crosses
//     in shared memory under a pair of named barriers;
//   * at the turn-around the x' of the last two chunks is still in the output buffers, so the back-substitution
//     starts without waiting for anything to come back from L2.
// Steps t = 0 .. 2*nch-1: forward chunk t, then backward chunk 2*nch-1-t. Step t has ring use t in stage t % NS.
#include "tridiag_tma.cuh"

namespace ekb {

template <typename R, int CPT, int NS, int LBT, bool CM, bool PP, bool CLB>
__global__ void __launch_bounds__(PP ? 96 : 64)
k_tridiag_bfb_stream(const __grid_constant__ CUtensorMap mdl, const __grid_constant__ CUtensorMap md,
                     const __grid_constant__ CUtensorMap mdu, const __grid_constant__ CUtensorMap mx, R* __restrict__ X,
                     int64_t ncol, int nlev, int64_t x_stride, int pd, int clob_arg, R* __restrict__ DLg, R* __restrict__ Dg,
                     int64_t a_stride) {
  const int clob = CLB ? clob_arg : 0;  // (the instantiation without side effects carries no multiplier registers)
  using V = Vec<R, CPT>;
  constexpr int LB = LBT, W = 32 * CPT, CH = LB * 32;
  constexpr unsigned CHUNK_BYTES = LB * W * sizeof(R);
  constexpr int NCW = PP ? 2 : 1;  // sweeping warps
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, cw = threadIdx.x >> 5;
  const int nfull = nlev / LB, ntail = nlev - nfull * LB, nch = nfull + (ntail ? 1 : 0);
  const int NU = 2 * nch;  // steps = ring uses

  uint64_t* rbar = reinterpret_cast<uint64_t*>(smem_raw);  // [NS] copy landed in the stage
  uint64_t* ebar = rbar + NS;                              // [NS] the stage's operands are in registers
  uint64_t* dbar = ebar + NS;                              // [2]  results of a step are in OB[b]
  uint64_t* obar = dbar + 2;                               // [2]  OB[b] may be overwritten
  uint64_t* pbar = obar + 2;                               // [nch] pivot-plane chunk landed
  constexpr size_t AL = CM ? 1024 : 128;  // swizzled chunks sit on 1 KB boundaries
  const size_t hs_off = ((size_t)(2 * NS + 4 + nch) * 8 + 15) & ~(size_t)15;
  V* HS = reinterpret_cast<V*>(smem_raw + hs_off);  // chain state handed from one sweeping warp to the other [3][32]
  const size_t off = (hs_off + 3 * 32 * sizeof(V) + AL - 1) & ~(AL - 1);
  V* P0 = reinterpret_cast<V*>(smem_raw + off);  // d -> d'
  V* RG = P0 + (size_t)nch * CH;                 // [NS][3][LB][32]
  V* OB = RG + (size_t)NS * 3 * CH;              // [2][LB][32]
  V* OBM = OB + 2 * CH;                          // [2][LB][32] multipliers on their way to dl (clob bit 0 only)

  const int c0 = blockIdx.x * W;
  const int row0 = blockIdx.y * nlev;
  if (threadIdx.x == 0) {
    for (int s = 0; s < 2 * NS + 4 + nch; ++s) mbar_init(&rbar[s], 1);
    fence_mbar_init();
  }
  __syncthreads();

  if (cw == NCW) {
    // ================================================================ io warp: one thread issues every copy
    if (lane != 0) return;
    const uint64_t pol_stream = l2_policy_evict_first(), pol_keep = l2_policy_evict_last();
    auto load_chunk = [&](void* dst, const CUtensorMap* m, int k, uint64_t* bar, uint64_t pol) {
      if (CM) tma_load_2d_hint(dst, m, k * LB, c0, bar, pol);
      else tma_load_2d_hint(dst, m, c0, row0 + k * LB, bar, pol);
    };
    auto store_chunk = [&](const CUtensorMap* m, int k, const void* src, uint64_t pol) {  // (the caller commits)
      if (CM) tma_store_2d_hint(m, k * LB, c0, src, pol);
      else tma_store_2d_hint(m, c0, row0 + k * LB, src, pol);
    };
    auto issue_use = [&](int u) {
      const int s = u % NS;
      V* st = RG + (size_t)s * 3 * CH;
      if (u < nch) {
        mbar_expect_tx(&rbar[s], 3u * CHUNK_BYTES);
        load_chunk(st, &mdl, u, &rbar[s], pol_stream);
        load_chunk(st + CH, &mdu, u, &rbar[s], pol_keep);  // read again by the back-substitution
        load_chunk(st + 2 * CH, &mx, u, &rbar[s], pol_stream);
      } else {
        const int jb = u - nch, k = nch - 1 - jb;
        mbar_expect_tx(&rbar[s], (jb >= 2 ? 2u : 1u) * CHUNK_BYTES);
        load_chunk(st + CH, &mdu, k, &rbar[s], pol_stream);
        if (jb >= 2) load_chunk(st + 2 * CH, &mx, k, &rbar[s], pol_stream);  // x', written by this CTA on the way up
      }
    };
    auto issue_plane = [&](int k) {
      mbar_expect_tx(&pbar[k], CHUNK_BYTES);
      load_chunk(P0 + (size_t)k * CH, &md, k, &pbar[k], pol_stream);
    };
    unsigned eph = 0, dph = 0;  // phase bit per barrier
    int next_u = 0, last_e = -1, fwd_stores = 0, pending = -1;
    bool flushed = nch <= 2;  // forward stores (chunks 0 .. nch-3) all complete
    // ring use u may go out once its stage has been released (use u-NS) and, if it reads x' back (backward step
    // jb >= 2, chunk <= nch-3), once the forward stores have completed (one blocking wait, after the last of them)
    auto pump = [&]() {
      while (next_u < NU && next_u - NS <= last_e) {
        if (next_u >= nch + 2 && !flushed) {
          if (fwd_stores < nch - 2) break;
          tma_store_wait_all<0>();
          flushed = true;
        }
        issue_use(next_u++);
      }
    };
    // results of step t: store them (if they leave by tensor store) and tell the sweeping warps when OB[b] is free
    auto handle_d = [&](int t) {
      const int k = t < nch ? t : 2 * nch - 1 - t, b = k & 1;
      mbar_wait(&dbar[b], (dph >> b) & 1u);
      dph ^= 1u << b;
      const bool st = t < nch ? (k <= nch - 3) : (CM || k != nfull);
      bool any = st;
      if (st) {
        store_chunk(&mx, k, OB + (size_t)b * CH, t < nch ? pol_keep : pol_stream);
        if (t < nch) ++fwd_stores;
      }
      if (clob && t < nch && k != nfull) {  // the factorisation goes back too (a ragged top chunk is stored by the sweeping warp)
        if (clob & 1) store_chunk(&mdl, k, OBM + (size_t)b * CH, pol_stream);
        if (clob & 2) store_chunk(&md, k, P0 + (size_t)k * CH, pol_stream);
        any = true;
      }
      if (any) tma_store_commit();
      if (pending >= 0) {
        if (any) tma_store_wait_read<1>();
        else tma_store_wait_read<0>();
        mbar_arrive(&obar[pending]);
        pending = -1;
      }
      if (st) pending = b;
      else mbar_arrive(&obar[b]);
    };
    for (int k = 0; k < pd && k < nch; ++k) issue_plane(k);
    pump();
    for (int t = 0; t < NU; ++t) {
      const int s = t % NS;
      mbar_wait(&ebar[s], (eph >> s) & 1u);
      eph ^= 1u << s;
      last_e = t;
      pump();
      if (t + pd < nch) issue_plane(t + pd);
      if (t >= 1) {
        handle_d(t - 1);
        pump();
      }
    }
    handle_d(NU - 1);
    tma_store_wait_read<0>();  // shared memory must outlive the stores' reads
    return;
  }

  // ================================================================ sweeping warps
  auto ld = [&](const V* chunk, int u) { return t_ld<R, CPT, LB, CM>(chunk, u, lane); };
  auto hand_wait = [&]() { asm volatile("bar.sync %0, 64;" ::"r"(1 + cw) : "memory"); };    // state for this warp is there
  auto hand_over = [&]() { asm volatile("bar.arrive %0, 64;" ::"r"(2 - cw) : "memory"); };  // state for the other warp is there
  auto release_stage = [&](int s) {  // every lane has the stage's operands in registers
    __syncwarp();
    if (lane == 0) mbar_arrive(&ebar[s]);
  };
  // m-th results written to OB[b] (m counts the steps on that buffer): the (m-1)-th must have been read out
  auto wait_ob = [&](int b, int m) {
    if (m >= 1) mbar_wait(&obar[b], (unsigned)(m - 1) & 1u);
  };
  auto publish = [&](int b, bool to_async_proxy) {  // results of a step are in OB[b]
    if (to_async_proxy) fence_proxy_async();
    __syncwarp();
    if (lane == 0) mbar_arrive(&dbar[b]);
  };

  // ---------------------------------------------------------------- forward sweep
  R dprev[CPT], xprev[CPT], uprev[CPT];
#pragma unroll
  for (int j = 0; j < CPT; ++j) {
    dprev[j] = (R)1;
    xprev[j] = (R)0;
    uprev[j] = (R)0;
  }
  for (int k = PP ? cw : 0; k < nch; k += NCW) {
    const int s = k % NS, b = k & 1;
    mbar_wait(&rbar[s], (unsigned)(k / NS) & 1u);
    mbar_wait(&pbar[k], 0u);
    const V* st = RG + (size_t)s * 3 * CH;
    V* p0 = P0 + (size_t)k * CH;
    V* ob = OB + (size_t)b * CH;
    V a[LB], bb[LB], c[LB], x[LB], o0[LB], o1[LB], om[LB];
#pragma unroll
    for (int u = 0; u < LB; ++u) {
      a[u] = ld(st, u);
      c[u] = ld(st + CH, u);
      x[u] = ld(st + 2 * CH, u);
      bb[u] = ld(p0, u);
    }
    release_stage(s);
    if (PP && k > 0) {  // the other warp has finished chunk k-1: take over the chain
      hand_wait();
      const V h0 = HS[lane], h1 = HS[32 + lane], h2 = HS[64 + lane];
#pragma unroll
      for (int j = 0; j < CPT; ++j) {
        dprev[j] = h0.v[j];
        xprev[j] = h1.v[j];
        uprev[j] = h2.v[j];
      }
    }
    R sd[CPT], sx[CPT], su[CPT];
#pragma unroll
    for (int j = 0; j < CPT; ++j) {
      sd[j] = dprev[j];
      sx[j] = xprev[j];
      su[j] = uprev[j];
    }
    bool bad = false;
    if constexpr (CLB) bfb_fwd_chunk_m<R, CPT, LB, false>(a, bb, c, x, o0, o1, om, dprev, xprev, uprev, k == 0, bad);
    else bfb_fwd_chunk<R, CPT, LB, false>(a, bb, c, x, o0, o1, dprev, xprev, uprev, k == 0, bad);
    if (__any_sync(0xffffffffu, bad)) {  // rare operands somewhere in the warp: redo with the full-range division
#pragma unroll
      for (int j = 0; j < CPT; ++j) {
        dprev[j] = sd[j];
        xprev[j] = sx[j];
        uprev[j] = su[j];
      }
      if constexpr (CLB) bfb_fwd_chunk_m<R, CPT, LB, true>(a, bb, c, x, o0, o1, om, dprev, xprev, uprev, k == 0, bad);
      else bfb_fwd_chunk<R, CPT, LB, true>(a, bb, c, x, o0, o1, dprev, xprev, uprev, k == 0, bad);
    }
    if (PP && k < nch - 1) {  // hand the chain on before anything else: the other warp's chunk starts now
      V h0, h1, h2;
#pragma unroll
      for (int j = 0; j < CPT; ++j) {
        h0.v[j] = dprev[j];
        h1.v[j] = xprev[j];
        h2.v[j] = uprev[j];
      }
      HS[lane] = h0;
      HS[32 + lane] = h1;
      HS[64 + lane] = h2;
      hand_over();
    }
#pragma unroll
    for (int u = 0; u < LB; ++u) t_st<R, CPT, LB, CM>(p0, u, lane, o0[u]);
    wait_ob(b, k >> 1);
#pragma unroll
    for (int u = 0; u < LB; ++u) t_st<R, CPT, LB, CM>(ob, u, lane, o1[u]);
    if (CLB && (clob & 1)) {
      V* obm = OBM + (size_t)b * CH;
#pragma unroll
      for (int u = 0; u < LB; ++u) t_st<R, CPT, LB, CM>(obm, u, lane, om[u]);
    }
    if (CLB && clob && !CM && k == nfull) {  // ragged top chunk: a tensor store would spill into the next matrix' rows
      const int64_t col = (int64_t)c0 + lane * CPT;
#pragma unroll
      for (int u = 0; u < LB; ++u) {
        if (u < ntail) {
          const size_t g = (size_t)(row0 + k * LB + u) * a_stride + col;
#pragma unroll
          for (int j = 0; j < CPT; ++j) {
            if (col + j < ncol) {
              if (clob & 1) __stcs(DLg + g + j, om[u].v[j]);
              if (clob & 2) __stcs(Dg + g + j, o0[u].v[j]);
            }
          }
        }
      }
    }
    if (PP && k == nch - 1) hand_over();  // turn-around: the top chunk's d' and x' must be in shared memory first
    publish(b, k <= nch - 3 || clob);     // x' of the last two chunks stays in OB for the way back
  }

  // ---------------------------------------------------------------- back substitution
  // PP: the chunks keep alternating between the two warps across the turn-around (step t belongs to warp t % 2).
  // What a step reads from shared memory was written by the other warp -- the pivots of its chunk always, x' of the
  // last two chunks -- before a hand_over that precedes this step's hand_wait; only for jb <= 1 is that hand_over
  // the one this very step waits for, so those two steps wait before they load and all later ones load first.
  R xv[CPT];
#pragma unroll
  for (int j = 0; j < CPT; ++j) xv[j] = (R)0;
  auto take_x = [&]() {
    hand_wait();
    const V h0 = HS[lane];
#pragma unroll
    for (int j = 0; j < CPT; ++j) xv[j] = h0.v[j];
  };
  auto give_x = [&]() {
    V h0;
#pragma unroll
    for (int j = 0; j < CPT; ++j) h0.v[j] = xv[j];
    HS[lane] = h0;
    hand_over();
  };
  for (int jb = PP ? ((nch + cw) & 1) : 0; jb < nch; jb += NCW) {
    const int k = nch - 1 - jb, u0 = nch + jb, s = u0 % NS, b = k & 1;
    // steps on OB[b] before this one: the forward chunks of parity b, then the backward chunks of parity b above k
    const int kmax = (nch - 1) - ((((nch - 1) & 1) != b) ? 1 : 0);
    const int m = ((nch - b + 1) >> 1) + ((kmax - k) >> 1);
    mbar_wait(&rbar[s], (unsigned)(u0 / NS) & 1u);
    const V* st = RG + (size_t)s * 3 * CH;
    const V* bu = st + CH;
    V* ob = OB + (size_t)b * CH;
    const V* xs = jb <= 1 ? ob : st + 2 * CH;  // x' of the last two chunks never left the chip
    const V* q0p = P0 + (size_t)k * CH;
    if (PP && jb == 0) hand_wait();  // xv stays 0
    if (PP && jb == 1) take_x();
    if (k == nfull) {
      // partial top chunk (ntail levels; jb == 0)
      const int64_t col = (int64_t)c0 + lane * CPT;
      V y[LB], m0[LB], ui[LB];
#pragma unroll
      for (int u = 0; u < LB; ++u) {
        if (u < ntail) {
          y[u] = ld(xs, u);
          m0[u] = ld(q0p, u);
          ui[u] = ld(bu, u);
        }
      }
      release_stage(s);
      wait_ob(b, m);
#pragma unroll
      for (int u = LB - 1; u >= 0; --u) {
        if (u < ntail) {
          V o;
#pragma unroll
          for (int j = 0; j < CPT; ++j) {
            const R uu = (u == ntail - 1) ? (R)0 : ui[u].v[j];  // du(nrow-1) is never used (:24-27)
            xv[j] = x_div(x_sub(y[u].v[j], x_mul(uu, xv[j])), m0[u].v[j]);  // :317-319
            o.v[j] = xv[j];
          }
          if (CM) {
            t_st<R, CPT, LB, CM>(ob, u, lane, o);
          } else {  // a tensor store would spill into the next matrix' rows -> plain streaming stores
            R* g = X + (size_t)(row0 + k * LB + u) * x_stride + col;
#pragma unroll
            for (int j = 0; j < CPT; ++j)
              if (col + j < ncol) __stcs(g + j, o.v[j]);
          }
        }
      }
      if (PP && jb < nch - 1) give_x();
      publish(b, CM);
    } else {
      V y[LB], m0[LB], ui[LB];
#pragma unroll
      for (int u = 0; u < LB; ++u) {
        y[u] = ld(xs, u);
        m0[u] = ld(q0p, u);
        ui[u] = ld(bu, u);
      }
      release_stage(s);
      if (jb == 0) {  // this chunk ends with the last level
#pragma unroll
        for (int j = 0; j < CPT; ++j) ui[LB - 1].v[j] = (R)0;
      }
      if (PP && jb > 1) take_x();
      R sxv[CPT];
      V ys[LB];
#pragma unroll
      for (int j = 0; j < CPT; ++j) sxv[j] = xv[j];
#pragma unroll
      for (int u = 0; u < LB; ++u) ys[u] = y[u];
      bool bad = false;
      bfb_bwd_chunk<R, CPT, LB, false>(y, m0, ui, xv, bad);
      if (__any_sync(0xffffffffu, bad)) {
#pragma unroll
        for (int j = 0; j < CPT; ++j) xv[j] = sxv[j];
#pragma unroll
        for (int u = 0; u < LB; ++u) y[u] = ys[u];
        bfb_bwd_chunk<R, CPT, LB, true>(y, m0, ui, xv, bad);
      }
      if (PP && jb < nch - 1) give_x();
      wait_ob(b, m);
#pragma unroll
      for (int u = 0; u < LB; ++u) t_st<R, CPT, LB, CM>(ob, u, lane, y[u]);
      publish(b, true);
    }
  }
}

// ------------------------------------------------------------------------------------------- host side
template <typename R, int NS, bool CM, bool PP, bool CLB>
static int launch_stream_(R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nmat, int64_t a_stride, int64_t x_stride,
                         int pd, size_t smem, cudaStream_t stream, const char* name, int clob) {
  constexpr int CPT = TmaCfg<R>::CPT, LB = TmaGeom<R, CM>::LB, W = 32 * CPT;
  CUtensorMap mdl, md, mdu, mx;
  bool ok;
  const int swz = CM && LB * sizeof(R) == 64 ? 1 : 0;
  if (CM) {  // [ncol][stride] rows of which the first nlev entries are the column: fastest dimension = level
    ok = make_tensor_map_2d<R>(&mdl, dl, nlev, ncol, a_stride, LB, W, swz) && make_tensor_map_2d<R>(&md, d, nlev, ncol, a_stride, LB, W, swz) &&
         make_tensor_map_2d<R>(&mdu, du, nlev, ncol, a_stride, LB, W, swz) && make_tensor_map_2d<R>(&mx, X, nlev, ncol, x_stride, LB, W, swz);
  } else {
    const int64_t rows = (int64_t)nmat * nlev;
    ok = make_tensor_map_2d<R>(&mdl, dl, ncol, rows, a_stride, W, LB, 0) && make_tensor_map_2d<R>(&md, d, ncol, rows, a_stride, W, LB, 0) &&
         make_tensor_map_2d<R>(&mdu, du, ncol, rows, a_stride, W, LB, 0) && make_tensor_map_2d<R>(&mx, X, ncol, rows, x_stride, W, LB, 0);
  }
  if (!ok) return EKAT_B200_ERR_UNSUPPORTED;
  auto k = k_tridiag_bfb_stream<R, CPT, NS, LB, CM, PP, CLB>;
  int rc = allow_smem(k, smem);
  if (rc) return rc;
  dim3 grid((unsigned)((ncol + W - 1) / W), (unsigned)nmat);
  k<<<grid, PP ? 96 : 64, smem, stream>>>(mdl, md, mdu, mx, X, ncol, nlev, x_stride, pd, clob, dl, d, a_stride);
  note_launch(name);
  return check_cuda(cudaGetLastError());
}

template <typename R, int NS, bool CM, bool PP>
static int launch_stream(R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nmat, int64_t a_stride, int64_t x_stride,
                         int pd, size_t smem, cudaStream_t stream, const char* name, int clob) {
  if (!CM && clob) return launch_stream_<R, NS, CM, PP, true>(dl, d, du, X, ncol, nlev, nmat, a_stride, x_stride, pd, smem, stream, name, clob);
  return launch_stream_<R, NS, CM, PP, false>(dl, d, du, X, ncol, nlev, nmat, a_stride, x_stride, pd, smem, stream, name, 0);
}

template <typename R, bool CM>
int tridiag_bfb_stream(R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nmat, int64_t a_stride, int64_t x_stride,
                       cudaStream_t stream, int clob) {
  if (CM && clob) return EKAT_B200_ERR_UNSUPPORTED;
  constexpr int CPT = TmaCfg<R>::CPT, W = 32 * CPT, LB = TmaGeom<R, CM>::LB;
  const Tuning tu = tuning();
  const int nch = (nlev + LB - 1) / LB;
  const size_t chunk = (size_t)LB * W * sizeof(R);
  const size_t al = CM ? 1024 : 128;
  auto smem_for = [&](int stages) {
    const size_t hs = (((size_t)(2 * stages + 4 + nch) * 8 + 15) & ~(size_t)15) + 3 * 32 * sizeof(R) * CPT;  // barriers + hand-over state
    return ((hs + al - 1) & ~(al - 1)) + (size_t)(nch + 3 * stages + 2 + ((clob & 1) ? 2 : 0)) * chunk;
  };
  // tiles per SM: shared memory, and four CTAs of 96 threads x 150 registers fill the register file (capping the
  // registers at 128 for a fifth tile costs more than the tile brings: measured 77% against 82%)
  auto tiles = [&](int stages) { return tmin(4, (int)((kMaxSmemPerCta + 1024) / (smem_for(stages) + 1024))); };
  // as many tiles per SM as possible, then as many stages as that residency leaves room for (2 .. 6)
  int ns = 2;
  if (smem_for(2) > kMaxSmemPerCta) return EKAT_B200_ERR_UNSUPPORTED;
  for (int c = 3; c <= 6; ++c)
    if (tiles(c) == tiles(2)) ns = c;
  if (tu.cols_per_cta >= 2 && tu.cols_per_cta <= 6 && smem_for(tu.cols_per_cta) <= kMaxSmemPerCta) ns = tu.cols_per_cta;
  const int pd = tu.prefetch > 0 ? tu.prefetch : 4;
  const bool pp = tu.variant != kVarStreamOneWarp;  // one sweeping warp (A/B measurements)
  const bool f64 = sizeof(R) == 8;
  const char* name = CM ? (f64 ? "tridiag_bfb_stream_cm<f64>" : "tridiag_bfb_stream_cm<f32>")
                        : (f64 ? "tridiag_bfb_stream<f64>" : "tridiag_bfb_stream<f32>");
#define EKB_STREAM(NS_)                                                                                                    \
  (pp ? launch_stream<R, NS_, CM, true>(dl, d, du, X, ncol, nlev, nmat, a_stride, x_stride, pd, smem_for(NS_), stream, name, clob) \
      : launch_stream<R, NS_, CM, false>(dl, d, du, X, ncol, nlev, nmat, a_stride, x_stride, pd, smem_for(NS_), stream, name, clob))
  switch (ns) {
    case 2: return EKB_STREAM(2);
    case 3: return EKB_STREAM(3);
    case 4: return EKB_STREAM(4);
    case 5: return EKB_STREAM(5);
    default: return EKB_STREAM(6);
  }
#undef EKB_STREAM
}

template int tridiag_bfb_stream<double, false>(double*, double*, double*, double*, int64_t, int, int, int64_t, int64_t, cudaStream_t, int);
template int tridiag_bfb_stream<float, false>(float*, float*, float*, float*, int64_t, int, int, int64_t, int64_t, cudaStream_t, int);
template int tridiag_bfb_stream<double, true>(double*, double*, double*, double*, int64_t, int, int, int64_t, int64_t, cudaStream_t, int);
template int tridiag_bfb_stream<float, true>(float*, float*, float*, float*, int64_t, int, int, int64_t, int64_t, cudaStream_t, int);

}  // namespace ekb
