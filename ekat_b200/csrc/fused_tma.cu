// fused_tma.cu -- BASELINE config 4 in ONE kernel: the shared-matrix Thomas solve of nrhs right-hand sides followed
// by the vertical remap of the nrhs solutions, a tile of 16 columns kept on chip, nothing but the inputs and the
// remapped fields crossing HBM (25,600 B per column at 128 levels x 10 right-hand sides, f64).
//
// Replaces, back to back, the per-column loops around ekat::tridiag::thomas(team, dl, d, du, X[nrow][nrhs])
// (reference src/algorithm/ekat_tridiag.hpp:177-214, :323-334) and LinInterp::setup / lin_interp on each solution
// (src/algorithm/ekat_lin_interp_impl.hpp:89-186). Level-major operands (columns fastest).
//
// Why this shape. The two-pass path (tridiag_tma_shared.cu, then lin_interp_lm.cu) sends the solutions through HBM
// once in each direction: 45.5 KB per column for 25.6 KB algorithmic. Here a persistent CTA (one per SM) keeps every
// plane [level][16 columns] of its current tile in shared memory -- dl -> multipliers, d -> reciprocal pivots,
// du -> c', the nrhs right-hand sides -> solutions, and the source coordinates x1: 14 planes of 16 KB at 128 levels
// f64 = 224 KB, all an SM has. There is no room for a second tile, so the pipeline runs plane by plane instead: each
// plane is handed to the next tile the moment the current tile is done with it (see k_solve_remap_roles).
//
// Arithmetic: thomas as in tridiag_tma_shared.cu (ratio form, <= 3e-15 relative to the reference); LinInterp the
// reference's four statements in order with IEEE division (div_spec fast path == __ddiv_rn, full-range redo
// otherwise) -- bit-identical to ekat::LinInterp applied to the same solutions.
//
// What was tried on the way (B200, f64, 1,572,864 x 128 -> 128, 10 rhs; two-pass = 14.7 ms):
//   one CTA per tile, three CTA-wide phases (search | factor + sweeps | remap)          13.6 ms
//   the same, 8-column tiles so that two CTAs share an SM                               14.6 ms
//   persistent CTA, CTA-wide phases, next tile's copies issued as planes free up        13.3 ms
//   role-split warps (this file)                                                        12.7 ms
//   + full-range division out of line, full-chunk fast paths, warp roles grouped by scheduler   10.8 ms
//   + branch-free search / remap so that a thread's seven targets interleave             9.8 ms
//   (- software-pipelined sweeps 10.6 ms; - L2 prefetch of the next tile's planes 10.5-13 ms: both dropped)
#include <cuda.h>

#include <type_traits>

#include "lin_interp_search.cuh"
#include "thomas_two_sided.cuh"

namespace ekb {


// The role-split form (the default): no CTA-wide barrier after the prologue. NIW remap warps, NSW solver warps and one
// factor warp each walk the CTA's tiles at their own pace and hand planes to each other through mbarriers:
//   remap warps   : wait x1 -> bracket search (the last one out requests x1 of the next tile) -> for every solution f:
//                   wait "solved f" -> remap and store field f -> the last warp out requests right-hand side f of the
//                   next tile into the same plane;
//   solver warps  : lane = (rhs, column); wait for the rhs plane to land, follow the factor warp chunk by chunk down the
//                   column, sweep back up, signal "solved r"; after their last rhs, signal "factor planes free";
//   factor warp   : wait "factor planes free" -> request dl, d, du of its next tile -> factorise in place, one signal
//                   per chunk of 8 levels.
// So the sweeps of tile i+1 run in the planes the remap of tile i has just left, under the remap's stores and the next
// copies, and HBM stays busy except during the bracket search.
template <typename R, int LOG2P, int TW, int NIW, int NSW, int NT>
__global__ void __launch_bounds__(32 * (NIW + NSW + 1), 1)
k_solve_remap_roles(const __grid_constant__ CUtensorMap mdl, const __grid_constant__ CUtensorMap md,
                    const __grid_constant__ CUtensorMap mdu, const __grid_constant__ CUtensorMap mb,
                    const __grid_constant__ CUtensorMap mx1, const R* __restrict__ x2, R* __restrict__ Y2,
                    R* __restrict__ X, int64_t ncol, int nlev, int km2, int nrhs, int64_t pitch, int P, int ntiles,
                    int hb) {
  constexpr int WARPS = NIW + NSW + 1, LB = kLB, RN = TmaCfg<R>::RN, TPS = 32 / TW;
  constexpr unsigned B = sizeof(R), ROW = TW * B;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int tid = threadIdx.x, lane = tid & 31;
  // Role of each warp: 0 .. NIW-1 remap, NIW .. NIW+NSW-1 solver, WARPS-1 factor. A warp issues from scheduler
  // (warp id % 4); with 10 + 5 + 1 warps the latency-bound warps -- the factor warp and the solvers of the planes that
  // are released last -- share scheduler 3, the other two solvers sit on scheduler 2 with two remap warps, and the
  // other eight remap warps fill schedulers 0 and 1, so the recurrences do not queue behind the remap's instructions.
  const int warp = (NIW == 10 && NSW == 5) ? (int)((0xF987E654DB32CA10ull >> (4 * (tid >> 5))) & 15) : tid >> 5;
  const int lc = lane % TW, sub = lane / TW;
  const int nch = (nlev + LB - 1) / LB;
  // barriers: [0] x1 landed, [1] dl+d+du landed, [2] factor planes free (NSW arrivals), then per rhs: landed[r],
  // solved[r], then per chunk: factored[k]; then the nrhs + 1 "last one out" counters
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw);
  uint64_t* abar = bars + 2;
  uint64_t* lbar = bars + 3;
  uint64_t* sbar = lbar + nrhs;
  uint64_t* fbar = sbar + nrhs;
  uint64_t* bbar = fbar + nch;  // lower half of dl, d, du landed (when they come as two boxes of hb levels)
  int* cnt = reinterpret_cast<int*>(bbar + 1);
  const size_t bar_bytes = ((size_t)(4 + 2 * nrhs + nch) * 8 + (size_t)(nrhs + 1) * 4 + 127) & ~(size_t)127;
  const size_t plane = (((size_t)nlev * ROW + 127) & ~(size_t)127) / B;  // elements per plane
  R* PM = reinterpret_cast<R*>(smem_raw + bar_bytes);  // dl -> multipliers
  R* PR = PM + plane;                                  // d  -> reciprocal pivots
  R* PC = PR + plane;                                  // du -> c'
  R* PX1 = PC + plane;                                 // x1
  R* PX = PX1 + plane;                                 // [nrhs] right-hand sides -> solutions
  const unsigned plane_bytes = (unsigned)nlev * ROW;
  const int first = blockIdx.x, step = (int)gridDim.x;  // grid <= ntiles
  // dl, d, du of a tile: one box of nlev levels per plane, or (hb < nlev) two boxes of hb levels with a barrier each,
  // so that the factor warp -- on the critical path between two tiles -- can start when the upper half is there
  auto load_A = [&](int tile) {
    const unsigned half = (unsigned)hb * ROW;
    mbar_expect_tx(&bars[1], 3u * half);
    tma_load_2d(PM, &mdl, tile * TW, 0, &bars[1]);
    tma_load_2d(PR, &md, tile * TW, 0, &bars[1]);
    tma_load_2d(PC, &mdu, tile * TW, 0, &bars[1]);
    if (hb < nlev) {
      mbar_expect_tx(bbar, 3u * half);
      tma_load_2d(PM + hb * TW, &mdl, tile * TW, hb, bbar);
      tma_load_2d(PR + hb * TW, &md, tile * TW, hb, bbar);
      tma_load_2d(PC + hb * TW, &mdu, tile * TW, hb, bbar);
    }
  };
  const int nsw = tmin(NSW, (nrhs + TPS - 1) / TPS);     // solver warps that have a right-hand side

  if (tid == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    mbar_init(bbar, 1);
    mbar_init(abar, nsw);
    for (int i = 0; i < 2 * nrhs + nch; ++i) mbar_init(&lbar[i], 1);
    for (int r = 0; r <= nrhs; ++r) cnt[r] = 0;
    fence_mbar_init();
    mbar_expect_tx(&bars[0], plane_bytes);
    tma_load_2d(PX1, &mx1, first * TW, 0, &bars[0]);
    load_A(first);
    for (int r = 0; r < nrhs; ++r) {
      mbar_expect_tx(&lbar[r], plane_bytes);
      tma_load_2d(PX + (size_t)r * plane, &mb, first * TW, r * nlev, &lbar[r]);
    }
  }
  __syncthreads();

  if (warp == WARPS - 1) {
    // ================================================================ factor warp (with TW = 16 both half-warps
    // compute the same columns and write the same values): ratio form of tridiag_tma_shared.cu, 8 levels at a time
    unsigned par = 0;
    for (int tile = first; tile < ntiles; tile += step, par ^= 1u) {
      if (tile != first) {
        mbar_wait(abar, par ^ 1u);  // the solvers are done with the previous tile's factors
        if (lane == 0) {
          fence_proxy_async();
          load_A(tile);
        }
      }
      mbar_wait(&bars[1], par);
      if (hb < nlev) mbar_wait(bbar, par);  // both ends are needed at once
      auto done_step = [&](int t) {
        __syncwarp();
        if (lane == 0) mbar_arrive(&fbar[t]);
      };
      if (nlev % LB == 0) two_sided_factor<R, true>(PM + lc, PR + lc, PC + lc, TW, nlev, done_step);
      else two_sided_factor<R, false>(PM + lc, PR + lc, PC + lc, TW, nlev, done_step);
    }
  } else if (warp >= NIW) {
    // ================================================================ solver warps: lane = (rhs within the warp, column)
    const int sw = warp - NIW;
    if (sw >= nsw) return;
    unsigned par = 0;
    for (int tile = first; tile < ntiles; tile += step, par ^= 1u) {
      for (int r0 = sw * TPS; r0 < nrhs; r0 += nsw * TPS) {
        const int r = r0 + sub;
        if (r < nrhs) {
          R* px = PX + (size_t)r * plane + lc;
          mbar_wait(&lbar[r], par);
          auto wait_step = [&](int t) { mbar_wait(&fbar[t], par); };
          if (nlev % LB == 0) two_sided_sweeps<R, true>(PM + lc, PR + lc, PC + lc, TW, px, TW, nlev, wait_step);
          else two_sided_sweeps<R, false>(PM + lc, PR + lc, PC + lc, TW, px, TW, nlev, wait_step);
        }
        __syncwarp();
        if (r < nrhs && lc == 0) mbar_arrive(&sbar[r]);
      }
      if (lane == 0) mbar_arrive(abar);
    }
  } else {
    // ================================================================ remap warps
    const unsigned x1a = smem_u32(PX1) + (unsigned)lc * B;
    const unsigned last_off = (unsigned)(nlev - 1) * ROW;
    const int j0 = warp * TPS + sub;
    const size_t tstride = (size_t)(NIW * TPS) * pitch;
    R xt[NT];
    {
      const int64_t c = (int64_t)first * TW + lc;
      const int64_t gc = c < ncol ? c : ncol - 1;
#pragma unroll
      for (int t = 0; t < NT; ++t) {
        const int j = tmin(warp * TPS + sub + t * (NIW * TPS), km2 - 1);  // surplus slots redo the last target
        xt[t] = __ldg(x2 + (size_t)j * pitch + gc);
      }
    }
    unsigned par = 0;
    for (int tile = first; tile < ntiles; tile += step, par ^= 1u) {
      const int64_t col = (int64_t)tile * TW + lc;
      const bool live = col < ncol;
      const int next = tile + step;
      const bool has_next = next < ntiles;

      // -------------------------------------------------------------- brackets
      unsigned off[NT];
      R dxt[NT], dx[NT], rc[NT];
      mbar_wait(&bars[0], par);
#pragma unroll
      for (int t = 0; t < NT; ++t) {
        // no branch on j < km2: the seven searches of a thread are independent chains of shared-memory probes and
        // must be free to interleave (surplus slots hold a copy of the last target and are never stored)
        const unsigned o = bracket_off<R, LOG2P, ROW>(x1a, nlev, P, xt[t]);
        const unsigned oph = o == last_off ? o - ROW : o + ROW;  // k1ph, ekat_lin_interp_impl.hpp:119-131
        const R xa = lds_at(x1a + o, R()), xb = lds_at(x1a + oph, R());
        off[t] = o;
        dxt[t] = x_sub(xt[t], xa);
        dx[t] = x_sub(xb, xa);
        rc[t] = rcp_refined(dx[t]);
        pin(dxt[t]);  // left to itself the compiler recomputes these per field instead of keeping them
        pin(dx[t]);
        pin(rc[t]);
      }
      __syncwarp();
      if (lane == 0) {  // the last warp out of the search asks for x1 of the next tile
        __threadfence_block();
        if (atomicAdd(&cnt[nrhs], 1) == NIW - 1) {
          cnt[nrhs] = 0;
          if (has_next) {
            __threadfence_block();
            fence_proxy_async();
            mbar_expect_tx(&bars[0], plane_bytes);
            tma_load_2d(PX1, &mx1, next * TW, 0, &bars[0]);
          }
        }
      }
      if (has_next) {
        const int64_t c = (int64_t)next * TW + lc;
        const int64_t gc = c < ncol ? c : ncol - 1;
#pragma unroll
        for (int t = 0; t < NT; ++t) {
          const int j = tmin(warp * TPS + sub + t * (NIW * TPS), km2 - 1);
          xt[t] = __ldg(x2 + (size_t)j * pitch + gc);
        }
      }

      // -------------------------------------------------------------- remap of every solution as it completes
      R* gy0 = Y2 + (size_t)j0 * pitch + col;
      for (int f = 0; f < nrhs; ++f) {
        mbar_wait(&sbar[f], par);
        if (X != nullptr) {  // (optional) the solution itself
          const R* px = PX + (size_t)f * plane + lc;
          R* gx = X + ((size_t)f * nlev) * pitch + col;
          for (int l = warp * TPS + sub; l < nlev; l += NIW * TPS)
            if (live) st_stream(gx + (size_t)l * pitch, px[l * TW]);
        }
        const unsigned y1a = smem_u32(PX + (size_t)f * plane) + (unsigned)lc * B;
        R* gy = gy0;  // row (f, j0) of this thread's column; the targets of a thread are NIW * TPS rows apart
        gy0 += (size_t)km2 * pitch;
        bool bad = false;
#pragma unroll
        for (int t = 0; t < NT; ++t) {
          const unsigned o = off[t];
          const unsigned oph = o == last_off ? o - ROW : o + ROW;
          const R ya = lds_at(y1a + o, R()), yb = lds_at(y1a + oph, R());
          R v = x_sub(yb, ya);                   // ekat_lin_interp_impl.hpp:153
          v = x_mul(v, dxt[t]);                  // :154
          v = div_spec(v, dx[t], rc[t], bad);    // :155 (fast-path quotient; `bad` records operands it is not proven for)
          v = x_add(v, ya);                      // :156
          if (live && j0 + t * (NIW * TPS) < km2) st_stream(gy + t * tstride, v);
        }
        if (__builtin_expect(bad, 0)) {  // rare operands: the same statements with the full-range division
#pragma unroll
          for (int t = 0; t < NT; ++t) {
            const unsigned o = off[t];
            const unsigned oph = o == last_off ? o - ROW : o + ROW;
            const R ya = lds_at(y1a + o, R()), yb = lds_at(y1a + oph, R());
            R v = x_sub(yb, ya);
            v = x_mul(v, dxt[t]);
            v = x_div_full(v, dx[t]);
            v = x_add(v, ya);
            if (live && j0 + t * (NIW * TPS) < km2) st_stream(gy + t * tstride, v);
          }
        }
        __syncwarp();
        if (lane == 0) {  // the last warp to leave solution f asks for right-hand side f of the next tile in its place
          __threadfence_block();
          if (atomicAdd(&cnt[f], 1) == NIW - 1) {
            cnt[f] = 0;
            if (has_next) {
              __threadfence_block();
              fence_proxy_async();
              mbar_expect_tx(&lbar[f], plane_bytes);
              tma_load_2d(PX + (size_t)f * plane, &mb, next * TW, f * nlev, &lbar[f]);
            }
          }
        }
      }
    }
  }
}

constexpr int kFtRemapWarps = 10, kFtSolverWarps = 5;  // + 1 factor warp = 512 threads

template <typename R, int TW>
static int launch_solve_remap(const R* dl, const R* d, const R* du, const R* Bm, const R* x1, const R* x2, R* Y2, R* X,
                              int64_t ncol, int nlev, int km2, int nrhs, int64_t pitch, size_t smem, cudaStream_t stream) {
  constexpr int PER_STEP = kFtRemapWarps * (32 / TW);       // targets of a column the remap warps cover per step
  constexpr int NT = (128 + PER_STEP - 1) / PER_STEP;       // km2 <= 128
  constexpr int THREADS = 32 * (kFtRemapWarps + kFtSolverWarps + 1);
  CUtensorMap mdl, md, mdu, mb, mx1;
  const int hb = (nlev % 2 == 0 && nlev >= 32 && tuning().prefetch != 1) ? nlev / 2 : nlev;  // levels per box of dl, d, du
  if (!make_tensor_map_2d<R>(&mdl, dl, ncol, nlev, pitch, TW, hb, 0) || !make_tensor_map_2d<R>(&md, d, ncol, nlev, pitch, TW, hb, 0) ||
      !make_tensor_map_2d<R>(&mdu, du, ncol, nlev, pitch, TW, hb, 0) ||
      !make_tensor_map_2d<R>(&mb, Bm, ncol, (int64_t)nrhs * nlev, pitch, TW, nlev, 0) ||
      !make_tensor_map_2d<R>(&mx1, x1, ncol, nlev, pitch, TW, nlev, 0))
    return EKAT_B200_ERR_UNSUPPORTED;
  int P = 1, log2p = 0;
  while (2 * P <= nlev) {
    P *= 2;
    ++log2p;
  }
  const int64_t ntiles = (ncol + TW - 1) / TW;
  const unsigned grid = (unsigned)tmin<int64_t>(ntiles, kNumSms);
  int rc;
#define EKB_FT_LAUNCH(L2_)                                                                                     \
  {                                                                                                            \
    auto k = k_solve_remap_roles<R, L2_, TW, kFtRemapWarps, kFtSolverWarps, NT>;                               \
    if ((rc = allow_smem(k, smem))) return rc;                                                                 \
    k<<<grid, THREADS, smem, stream>>>(mdl, md, mdu, mb, mx1, x2, Y2, X, ncol, nlev, km2, nrhs, pitch, P,       \
                                       (int)ntiles, hb);                                                       \
  }
  if (log2p == 6) EKB_FT_LAUNCH(6)
  else if (log2p == 7) EKB_FT_LAUNCH(7)
  else EKB_FT_LAUNCH(0)
#undef EKB_FT_LAUNCH
  note_launch(sizeof(R) == 8 ? "solve_remap_tma<f64>" : "solve_remap_tma<f32>");
  return check_cuda(cudaGetLastError());
}

// Returns EKAT_B200_ERR_UNSUPPORTED when the tile does not fit on chip, the operands do not meet TMA's requirements or
// the shape is one the two-pass path serves better (a single right-hand side: nothing to amortise the search over).
template <typename R>
int solve_remap_tma(const R* dl, const R* d, const R* du, const R* Bm, const R* x1, const R* x2, R* Y2, R* X, int64_t ncol,
                    int nlev, int km2, int nrhs, int64_t pitch, cudaStream_t stream) {
  if (ncol == 0 || nrhs == 0) return EKAT_B200_OK;
  auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  if (!al(dl) || !al(d) || !al(du) || !al(Bm) || !al(x1) || (pitch * sizeof(R)) % 16) return EKAT_B200_ERR_UNSUPPORTED;
  if (nlev < 2 || nlev > 256 || km2 < 1 || km2 > 128) return EKAT_B200_ERR_UNSUPPORTED;
  if (ncol >= (int64_t)1 << 31 || (int64_t)nrhs * nlev >= (int64_t)1 << 31) return EKAT_B200_ERR_UNSUPPORTED;
  // few right-hand sides: too little to amortise the search and the factorisation over, the two-pass path is faster
  // (345,600 x 128, two-pass vs this kernel: f64 3 rhs 1.20 vs 1.49 ms, 4 rhs 1.45 vs 1.55, 5 rhs 1.61 vs 1.56,
  //  6 rhs 2.15 vs 1.63; f32 1 rhs 0.54 vs 0.50, 2 rhs 0.79 vs 0.57)
  if (nrhs < (sizeof(R) == 8 ? 5 : 2) && tuning().cols_per_cta == 0) return EKAT_B200_ERR_UNSUPPORTED;
  const int nch = (nlev + kLB - 1) / kLB;
  auto smem_for = [&](int tw) {
    const size_t plane = ((size_t)nlev * tw * sizeof(R) + 127) & ~(size_t)127;
    return (((size_t)(4 + 2 * nrhs + nch) * 8 + (size_t)(nrhs + 1) * 4 + 127) & ~(size_t)127) + (size_t)(4 + nrhs) * plane;
  };
  // 128-byte rows: 16 columns of f64, 32 of f32 (tuning: cols_per_cta = 16 forces 16 for f32 as well)
  if constexpr (sizeof(R) == 4) {
    if (tuning().cols_per_cta != 16 && smem_for(32) <= kMaxSmemPerCta)
      return launch_solve_remap<R, 32>(dl, d, du, Bm, x1, x2, Y2, X, ncol, nlev, km2, nrhs, pitch, smem_for(32), stream);
  }
  if (smem_for(16) > kMaxSmemPerCta) return EKAT_B200_ERR_UNSUPPORTED;
  return launch_solve_remap<R, 16>(dl, d, du, Bm, x1, x2, Y2, X, ncol, nlev, km2, nrhs, pitch, smem_for(16), stream);
}

template int solve_remap_tma<double>(const double*, const double*, const double*, const double*, const double*, const double*,
                                     double*, double*, int64_t, int, int, int, int64_t, cudaStream_t);
template int solve_remap_tma<float>(const float*, const float*, const float*, const float*, const float*, const float*,
                                    float*, float*, int64_t, int, int, int, int64_t, cudaStream_t);

}  // namespace ekb
