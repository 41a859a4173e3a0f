// lin_interp.cu -- fused bracket search + linear interpolation, reference layout.
//
// Replaces the per-column loops around ekat::LinInterp::setup and ::lin_interp
// (reference: src/algorithm/ekat_lin_interp_impl.hpp:160-186 setup_impl with
// src/kokkos/ekat_upper_bound.hpp:28-50, and :89-158 lin_interp_impl).
//
// Layout is the reference's: x1,y1 [ncol][pitch1], x2,y2 [ncol][pitch2], so a tile
// of CB whole columns is ONE contiguous range per array. A persistent CTA streams
// tiles through a 2-stage shared-memory ring with TMA bulk copies
// (cp.async.bulk + mbarrier complete_tx): while the warps search/interpolate tile n
// out of stage n&1, the copy engine fills the other stage with tile n+1. Source
// coordinates, source fields and target coordinates are all staged; the index map
// never leaves the SM unless the caller asks for it (m_indx_map parity).
// Each lane owns one target level: a 32-wide group of consecutive targets of one
// column per warp step, so x2 reads, the optional map and y2 writes are coalesced
// and the bracket gathers hit neighbouring shared-memory words.
//
// Arithmetic is the reference's, unfused and in its order (:153-156):
//     y2  = y1[k1ph] - y1[k1];  y2 *= x2 - x1[k1];  y2 /= x1[k1ph] - x1[k1];  y2 += y1[k1]
// so results are bit-identical to the reference at -ffp-contract=off, for any Pack size
// (pack padding is skipped: only the first km1 / km2 entries of a row are touched as data).
#include <cstdio>
#include <cstdlib>

#include "lin_interp_search.cuh"

namespace ekb {

enum LiMode { LI_FUSED = 0, LI_SETUP = 1, LI_APPLY = 2 };

struct LiArgs {
  const void *x1, *x2, *y1;
  void* y2;
  int* indx;
  const int* indx_in;
  int64_t ncol, pitch1, pitch2, y1_fstride, y2_fstride;
  int km1, km2, nfields, CB, use_tma, P;
};

template <typename R, int MODE, int LOG2P>
__global__ void __launch_bounds__(256) k_lin_interp(const LiArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t bars[2];

  const R* __restrict__ gx1 = static_cast<const R*>(a.x1);
  const R* __restrict__ gx2 = static_cast<const R*>(a.x2);
  const R* __restrict__ gy1 = static_cast<const R*>(a.y1);
  R* __restrict__ gy2 = static_cast<R*>(a.y2);
  const int km1 = a.km1, km2 = a.km2, CB = a.CB;
  const int nf = MODE == LI_SETUP ? 0 : a.nfields;
  const bool need_x2 = true;  // x2 drives the search (setup/fused) and enters the formula (apply/fused)
  // stage layout: x1[CB*pitch1] | y1[nf][CB*pitch1] | x2[CB*pitch2], each 16B-aligned
  const size_t n1 = (size_t)CB * a.pitch1, n2 = (size_t)CB * a.pitch2;
  const size_t n1p = (n1 * sizeof(R) + 15) / 16 * 16 / sizeof(R), n2p = (n2 * sizeof(R) + 15) / 16 * 16 / sizeof(R);
  const size_t stage_elems = n1p * (1 + nf) + n2p;
  R* stage0 = reinterpret_cast<R*>(smem_raw);

  const int64_t ntiles = (a.ncol + CB - 1) / CB;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;

  if (tid == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    fence_mbar_init();
  }
  __syncthreads();

  // issue the loads of tile `tile` into stage `s` (TMA when the tile is full and aligned)
  auto issue = [&](int64_t tile, int s) {
    const int64_t c0 = tile * CB;
    const bool tma = a.use_tma && c0 + CB <= a.ncol;
    if (tma && tid != 0) return true;  // one thread feeds the copy engine; nobody else spends an instruction here
    const int nc = (int)tmin<int64_t>(CB, a.ncol - c0);
    R* sx1 = stage0 + (size_t)s * stage_elems;
    R* sy1 = sx1 + n1p;
    R* sx2 = sy1 + n1p * nf;
    if (tma) {
      {
        fence_proxy_async();
        const unsigned b1 = (unsigned)(n1 * sizeof(R)), b2 = (unsigned)(n2 * sizeof(R));
        mbar_expect_tx(&bars[s], b1 * (1 + nf) + (need_x2 ? b2 : 0));
        tma_load_1d(sx1, gx1 + (size_t)c0 * a.pitch1, b1, &bars[s]);
        for (int f = 0; f < nf; ++f)
          tma_load_1d(sy1 + n1p * f, gy1 + (size_t)f * a.y1_fstride + (size_t)c0 * a.pitch1, b1, &bars[s]);
        if (need_x2) tma_load_1d(sx2, gx2 + (size_t)c0 * a.pitch2, b2, &bars[s]);
      }
    } else {
      // ragged or unaligned tile: plain coalesced loads (only real data, never the pack padding)
      for (int c = warp; c < nc; c += nwarp) {
        for (int k = lane; k < km1; k += 32) {
          sx1[(size_t)c * a.pitch1 + k] = gx1[(size_t)(c0 + c) * a.pitch1 + k];
          for (int f = 0; f < nf; ++f)
            sy1[n1p * f + (size_t)c * a.pitch1 + k] = gy1[(size_t)f * a.y1_fstride + (size_t)(c0 + c) * a.pitch1 + k];
        }
        for (int k = lane; k < km2; k += 32) sx2[(size_t)c * a.pitch2 + k] = gx2[(size_t)(c0 + c) * a.pitch2 + k];
      }
    }
    return tma;
  };

  unsigned phase_bits = 0;  // bit s = parity to wait for on stage s
  int64_t tile = blockIdx.x;
  bool cur_tma = false;
  if (tile < ntiles) cur_tma = issue(tile, 0);
  for (int it = 0; tile < ntiles; tile += gridDim.x, ++it) {
    const int s = it & 1;
    const int64_t next = tile + gridDim.x;
    bool next_tma = false;
    // stage s^1 was fully consumed before the __syncthreads that ended the previous iteration
    if (next < ntiles) next_tma = issue(next, s ^ 1);
    if (cur_tma) {
      mbar_wait(&bars[s], (phase_bits >> s) & 1u);
      phase_bits ^= 1u << s;
    } else {
      __syncthreads();  // plain loads of this stage were written by other warps (previous iteration / prologue)
    }

    const int64_t c0 = tile * CB;
    const int nc = (int)tmin<int64_t>(CB, a.ncol - c0);
    const R* sx1 = stage0 + (size_t)s * stage_elems;
    const R* sy1 = sx1 + n1p;
    const R* sx2 = sy1 + n1p * nf;
    // one warp per column (CB is a multiple of the warp count whenever a stage can hold that many columns); lanes walk
    // the targets 32 at a time. Everything per target is [32-bit shared address + byte offset]: per-column bases are
    // formed once, the loop carries three running pointers and no index arithmetic.
    const int p1 = (int)a.pitch1, p2 = (int)a.pitch2, P = a.P;
    constexpr unsigned B = sizeof(R);
    const unsigned last_off = (unsigned)(km1 - 1) * B;
    const bool want_map = MODE != LI_APPLY && a.indx != nullptr;
    for (int c = warp; c < nc; c += nwarp) {
      const unsigned x1a = smem_u32(sx1 + c * p1), y1a = smem_u32(sy1 + c * p1);
      unsigned x2a = smem_u32(sx2 + c * p2) + (unsigned)lane * B;
      const size_t gcol = (size_t)(c0 + c);
      R* gy = gy2 + gcol * p2 + lane;
      int* gi = want_map ? a.indx + gcol * km2 + lane : nullptr;
      const int* gin = MODE == LI_APPLY ? a.indx_in + gcol * km2 + lane : nullptr;
      for (int k2 = lane; k2 < km2; k2 += 32, x2a += 32 * B, gy += 32) {
        const R xt = lds_at(x2a, R());
        unsigned off;  // byte offset of the bracket k1 within the column
        if (MODE == LI_APPLY) off = (unsigned)gin[k2 - lane] * B;
        else off = bracket_off<R, LOG2P, sizeof(R)>(x1a, km1, P, xt);
        if (want_map) gi[k2 - lane] = (int)(off / B);
        if (MODE != LI_SETUP) {
          const unsigned offph = off == last_off ? off - B : off + B;  // k1ph, :119-131
          const R xa = lds_at(x1a + off, R()), xb = lds_at(x1a + offph, R());
          const R dxt = x_sub(xt, xa), dx = x_sub(xb, xa);
          unsigned yf = y1a;
          R* gyf = gy;
          for (int f = 0; f < nf; ++f) {
            const R ya = lds_at(yf + off, R()), yb = lds_at(yf + offph, R());
            R v = x_sub(yb, ya);  // :153
            v = x_mul(v, dxt);    // :154
            v = x_div(v, dx);     // :155
            v = x_add(v, ya);     // :156
            st_stream(gyf, v);
            yf += (unsigned)(n1p * B);
            gyf += a.y2_fstride;
          }
        }
      }
    }
    __syncthreads();  // everyone is done with stage s before it is refilled two iterations from now
    cur_tma = next_tma;
  }
}

// The same kernel with a lane's targets taken U at a time and no branch between them: U independent chains of probes,
// then U interpolations whose divisions go through the speculative fast path (common.cuh div_spec: equal to __ddiv_rn
// unless it says otherwise, in which case the batch is redone with the full-range routine). More registers (three
// CTAs per SM instead of four), so it pays where k_lin_interp is instruction-bound rather than HBM-bound: fewer source
// levels than targets (f64 72 -> 128: 79 -> 82% of the HBM peak) and f32 with equal counts (128 -> 128: 60 -> 67%). One
// target at a time stays the form elsewhere (f64 128 -> 128: 98%; batched 79-92%).
template <typename R, int MODE, int LOG2P, int U>
__global__ void __launch_bounds__(256, 3) k_lin_interp_batch(const LiArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t bars[2];

  const R* __restrict__ gx1 = static_cast<const R*>(a.x1);
  const R* __restrict__ gx2 = static_cast<const R*>(a.x2);
  const R* __restrict__ gy1 = static_cast<const R*>(a.y1);
  R* __restrict__ gy2 = static_cast<R*>(a.y2);
  const int km1 = a.km1, km2 = a.km2, CB = a.CB;
  const int nf = MODE == LI_SETUP ? 0 : a.nfields;
  const bool need_x2 = true;  // x2 drives the search (setup/fused) and enters the formula (apply/fused)
  // stage layout: x1[CB*pitch1] | y1[nf][CB*pitch1] | x2[CB*pitch2], each 16B-aligned
  const size_t n1 = (size_t)CB * a.pitch1, n2 = (size_t)CB * a.pitch2;
  const size_t n1p = (n1 * sizeof(R) + 15) / 16 * 16 / sizeof(R), n2p = (n2 * sizeof(R) + 15) / 16 * 16 / sizeof(R);
  const size_t stage_elems = n1p * (1 + nf) + n2p;
  R* stage0 = reinterpret_cast<R*>(smem_raw);

  const int64_t ntiles = (a.ncol + CB - 1) / CB;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;

  if (tid == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    fence_mbar_init();
  }
  __syncthreads();

  // issue the loads of tile `tile` into stage `s` (TMA when the tile is full and aligned)
  auto issue = [&](int64_t tile, int s) {
    const int64_t c0 = tile * CB;
    const bool tma = a.use_tma && c0 + CB <= a.ncol;
    if (tma && tid != 0) return true;  // one thread feeds the copy engine; nobody else spends an instruction here
    const int nc = (int)tmin<int64_t>(CB, a.ncol - c0);
    R* sx1 = stage0 + (size_t)s * stage_elems;
    R* sy1 = sx1 + n1p;
    R* sx2 = sy1 + n1p * nf;
    if (tma) {
      {
        fence_proxy_async();
        const unsigned b1 = (unsigned)(n1 * sizeof(R)), b2 = (unsigned)(n2 * sizeof(R));
        mbar_expect_tx(&bars[s], b1 * (1 + nf) + (need_x2 ? b2 : 0));
        tma_load_1d(sx1, gx1 + (size_t)c0 * a.pitch1, b1, &bars[s]);
        for (int f = 0; f < nf; ++f)
          tma_load_1d(sy1 + n1p * f, gy1 + (size_t)f * a.y1_fstride + (size_t)c0 * a.pitch1, b1, &bars[s]);
        if (need_x2) tma_load_1d(sx2, gx2 + (size_t)c0 * a.pitch2, b2, &bars[s]);
      }
    } else {
      // ragged or unaligned tile: plain coalesced loads (only real data, never the pack padding)
      for (int c = warp; c < nc; c += nwarp) {
        for (int k = lane; k < km1; k += 32) {
          sx1[(size_t)c * a.pitch1 + k] = gx1[(size_t)(c0 + c) * a.pitch1 + k];
          for (int f = 0; f < nf; ++f)
            sy1[n1p * f + (size_t)c * a.pitch1 + k] = gy1[(size_t)f * a.y1_fstride + (size_t)(c0 + c) * a.pitch1 + k];
        }
        for (int k = lane; k < km2; k += 32) sx2[(size_t)c * a.pitch2 + k] = gx2[(size_t)(c0 + c) * a.pitch2 + k];
      }
    }
    return tma;
  };

  unsigned phase_bits = 0;  // bit s = parity to wait for on stage s
  int64_t tile = blockIdx.x;
  bool cur_tma = false;
  if (tile < ntiles) cur_tma = issue(tile, 0);
  for (int it = 0; tile < ntiles; tile += gridDim.x, ++it) {
    const int s = it & 1;
    const int64_t next = tile + gridDim.x;
    bool next_tma = false;
    // stage s^1 was fully consumed before the __syncthreads that ended the previous iteration
    if (next < ntiles) next_tma = issue(next, s ^ 1);
    if (cur_tma) {
      mbar_wait(&bars[s], (phase_bits >> s) & 1u);
      phase_bits ^= 1u << s;
    } else {
      __syncthreads();  // plain loads of this stage were written by other warps (previous iteration / prologue)
    }

    const int64_t c0 = tile * CB;
    const int nc = (int)tmin<int64_t>(CB, a.ncol - c0);
    const R* sx1 = stage0 + (size_t)s * stage_elems;
    const R* sy1 = sx1 + n1p;
    const R* sx2 = sy1 + n1p * nf;
    // one warp per column (CB is a multiple of the warp count whenever a stage can hold that many columns); lanes walk
    // the targets 32 at a time. Everything per target is [32-bit shared address + byte offset]: per-column bases are
    // formed once, the loop carries three running pointers and no index arithmetic.
    const int p1 = (int)a.pitch1, p2 = (int)a.pitch2, P = a.P;
    constexpr unsigned B = sizeof(R);
    const unsigned last_off = (unsigned)(km1 - 1) * B;
    const bool want_map = MODE != LI_APPLY && a.indx != nullptr;
    for (int c = warp; c < nc; c += nwarp) {
      const unsigned x1a = smem_u32(sx1 + c * p1), y1a = smem_u32(sy1 + c * p1);
      unsigned x2a = smem_u32(sx2 + c * p2) + (unsigned)lane * B;
      const size_t gcol = (size_t)(c0 + c);
      R* gy = gy2 + gcol * p2 + lane;
      int* gi = want_map ? a.indx + gcol * km2 + lane : nullptr;
      const int* gin = MODE == LI_APPLY ? a.indx_in + gcol * km2 + lane : nullptr;
      for (int k2 = lane; k2 < km2; k2 += 32 * U, x2a += 32 * U * B, gy += 32 * U) {
        R xt[U];
        unsigned off[U];
        bool live[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          live[u] = k2 + 32 * u < km2;
          xt[u] = lds_at(live[u] ? x2a + 32 * u * B : x2a, R());
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          if (MODE == LI_APPLY) off[u] = (unsigned)gin[live[u] ? k2 - lane + 32 * u : k2 - lane] * B;
          else off[u] = bracket_off<R, LOG2P, sizeof(R)>(x1a, km1, P, xt[u]);  // byte offset of the bracket k1 within the column
          if (want_map && live[u]) gi[k2 - lane + 32 * u] = (int)(off[u] / B);
        }
        if (MODE != LI_SETUP) {
          unsigned offph[U];
          R dxt[U], dx[U], rc[U];
#pragma unroll
          for (int u = 0; u < U; ++u) {
            offph[u] = off[u] == last_off ? off[u] - B : off[u] + B;  // k1ph, :119-131
            const R xa = lds_at(x1a + off[u], R()), xb = lds_at(x1a + offph[u], R());
            dxt[u] = x_sub(xt[u], xa);
            dx[u] = x_sub(xb, xa);
            rc[u] = rcp_refined(dx[u]);
          }
          unsigned yf = y1a;
          R* gyf = gy;
          for (int f = 0; f < nf; ++f) {
            bool bad = false;
            R ya[U], num[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
              ya[u] = lds_at(yf + off[u], R());
              const R yb = lds_at(yf + offph[u], R());
              R v = x_sub(yb, ya[u]);                    // :153
              num[u] = x_mul(v, dxt[u]);                 // :154
              v = div_spec(num[u], dx[u], rc[u], bad);   // :155
              v = x_add(v, ya[u]);                       // :156
              if (live[u]) st_stream(gyf + 32 * u, v);
            }
            if (__builtin_expect(bad, 0)) {
#pragma unroll
              for (int u = 0; u < U; ++u) {
                const R v = x_add(x_div_full(num[u], dx[u]), ya[u]);
                if (live[u]) st_stream(gyf + 32 * u, v);
              }
            }
            yf += (unsigned)(n1p * B);
            gyf += a.y2_fstride;
          }
        }
      }
    }
    __syncthreads();  // everyone is done with stage s before it is refilled two iterations from now
    cur_tma = next_tma;
  }
}

template <typename R>
int lin_interp(int mode, const R* x1, const R* x2, const R* y1, R* y2, int* indx, const int* indx_in,
               int64_t ncol, int km1, int km2, int64_t pitch1, int64_t pitch2, int nfields, int64_t y1_fstride,
               int64_t y2_fstride, cudaStream_t stream) {
  if (ncol == 0 || km2 == 0) return EKAT_B200_OK;
  if (mode != LI_SETUP && nfields == 0) {
    if (mode == LI_APPLY || !indx) return EKAT_B200_OK;
    mode = LI_SETUP;
  }
  const int nf = mode == LI_SETUP ? 0 : nfields;
  const size_t col_bytes = ((size_t)pitch1 * (1 + nf) + pitch2) * sizeof(R);
  // columns per tile: a multiple of the 8 warps, stages of at most ~36 KB (2-4 CTAs of 2 stages each per SM) and at
  // most 24 columns (measured, f64 / f32: 72->128 best at 16 / 24 columns, 128->128 at 8 / 16)
  const Tuning tu = tuning();
  int CB = tu.cols_per_cta > 0 ? tu.cols_per_cta : (int)((36 * 1024) / col_bytes);
  if (tu.cols_per_cta <= 0 && CB > 24) CB = 24;
  if (CB > 64) CB = 64;
  if (tu.cols_per_cta <= 0) CB = CB >= 8 ? (CB & ~7) : 8;  // a multiple of the 8 warps: one column per warp per round
  if (CB > 1) CB &= ~1;  // even -> tile byte counts are multiples of 16 for f64
  if (sizeof(R) == 4 && CB > 3) CB &= ~3;
  if (CB < 1) CB = 1;
  size_t smem;
  for (;;) {
    const size_t n1p = ((size_t)CB * pitch1 * sizeof(R) + 15) / 16 * 16, n2p = ((size_t)CB * pitch2 * sizeof(R) + 15) / 16 * 16;
    smem = 2 * (n1p * (1 + nf) + n2p);
    if (smem <= kMaxSmemPerCta || CB == 1) break;
    CB /= 2;
  }
  if (smem > kMaxSmemPerCta) return EKAT_B200_ERR_TOO_LARGE;
  auto aligned16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  const bool sizes_ok = ((size_t)CB * pitch1 * sizeof(R)) % 16 == 0 && ((size_t)CB * pitch2 * sizeof(R)) % 16 == 0 &&
                        ((size_t)y1_fstride * sizeof(R)) % 16 == 0;
  const bool use_tma = tu.variant != kVarNoTmaStaging && sizes_ok && aligned16(x1) && aligned16(x2) && (nf == 0 || aligned16(y1));

  LiArgs a;
  a.x1 = x1; a.x2 = x2; a.y1 = y1; a.y2 = y2; a.indx = indx; a.indx_in = indx_in;
  a.ncol = ncol; a.pitch1 = pitch1; a.pitch2 = pitch2; a.y1_fstride = y1_fstride; a.y2_fstride = y2_fstride;
  a.km1 = km1; a.km2 = km2; a.nfields = nfields; a.CB = CB; a.use_tma = use_tma ? 1 : 0;
  int P = 1, log2p = 0;
  while (2 * P <= km1) { P *= 2; ++log2p; }
  a.P = P;

  const int64_t ntiles = (ncol + CB - 1) / CB;
  int per_sm = (int)(kMaxSmemPerCta / (smem + 1024));
  if (per_sm < 1) per_sm = 1;
  if (per_sm > 8) per_sm = 8;
  const unsigned grid = (unsigned)tmin<int64_t>(ntiles, (int64_t)kNumSms * per_sm);
  int rc;
  const char* name;
  if (getenv("EKAT_B200_DEBUG"))
    fprintf(stderr, "[ekat_b200] lin_interp mode=%d ncol=%lld km1=%d km2=%d p1=%lld p2=%lld nf=%d CB=%d smem=%zu grid=%u tma=%d\n",
            mode, (long long)ncol, km1, km2, (long long)pitch1, (long long)pitch2, nf, CB, smem, grid, (int)use_tma);
  // batched form for f32 and for f64 with fewer source levels than targets (see k_lin_interp_batch); tuning variant
  // 21 / 22 force the one-at-a-time / the batched form. Its persistent grid is what is actually resident: registers
  // bind before shared memory does, and a CTA that only starts when another has finished all its tiles would run alone.
  // (B200, 1.57 M columns, one-at-a-time vs batched, % of the HBM peak: f64 72->128 79 vs 82, 128->128 94-98 vs 79-92,
  //  128->72 73 vs 71; f32 128->128 60 vs 67, 72->128 53 vs 52, 128->72 58 vs 49)
  const bool batched = tu.variant == kVarLiBatched || (tu.variant != kVarLiUnbatched && (km1 < km2 || (sizeof(R) == 4 && km1 == km2)));
#define EKB_LI_LAUNCH_B(MODE_, L2_, U_)                                                                      \
  {                                                                                                          \
    auto k = k_lin_interp_batch<R, MODE_, L2_, U_>;                                                          \
    if ((rc = allow_smem(k, smem))) return rc;                                                               \
    int nb = 1;                                                                                              \
    if ((rc = check_cuda(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k, 256, smem)))) return rc;      \
    nb = nb < 1 ? 1 : nb > 8 ? 8 : nb;                                                                       \
    k<<<(unsigned)tmin<int64_t>(ntiles, (int64_t)kNumSms * nb), 256, smem, stream>>>(a);                     \
  }
#define EKB_LI_LAUNCH(MODE_, L2_)                      \
  if (batched) {                                       \
    if constexpr (sizeof(R) == 4) EKB_LI_LAUNCH_B(MODE_, L2_, 4) \
    else EKB_LI_LAUNCH_B(MODE_, L2_, 2)                \
  } else {                                             \
    auto k = k_lin_interp<R, MODE_, L2_>;              \
    if ((rc = allow_smem(k, smem))) return rc;         \
    k<<<grid, 256, smem, stream>>>(a);                 \
  }
#define EKB_LI_MODE(MODE_)                             \
  if (log2p == 6) EKB_LI_LAUNCH(MODE_, 6)              \
  else if (log2p == 7) EKB_LI_LAUNCH(MODE_, 7)         \
  else EKB_LI_LAUNCH(MODE_, 0)
  if (mode == LI_FUSED) {
    EKB_LI_MODE(LI_FUSED)
    name = sizeof(R) == 8 ? "lin_interp<f64,fused>" : "lin_interp<f32,fused>";
  } else if (mode == LI_SETUP) {
    EKB_LI_MODE(LI_SETUP)
    name = sizeof(R) == 8 ? "lin_interp<f64,setup>" : "lin_interp<f32,setup>";
  } else {
    EKB_LI_LAUNCH(LI_APPLY, 0)
    name = sizeof(R) == 8 ? "lin_interp<f64,apply>" : "lin_interp<f32,apply>";
  }
#undef EKB_LI_MODE
#undef EKB_LI_LAUNCH
#undef EKB_LI_LAUNCH_B
  note_launch(name);
  return check_cuda(cudaGetLastError());
}

template int lin_interp<double>(int, const double*, const double*, const double*, double*, int*, const int*, int64_t,
                                int, int, int64_t, int64_t, int, int64_t, int64_t, cudaStream_t);
template int lin_interp<float>(int, const float*, const float*, const float*, float*, int*, const int*, int64_t, int,
                               int, int64_t, int64_t, int, int64_t, int64_t, cudaStream_t);

}  // namespace ekb
