// lin_interp_lm.cu -- fused bracket search + linear interpolation on LEVEL-MAJOR operands
//     x1 [km1][pitch], x2 [km2][pitch], y1 [nfields][km1][pitch], y2 [nfields][km2][pitch]      (columns fastest)
// the layout the tridiagonal kernels are native in, so that "diffuse then remap" (BASELINE config 4) needs no
// transpose between the two. Same arithmetic as lin_interp.cu: the reference's predicate `v >= *it`
// (src/kokkos/ekat_upper_bound.hpp:37-48), "if (idx > 0) --idx" (src/algorithm/ekat_lin_interp_impl.hpp:179-181) and
// the four statements of :153-156 in order, unfused -> bit-identical to ekat::LinInterp for any pack size.
//
// A CTA owns a tile of 32 columns: lane = column, the 8 warps split the target levels. The tile's source coordinates
// and one source field at a time (double-buffered) arrive by 2D TMA tensor copies as [level][32 columns], so a lane
// only ever touches its own shared-memory bank (every gather is conflict-free whatever the brackets are), target
// coordinates are read and results written as coalesced 256-byte rows, and the bracket of a target is searched once
// and reused for every field (the reference's setup-once / interpolate-many contract, ekat_lin_interp.hpp:13-16).
// y2 may alias y1 when km2 == km1: a field's tile is complete in shared memory before its targets are written.
#include <cuda.h>

#include "lin_interp_search.cuh"
#include "tridiag_tma.cuh"

namespace ekb {

constexpr int kLmWarps = 8;

// TW = tile width in columns (16 or 32): lane = (target parity, column) for TW = 16. NSY field stages in flight.
// NTP > 0: per-target state in registers across the fields, NTP targets per thread (km2 <= NTP * kLmWarps * 32 / TW);
// NTP == 0: the index map in shared memory.
template <typename R, int LOG2P, int TW, int NTP>
__global__ void __launch_bounds__(32 * kLmWarps, NTP == 0 ? 4 : NTP <= 8 ? 3 : 2)
k_lin_interp_lm(const __grid_constant__ CUtensorMap mx1, const __grid_constant__ CUtensorMap my1, const R* __restrict__ x2,
                R* __restrict__ y2, int64_t ncol, int km1, int km2, int64_t pitch, int nf, int P, int NSY) {
  constexpr unsigned B = sizeof(R), ROW = TW * B;
  constexpr int TPS = 32 / TW;  // targets per warp step
  constexpr bool REG = NTP > 0;
  constexpr int NT = REG ? NTP : 1;  // REG: targets per thread
  extern __shared__ __align__(128) unsigned char smem_raw[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw);  // [0] x1 landed, [1 + s] field stage s landed
  const size_t tile = (((size_t)km1 * ROW + 127) & ~(size_t)127) / B;  // elements per staged tile (TMA wants 128-byte aligned boxes)
  R* sx1 = reinterpret_cast<R*>(smem_raw + 128);           // [km1][TW]
  R* sy = sx1 + tile;                                      // [NSY][km1][TW]
  unsigned short* smap = reinterpret_cast<unsigned short*>(sy + (size_t)NSY * tile);  // [km2][TW] brackets

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int lc = lane % TW, sub = lane / TW;
  const int c0 = blockIdx.x * TW;
  const int64_t col = (int64_t)c0 + lc;
  const bool live = col < ncol;
  const int64_t gcol = live ? col : ncol - 1;
  const unsigned tile_bytes = (unsigned)km1 * ROW;

  auto load_tile = [&](R* dst, const CUtensorMap* m, int row0, uint64_t* bar) {  // one box: km1 (<= 256) rows of TW columns
    mbar_expect_tx(bar, tile_bytes);
    tma_load_2d(dst, m, c0, row0, bar);
  };
  if (tid == 0) {
    for (int s = 0; s <= NSY; ++s) mbar_init(&bars[s], 1);
    fence_mbar_init();
    load_tile(sx1, &mx1, 0, &bars[0]);
    for (int f = 0; f < NSY && f < nf; ++f) load_tile(sy + (size_t)f * tile, &my1, f * km1, &bars[1 + f]);
  }
  __syncthreads();

  const unsigned x1a = smem_u32(sx1) + (unsigned)lc * B;
  const unsigned last_off = (unsigned)(km1 - 1) * ROW;
  mbar_wait(&bars[0], 0u);
  if (REG) {
    // ---- km2 <= NT * 8 * TPS targets: a thread owns at most NT (target, column) pairs for EVERY field, so everything
    // that depends on the coordinates only stays in registers across the field loop: the bracket, x2 - x1[k1], the
    // interval x1[k1ph] - x1[k1] and the refined reciprocal of the interval -- the part of the IEEE division that does
    // not depend on the numerator (common.cuh: div_given_rcp is bit-identical to __ddiv_rn). A field then costs two
    // gathers, a subtraction, a product, the quotient's three fused steps and an addition per target.
    // No branch on j < km2 or on the division's rare operands inside the loops: a thread's NT targets are independent
    // chains (shared-memory probes, then fp64 latency) and must be free to interleave. Surplus slots redo the last
    // target and are never stored; div_spec records operands its fast path is not proven for and the field is then
    // redone with the full-range division (common.cuh).
    unsigned off[NT];
    R dxt[NT], dx[NT], rc[NT];
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      const int j = tmin(warp * TPS + sub + t * (kLmWarps * TPS), km2 - 1);
      const R xt = __ldg(x2 + (size_t)j * pitch + gcol);
      const unsigned o = bracket_off<R, LOG2P, ROW>(x1a, km1, P, xt);
      const unsigned oph = o == last_off ? o - ROW : o + ROW;  // k1ph, :119-131
      const R xa = lds_at(x1a + o, R()), xb = lds_at(x1a + oph, R());
      off[t] = o;
      dxt[t] = x_sub(xt, xa);
      dx[t] = x_sub(xb, xa);
      rc[t] = rcp_refined(dx[t]);
      pin(dxt[t]);  // left to itself the compiler recomputes these per field instead of keeping them
      pin(dx[t]);
      pin(rc[t]);
    }
    const int j0 = warp * TPS + sub;
    const size_t tstride = (size_t)(kLmWarps * TPS) * pitch;
    R* gy = y2 + (size_t)j0 * pitch + col;
    int s = 0;
    unsigned ph = 0;
    for (int f = 0; f < nf; ++f, gy += (size_t)km2 * pitch) {
      mbar_wait(&bars[1 + s], ph);
      const unsigned y1a = smem_u32(sy + (size_t)s * tile) + (unsigned)lc * B;
      bool bad = false;
#pragma unroll
      for (int t = 0; t < NT; ++t) {
        const unsigned o = off[t];
        const unsigned oph = o == last_off ? o - ROW : o + ROW;
        const R ya = lds_at(y1a + o, R()), yb = lds_at(y1a + oph, R());
        R v = x_sub(yb, ya);                   // :153
        v = x_mul(v, dxt[t]);                  // :154
        v = div_spec(v, dx[t], rc[t], bad);    // :155
        v = x_add(v, ya);                      // :156
        if (live && j0 + t * (kLmWarps * TPS) < km2) st_stream(gy + t * tstride, v);
      }
      if (__builtin_expect(bad, 0)) {
#pragma unroll
        for (int t = 0; t < NT; ++t) {
          const unsigned o = off[t];
          const unsigned oph = o == last_off ? o - ROW : o + ROW;
          const R ya = lds_at(y1a + o, R()), yb = lds_at(y1a + oph, R());
          R v = x_sub(yb, ya);
          v = x_mul(v, dxt[t]);
          v = x_div_full(v, dx[t]);
          v = x_add(v, ya);
          if (live && j0 + t * (kLmWarps * TPS) < km2) st_stream(gy + t * tstride, v);
        }
      }
      __syncthreads();  // every warp is done with stage s
      if (tid == 0 && f + NSY < nf) load_tile(sy + (size_t)s * tile, &my1, (f + NSY) * km1, &bars[1 + s]);
      if (++s == NSY) {
        s = 0;
        ph ^= 1u;
      }
    }
    return;
  }

  // ---- setup: one bracket per (target level, column); each thread later reads back only what it wrote itself
  for (int j = warp * TPS + sub; j < km2; j += kLmWarps * TPS) {
    const R xt = __ldg(x2 + (size_t)j * pitch + gcol);
    const unsigned off = bracket_off<R, LOG2P, ROW>(x1a, km1, P, xt);
    smap[j * TW + lc] = (unsigned short)(off / ROW);
  }

  // ---- interpolation, field by field
  int s = 0;
  unsigned ph = 0;
  for (int f = 0; f < nf; ++f) {
    mbar_wait(&bars[1 + s], ph);
    const unsigned y1a = smem_u32(sy + (size_t)s * tile) + (unsigned)lc * B;
    R* gy = y2 + ((size_t)f * km2) * pitch + col;
    for (int j = warp * TPS + sub; j < km2; j += kLmWarps * TPS) {
      const R xt = __ldg(x2 + (size_t)j * pitch + gcol);
      const unsigned off = (unsigned)smap[j * TW + lc] * ROW;
      const unsigned offph = off == last_off ? off - ROW : off + ROW;  // k1ph, :119-131
      const R xa = lds_at(x1a + off, R()), xb = lds_at(x1a + offph, R());
      const R ya = lds_at(y1a + off, R()), yb = lds_at(y1a + offph, R());
      const R dxt = x_sub(xt, xa), dx = x_sub(xb, xa);
      R v = x_sub(yb, ya);  // :153
      v = x_mul(v, dxt);    // :154
      v = x_div(v, dx);     // :155
      v = x_add(v, ya);     // :156
      if (live) st_stream(gy + (size_t)j * pitch, v);
    }
    __syncthreads();  // every warp is done with stage s
    if (tid == 0 && f + NSY < nf) load_tile(sy + (size_t)s * tile, &my1, (f + NSY) * km1, &bars[1 + s]);
    if (++s == NSY) {
      s = 0;
      ph ^= 1u;
    }
  }
}

// Returns EKAT_B200_ERR_UNSUPPORTED when the operands do not meet TMA's requirements or a tile does not fit on chip.
template <typename R>
int lin_interp_lm(const R* x1, const R* x2, const R* y1, R* y2, int64_t ncol, int km1, int km2, int64_t pitch, int nfields,
                  cudaStream_t stream) {
  if (ncol == 0 || km2 == 0 || nfields == 0) return EKAT_B200_OK;
  auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  if (!al(x1) || !al(y1) || (pitch * sizeof(R)) % 16) return EKAT_B200_ERR_UNSUPPORTED;
  if (ncol >= (int64_t)1 << 31 || (int64_t)nfields * km1 >= (int64_t)1 << 31 || km1 > 256) return EKAT_B200_ERR_UNSUPPORTED;  // one TMA box per tile
  // tile width and field stages: 16-column tiles (128-byte rows for f64), two field stages -> 52 KB, four CTAs per SM;
  // tuning: cols_per_cta = 16 | 32, prefetch = stages
  const Tuning tu = tuning();
  // 128-byte rows: 16 columns of f64, 32 of f32 (f32, 128 -> 64, 10 fields: 93% of the HBM peak against 48% with 16)
  // -- but 128 targets of 32 columns are 16 per thread, 128 registers and two CTAs per SM, which only pays when there
  // are fields to spread the search over (f32 128 -> 128: 1 field 35% against 47%, 10 fields 65% against 49%)
  const int tw = tu.cols_per_cta == 32 || tu.cols_per_cta == 16 ? tu.cols_per_cta
                 : (sizeof(R) == 4 && (km2 <= 8 * kLmWarps || nfields >= 4)) ? 32 : 16;
  int nsy = tu.prefetch >= 1 && tu.prefetch <= 8 ? tu.prefetch : 2;  // measured: residency (4 CTAs per SM) beats deeper staging
  if (nsy > nfields) nsy = nfields;
  auto smem_for = [&](int st) {
    return 128 + (size_t)(1 + st) * (((size_t)km1 * tw * sizeof(R) + 127) & ~(size_t)127) + (size_t)km2 * tw * sizeof(unsigned short);
  };
  while (nsy > 1 && smem_for(nsy) > kMaxSmemPerCta / 2) --nsy;
  const size_t smem = smem_for(nsy);
  if (smem > kMaxSmemPerCta) return EKAT_B200_ERR_UNSUPPORTED;
  CUtensorMap mx1, my1;
  const int boxr = km1;
  if (!make_tensor_map_2d<R>(&mx1, x1, ncol, km1, pitch, tw, boxr, 0) ||
      !make_tensor_map_2d<R>(&my1, y1, ncol, (int64_t)nfields * km1, pitch, tw, boxr, 0))
    return EKAT_B200_ERR_UNSUPPORTED;
  int P = 1, log2p = 0;
  while (2 * P <= km1) {
    P *= 2;
    ++log2p;
  }
  const unsigned grid = (unsigned)((ncol + tw - 1) / tw);
  // per-target state in registers whenever a thread's targets fit (km2 <= 8 per thread): f64 128 -> 128, 1 / 2 / 4 / 10
  // fields: 69 / 76 / 81 / 82% of the HBM peak against 53 / 57 / 63 / 76% for the shared-memory map with its rolled,
  // branching loop. tuning variant 1 forces the shared-memory map.
  const int per_step = kLmWarps * (32 / tw);
  const int ntp = tu.variant == kVarNoTmaStaging ? 0 : km2 <= 8 * per_step ? 8 : (sizeof(R) == 4 && km2 <= 16 * per_step) ? 16 : 0;
  int rc;
#define EKB_LM_LAUNCH3(L2_, TW_, NTP_)                                                                 \
  {                                                                                                    \
    auto k = k_lin_interp_lm<R, L2_, TW_, NTP_>;                                                       \
    if ((rc = allow_smem(k, smem))) return rc;                                                         \
    k<<<grid, 32 * kLmWarps, smem, stream>>>(mx1, my1, x2, y2, ncol, km1, km2, pitch, nfields, P, nsy); \
  }
#define EKB_LM_LAUNCH2(L2_, TW_)                                 \
  {                                                              \
    if (ntp == 8) EKB_LM_LAUNCH3(L2_, TW_, 8)                    \
    else if (ntp == 16) {                                        \
      if constexpr (sizeof(R) == 4) EKB_LM_LAUNCH3(L2_, TW_, 16) \
    } else EKB_LM_LAUNCH3(L2_, TW_, 0)                           \
  }
#define EKB_LM_LAUNCH(L2_)            \
  if (tw == 32) EKB_LM_LAUNCH2(L2_, 32) \
  else EKB_LM_LAUNCH2(L2_, 16)
  if (log2p == 6) EKB_LM_LAUNCH(6)
  else if (log2p == 7) EKB_LM_LAUNCH(7)
  else EKB_LM_LAUNCH(0)
#undef EKB_LM_LAUNCH
#undef EKB_LM_LAUNCH2
#undef EKB_LM_LAUNCH3
  note_launch(sizeof(R) == 8 ? "lin_interp_lm<f64>" : "lin_interp_lm<f32>");
  return check_cuda(cudaGetLastError());
}

template int lin_interp_lm<double>(const double*, const double*, const double*, double*, int64_t, int, int, int64_t, int, cudaStream_t);
template int lin_interp_lm<float>(const float*, const float*, const float*, float*, int64_t, int, int, int64_t, int, cudaStream_t);

}  // namespace ekb
