// tridiag_cr_blk.cu -- batched cyclic reduction, level-major operands, one matrix and one right-hand side per
// column, many columns of at most 128 levels (BASELINE config 3: cr over 345,600 columns x 128 levels; config 1's 72).
//
// Replaces the per-column loops around ekat::tridiag::cr(team, dl, d, du, X) (reference
// src/algorithm/ekat_tridiag.hpp:385-458). Every row update is the reference's, operation for operation and
// separately rounded (:404-419 down, :427-437 bottom, :446-452 up), so the results are bit-identical to the
// reference's CR at -ffp-contract=off. What is re-thought is who computes which row when:
//
//  * k_tridiag_cr_tma gives every level of the reduction its own pass over shared memory and its own CTA-wide
//    barrier: 13 barriers for 128 levels, one row update per thread between them, ~95 instructions per update of
//    which most are address arithmetic, predicate selects and the compiler's division; ncu: issue slots 49% busy
//    with the arithmetic chain of one update exposed between two barriers.
//  * Here a thread takes a BLOCK of four rows of the current level (stride S) and carries it through TWO levels
//    in registers: it loads the seven rows r0-3S .. r0+3S once, updates r0-2S, r0, r0+2S at stride S and r0 at
//    stride 2S. The update of r0-2S repeats what the thread of the block to the left computes (same operands,
//    same operations, same bits) -- that one redundant update buys a stage without any exchange between threads,
//    so 128 rows need three stages = three barriers on the way down instead of six, the three updates of a level
//    are independent chains in one thread, and the four divisors of a stage's first level are refined once for
//    the six quotients that use them.
//  * Only row r0 is written back during a stage (no other thread reads it in that stage: blocks read rows
//    r0 + kS, k = -3..3, k != 0 of their neighbours). The level-S state of row r0+2S, which the way up needs,
//    stays in the thread's registers until the same thread solves that row on the way up.
//  * Divisions are IEEE quotients computed as refined reciprocal + correction, speculatively (div_spec): a thread
//    whose operands leave the range where that is proven equal to the IEEE quotient (CR's off-diagonals shrink
//    geometrically, so deep levels can get there) repeats its block with the compiler's full-range division.
//  * The CTA is persistent (one per SM, 16 columns x 32 row blocks = 512 threads) with two tile buffers: the four
//    operand planes of the next tile arrive by TMA while this one is reduced, and the solution plane leaves by a
//    TMA tensor store that the next tile's reduction overlaps.
// An odd number of levels is handled by one plain level at stride 1 in front of the stages (and after them on the
// way up). HBM traffic is the algorithmic minimum (read dl, d, du, X once, write X once).
#include <cuda.h>

#include "tridiag_tma.cuh"

namespace ekb {

constexpr int kBlkTW = 16;        // columns per tile (one 128-byte row of f64)
constexpr int kBlkSlots = 32;     // row blocks per tile
constexpr int kBlkMaxStages = 3;  // nrow <= 128: at most six levels
constexpr int kBlkMaxRows = 128;

template <typename R>
struct CrRow {
  R l, d, u, x;
};

template <typename R, bool SLOW>
__device__ __forceinline__ R cr_div(R a, R b, R rb, bool& bad) {
  if (SLOW) return x_div(a, b);
  return div_spec(a, b, rb, bad);
}

// One row of a down-sweep level (:404-419). m / p: rows i-os / i+os in their state after the previous level; rm / rp:
// refined reciprocals of their diagonals (unused when SLOW). A missing neighbour is the row itself times a zero
// factor, exactly as the reference's clamped index does it -- including that d(i) then sees the NEW du(i) / dl(i).
template <typename R, bool SLOW>
__device__ __forceinline__ CrRow<R> cr_down(const CrRow<R>& me, const CrRow<R>& m, const CrRow<R>& p, bool im_ok, bool ip_ok,
                                            R rm, R rp, bool& bad) {
  bool b1 = false, b2 = false;
  R f1 = cr_div<R, SLOW>(-me.l, m.d, rm, b1);
  R f2 = cr_div<R, SLOW>(-me.u, p.d, rp, b2);
  f1 = im_ok ? f1 : (R)0;  // :408-409
  f2 = ip_ok ? f2 : (R)0;
  bad = bad || (im_ok && b1) || (ip_ok && b2);
  CrRow<R> o;
  o.l = x_mul(f1, im_ok ? m.l : me.l);  // :415
  o.u = x_mul(f2, ip_ok ? p.u : me.u);  // :416
  o.d = x_add(me.d, x_add(x_mul(f1, im_ok ? m.u : o.u), x_mul(f2, ip_ok ? p.l : o.l)));  // :417
  o.x = x_add(me.x, x_add(x_mul(f1, im_ok ? m.x : me.x), x_mul(f2, ip_ok ? p.x : me.x)));  // :418
  return o;
}

// One row of an up-sweep level (:446-452): xm / xp are the solved neighbours.
template <typename R, bool SLOW>
__device__ __forceinline__ R cr_up(const CrRow<R>& me, R xm, R xp, bool im_ok, bool ip_ok, R rd, bool& bad) {
  R f = 0;
  f = x_add(f, im_ok ? x_mul(me.l, xm) : (R)0);  // :450-451
  f = x_add(f, ip_ok ? x_mul(me.u, xp) : (R)0);
  return cr_div<R, SLOW>(x_sub(me.x, f), me.d, rd, bad);  // :452
}

template <typename R>
__device__ __forceinline__ CrRow<R> cr_ld(const R* pl, const R* pd, const R* pu, const R* px, int i) {
  CrRow<R> r;
  r.l = pl[i * kBlkTW];
  r.d = pd[i * kBlkTW];
  r.u = pu[i * kBlkTW];
  r.x = px[i * kBlkTW];
  return r;
}

// Two levels (strides S and 2S) of the block of rows r0, r0+S, r0+2S, r0+3S. out0: row r0 after both levels;
// out2: row r0+2S after the first (meaningful if it exists). `left`: r0 > 0, so rows r0-S, r0-2S, r0-3S exist.
template <typename R, bool SLOW>
__device__ __forceinline__ void cr_stage_down(const R* pl, const R* pd, const R* pu, const R* px, int r0, int S, int nrow,
                                              bool left, CrRow<R>& out0, CrRow<R>& out2, bool& bad) {
  const bool e1 = r0 + S < nrow, e2 = r0 + 2 * S < nrow, e3 = r0 + 3 * S < nrow;
  // rows that do not exist are read as row r0: finite values that only ever meet a zero factor or a select
  const CrRow<R> a0 = cr_ld(pl, pd, pu, px, r0);
  const CrRow<R> am1 = cr_ld(pl, pd, pu, px, left ? r0 - S : r0);
  const CrRow<R> am2 = cr_ld(pl, pd, pu, px, left ? r0 - 2 * S : r0);
  const CrRow<R> am3 = cr_ld(pl, pd, pu, px, left ? r0 - 3 * S : r0);
  const CrRow<R> a1 = cr_ld(pl, pd, pu, px, e1 ? r0 + S : r0);
  const CrRow<R> a2 = cr_ld(pl, pd, pu, px, e2 ? r0 + 2 * S : r0);
  const CrRow<R> a3 = cr_ld(pl, pd, pu, px, e3 ? r0 + 3 * S : r0);
  R rm3 = 0, rm1 = 0, r1 = 0, r3 = 0;
  if (!SLOW) {
    rm3 = rcp_refined(am3.d);
    rm1 = rcp_refined(am1.d);
    r1 = rcp_refined(a1.d);
    r3 = rcp_refined(a3.d);
  }
  bool bA = false, bB = false, bC = false, bD = false;
  const CrRow<R> A = cr_down<R, SLOW>(am2, am3, am1, true, true, rm3, rm1, bA);  // row r0-2S (the left block's update, repeated)
  const CrRow<R> B = cr_down<R, SLOW>(a0, am1, a1, left, e1, rm1, r1, bB);       // row r0
  const CrRow<R> C = cr_down<R, SLOW>(a2, a1, a3, true, e3, r1, r3, bC);         // row r0+2S
  R rA = 0, rC = 0;
  if (!SLOW) {
    rA = rcp_refined(A.d);
    rC = rcp_refined(C.d);
  }
  out0 = cr_down<R, SLOW>(B, A, C, left, e2, rA, rC, bD);  // row r0 at stride 2S
  out2 = C;
  bad = (left && bA) || bB || (e2 && bC) || bD;
}

// The same block on the way up: x(r0) and x(r0+4S) are solved; row r0+2S at stride 2S (state `sv` from the way
// down), then rows r0+S and r0+3S at stride S (state in shared memory since the previous stage).
template <typename R, bool SLOW>
__device__ __forceinline__ void cr_stage_up(const R* pl, const R* pd, const R* pu, const R* px, int r0, int S, int nrow,
                                            const CrRow<R>& sv, R& x1, R& x2, R& x3, bool& bad) {
  const bool e1 = r0 + S < nrow, e2 = r0 + 2 * S < nrow, e3 = r0 + 3 * S < nrow, e4 = r0 + 4 * S < nrow;
  const R x0 = px[r0 * kBlkTW], x4 = px[(e4 ? r0 + 4 * S : r0) * kBlkTW];
  const CrRow<R> a1 = cr_ld(pl, pd, pu, px, e1 ? r0 + S : r0);
  const CrRow<R> a3 = cr_ld(pl, pd, pu, px, e3 ? r0 + 3 * S : r0);
  R rs = 0, r1 = 0, r3 = 0;
  if (!SLOW) {
    rs = rcp_refined(sv.d);
    r1 = rcp_refined(a1.d);
    r3 = rcp_refined(a3.d);
  }
  bool b2 = false, b1 = false, b3 = false;
  x2 = cr_up<R, SLOW>(sv, x0, x4, true, e4, rs, b2);
  x1 = cr_up<R, SLOW>(a1, x0, x2, true, e2, r1, b1);
  x3 = cr_up<R, SLOW>(a3, x2, x4, true, e4, r3, b3);
  bad = (e2 && b2) || (e1 && b1) || (e3 && b3);
}

template <typename R>
__global__ void __launch_bounds__(kBlkTW* kBlkSlots, 1)
k_tridiag_cr_blk(const __grid_constant__ CUtensorMap mdl, const __grid_constant__ CUtensorMap md,
                 const __grid_constant__ CUtensorMap mdu, const __grid_constant__ CUtensorMap mx, int nrow, int ntiles, int nst,
                 int tail) {
  constexpr int TW = kBlkTW, L = TW;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw);  // [2] the four planes of a buffer have landed
  const size_t plane = (((size_t)nrow * TW * sizeof(R) + 127) & ~(size_t)127) / sizeof(R);  // TMA boxes land 128-byte aligned
  R* buf0 = reinterpret_cast<R*>(smem_raw + 128);
  const int tid = threadIdx.x;
  const int c = tid % TW, w = tid / TW;  // column of the tile, row block
  const unsigned tile_bytes = (unsigned)(4 * (size_t)nrow * TW * sizeof(R));
  auto load_tile = [&](int t, int b) {  // thread 0
    R* p = buf0 + (size_t)b * 4 * plane;
    mbar_expect_tx(&full[b], tile_bytes);
    tma_load_2d(p, &mdl, t * TW, 0, &full[b]);
    tma_load_2d(p + plane, &md, t * TW, 0, &full[b]);
    tma_load_2d(p + 2 * plane, &mdu, t * TW, 0, &full[b]);
    tma_load_2d(p + 3 * plane, &mx, t * TW, 0, &full[b]);
  };
  if (tid == 0) {
    mbar_init(&full[0], 1);
    mbar_init(&full[1], 1);
    fence_mbar_init();
    load_tile(blockIdx.x, 0);
  }
  __syncthreads();

  int it = 0;
  for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++it) {
    const int b = it & 1;
    if (tid == 0 && t + (int)gridDim.x < ntiles) {
      tma_store_wait_read<0>();  // the store of the previous tile has read the other buffer
      load_tile(t + gridDim.x, b ^ 1);
    }
    R* pl = buf0 + (size_t)b * 4 * plane + c;
    R* pd = pl + plane;
    R* pu = pd + plane;
    R* px = pu + plane;
    mbar_wait(&full[b], (unsigned)(it >> 1) & 1u);

    int S = 1;
    if (tail) {  // an odd number of levels: the first one on its own (:401-423 at os = 1)
      for (int i = 2 * w; i < nrow; i += 2 * kBlkSlots) {
        const bool im_ok = i >= 1, ip_ok = i + 1 < nrow;
        const CrRow<R> me = cr_ld(pl, pd, pu, px, i), m = cr_ld(pl, pd, pu, px, im_ok ? i - 1 : i),
                       p = cr_ld(pl, pd, pu, px, ip_ok ? i + 1 : i);
        bool bad = false;
        const CrRow<R> o = cr_down<R, true>(me, m, p, im_ok, ip_ok, (R)0, (R)0, bad);
        pl[i * L] = o.l;
        pd[i * L] = o.d;
        pu[i * L] = o.u;
        px[i * L] = o.x;
      }
      __syncthreads();
      S = 2;
    }
    CrRow<R> sv[kBlkMaxStages];
#pragma unroll
    for (int s = 0; s < kBlkMaxStages; ++s) {
      if (s < nst) {
        const int r0 = 4 * w * S;
        if (r0 < nrow) {
          CrRow<R> o0, o2;
          bool bad = false;
          cr_stage_down<R, false>(pl, pd, pu, px, r0, S, nrow, w > 0, o0, o2, bad);
          if (bad) cr_stage_down<R, true>(pl, pd, pu, px, r0, S, nrow, w > 0, o0, o2, bad);
          pl[r0 * L] = o0.l;
          pd[r0 * L] = o0.d;
          pu[r0 * L] = o0.u;
          px[r0 * L] = o0.x;
          sv[s] = o2;
        }
        __syncthreads();
        S *= 4;
      }
    }
    if (w == 0) {  // bottom 1x1 or 2x2 solve :427-437 (S is the first stride that is not a level)
      if (S >= nrow) {
        px[0] = x_div(px[0], pd[0]);
      } else {
        const R d0 = pd[0], dos = pd[S * L], du0 = pu[0], dlos = pl[S * L];
        const R det = x_sub(x_mul(d0, dos), x_mul(du0, dlos));
        const R x0 = px[0], x1 = px[S * L];
        px[0] = x_div(x_sub(x_mul(dos, x0), x_mul(du0, x1)), det);
        px[S * L] = x_div(x_sub(x_mul(d0, x1), x_mul(dlos, x0)), det);
      }
    }
    __syncthreads();  // :438
#pragma unroll
    for (int s = kBlkMaxStages - 1; s >= 0; --s) {
      if (s < nst) {
        S /= 4;
        const int r0 = 4 * w * S;
        if (r0 < nrow) {
          R x1, x2, x3;
          bool bad = false;
          cr_stage_up<R, false>(pl, pd, pu, px, r0, S, nrow, sv[s], x1, x2, x3, bad);
          if (bad) cr_stage_up<R, true>(pl, pd, pu, px, r0, S, nrow, sv[s], x1, x2, x3, bad);
          if (r0 + S < nrow) px[(r0 + S) * L] = x1;
          if (r0 + 2 * S < nrow) px[(r0 + 2 * S) * L] = x2;
          if (r0 + 3 * S < nrow) px[(r0 + 3 * S) * L] = x3;
        }
        __syncthreads();
      }
    }
    if (tail) {  // the last level on the way up: the odd rows (:446-452 at os = 1)
      for (int i = 2 * w + 1; i < nrow; i += 2 * kBlkSlots) {
        const bool ip_ok = i + 1 < nrow;
        const CrRow<R> me = cr_ld(pl, pd, pu, px, i);
        bool bad = false;
        px[i * L] = cr_up<R, true>(me, px[(i - 1) * L], px[(ip_ok ? i + 1 : i) * L], true, ip_ok, (R)0, bad);
      }
    }
    fence_proxy_async();  // this thread's writes of the solution plane -> visible to the tensor store
    __syncthreads();
    if (tid == 0) {
      tma_store_2d(&mx, t * TW, 0, px - c);
      tma_store_commit();
    }
  }
  if (tid == 0) tma_store_wait_read<0>();
}

// Returns EKAT_B200_ERR_UNSUPPORTED when the operands do not meet TMA's requirements or the shape is not this
// kernel's (more than 128 levels, few columns); the caller then tries k_tridiag_cr_tma and the generic tile kernel.
template <typename R>
int tridiag_cr_blk(R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int64_t a_stride, int64_t x_stride, cudaStream_t stream) {
  auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  if (!al(dl) || !al(d) || !al(du) || !al(X) || (a_stride * sizeof(R)) % 16 || (x_stride * sizeof(R)) % 16)
    return EKAT_B200_ERR_UNSUPPORTED;
  if (nlev > kBlkMaxRows) return EKAT_B200_ERR_UNSUPPORTED;
  if (ncol >= (int64_t)1 << 31 || ncol < 2 * kNumSms * kBlkTW) return EKAT_B200_ERR_UNSUPPORTED;
  int nl = 0;
  for (int os = 1; 2 * os < nlev; os *= 2) ++nl;
  const int nst = nl / 2, tail = nl & 1;
  const size_t plane = ((size_t)nlev * kBlkTW * sizeof(R) + 127) & ~(size_t)127;
  const size_t smem = 128 + 2 * 4 * plane;
  if (smem > kMaxSmemPerCta) return EKAT_B200_ERR_UNSUPPORTED;
  CUtensorMap mdl, md, mdu, mx;
  if (!make_tensor_map_2d<R>(&mdl, dl, ncol, nlev, a_stride, kBlkTW, nlev, 0) || !make_tensor_map_2d<R>(&md, d, ncol, nlev, a_stride, kBlkTW, nlev, 0) ||
      !make_tensor_map_2d<R>(&mdu, du, ncol, nlev, a_stride, kBlkTW, nlev, 0) || !make_tensor_map_2d<R>(&mx, X, ncol, nlev, x_stride, kBlkTW, nlev, 0))
    return EKAT_B200_ERR_UNSUPPORTED;
  auto k = k_tridiag_cr_blk<R>;
  int rc = allow_smem(k, smem);
  if (rc) return rc;
  const int64_t ntiles = (ncol + kBlkTW - 1) / kBlkTW;
  const unsigned grid = (unsigned)tmin<int64_t>(ntiles, kNumSms);
  k<<<grid, kBlkTW * kBlkSlots, smem, stream>>>(mdl, md, mdu, mx, nlev, (int)ntiles, nst, tail);
  note_launch(sizeof(R) == 8 ? "tridiag_cr_blk<f64>" : "tridiag_cr_blk<f32>");
  return check_cuda(cudaGetLastError());
}

template int tridiag_cr_blk<double>(double*, double*, double*, double*, int64_t, int, int64_t, int64_t, cudaStream_t);
template int tridiag_cr_blk<float>(float*, float*, float*, float*, int64_t, int, int64_t, int64_t, cudaStream_t);

}  // namespace ekb
