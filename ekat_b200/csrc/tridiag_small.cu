// tridiag_small.cu -- Thomas for short columns / small batches: the whole tile on chip, solved from both ends.
//
// Replaces the per-column loop around ekat::tridiag::thomas (reference src/algorithm/ekat_tridiag.hpp:216-227,
// :323-381) where the batch is latency-bound rather than bandwidth-bound: BASELINE config 1 (21,600 columns x 72
// levels: 675 tiles of 32 columns for 148 SMs, 62 MB, L2-resident in a time loop). There a launch costs a few
// dependent round trips, not bytes. Measured on B200 with time stamps inside the kernel (tools/lat_bench.cu,
// profiles/r02_lat_bench.txt): on the pipelined kernels (tridiag_tma.cu, tridiag_tma_shared.cu) a single tile takes
// ~4.2 us -- a ring hand-off (TMA round trip, mbarrier, proxy fence) per chunk and sweep -- and the batch 13-14 us
// (two waves of tiles), against 9.5 us at the HBM roofline. So here:
//  * all four operand planes of a tile are requested up front, in a few large TMA boxes (GL levels each, one
//    mbarrier per box row group): one memory round trip per tile, ~0.5 us of issue, no ring;
//  * the column is eliminated FROM BOTH ENDS at once ("burn at both ends"): warp 0 runs the forward recurrence down
//    from level 0, warp 1 the mirrored recurrence up from level n-1 (dl and du swap roles), each over half of the
//    levels; they meet in the middle, exchange two numbers per column through shared memory (the run's only
//    hand-off: one named barrier), solve the 2x2 interface system and back-substitute outwards, again concurrently.
//    The dependent chain and the instruction count per warp are halved;
//  * the recurrence is the continuant ("ratio") form of tridiag_tma.cu (no division on the dependent chain,
//    renormalised every RN levels) with the divisions LEFT OUT of the elimination: it stores the leading minors q[i]
//    (over d) and the scaled right-hand side z[i] (over x) -- five fp64 operations per level. The substitution, whose
//    own chain is one FMA per level, forms 1/q[i], c'[i] and y[i] = z[i]/q[i] on the side;
//  * chunks of 8 levels are staged in registers and the next chunk's shared-memory loads are issued before the
//    current chunk's stores, so no load queues behind a store to the same plane (145 cycles per level otherwise);
//  * the solution is stored straight from registers with coalesced streaming stores -- no tensor store, no proxy
//    fence, no bulk-group wait; full tiles store without a branch per level.
// Shared memory: 4 planes x nlev x 32 columns (72 levels f64: 74 KB -> 3 tiles per SM; f32: 6). Used for batches of
// at most a few waves of tiles when >= 2 tiles fit an SM; everything else goes to the pipelined kernels, which need
// less shared memory per column and stream better. One matrix per right-hand side (nmat matrices stacked in the row
// dimension, blockIdx.y) or one of each. Agreement with the reference: tolerance level (1e-13 / 1e-5), as thomas.
#include <cuda.h>

#include "tridiag_tma.cuh"

namespace ekb {

constexpr int kSmallGroups = 3;  // TMA boxes per plane

// One half of a column tile. UP = false: the top half, levels ascending from 0 (the classical forward elimination);
// UP = true: the bottom half, levels descending from nlev-1 (the same recurrence on the reversed system: du plays
// dl's part and vice versa). Chunks [kfirst, kfirst + nk) of 8 levels in processing order; within a chunk, slot
// s = 0..7 is level k*8 + s (top) or k*8 + 7 - s (bottom), so both halves run identical code on slot-ordered registers.
// MODE 0: the matrix is given (dl, d, du planes). MODE 1: the matrix is the implicit-diffusion stencil built on the
// fly from the exchange coefficients w (in the sdu plane, see k_tridiag_small): dl[i] = -w[i-1], du[i] = -w[i],
// d[i] = 1 + w[i] + w[i-1] -- sdl is unused and q overwrites the coordinate plane.
template <typename R, int RN, bool UP, int MODE>
__device__ __forceinline__ void small_half(R* __restrict__ sdl, R* __restrict__ sd, R* __restrict__ sdu, R* __restrict__ sx,
                                           uint64_t* bar, R* xch, int lane, int nlev, int nch, int gch, int kt, R* __restrict__ gx,
                                           int64_t x_stride, bool full_tile, bool live) {
  constexpr int LB = kLB;
  const int nk = UP ? nch - kt : kt;              // chunks of this half
  const bool other = UP ? true : kt < nch;        // does the other half exist?
  auto chunk_of = [&](int j) { return UP ? nch - 1 - j : j; };  // j-th chunk in processing order
  auto lev_of = [&](int k, int s) { return k * LB + (UP ? LB - 1 - s : s); };
  // slot-ordered staging of (a, b, c, x): a couples to the level processed before, c to the one processed next.
  // Entries the reference never uses (dl(0), du(n-1), ekat_tridiag.hpp:24-27) and levels past the end of a partial
  // top chunk are neutralised here (identity rows), so the sweeps need no special cases.
  auto stage = [&](int k, R (&ra)[LB], R (&rb)[LB], R (&rc)[LB], R (&rx)[LB]) {
#pragma unroll
    for (int s = 0; s < LB; ++s) {
      const int i = lev_of(k, s);
      const size_t o = (size_t)i * 32 + lane;
      if (MODE == 1) {
        const R wi = sdu[o], wm = i > 0 ? sdu[o - 32] : (R)0;  // w[-1] = w[n-1] = 0: the unused ends vanish by themselves
        ra[s] = UP ? -wi : -wm;
        rc[s] = UP ? -wm : -wi;
        rb[s] = x_add(x_add((R)1, wi), wm);
      } else {
        ra[s] = UP ? sdu[o] : sdl[o];
        rb[s] = sd[o];
        rc[s] = UP ? sdl[o] : sdu[o];
      }
      rx[s] = sx[o];
      if (MODE == 0 && i == 0) (UP ? rc[s] : ra[s]) = (R)0;
      if (MODE == 0 && i == nlev - 1) (UP ? ra[s] : rc[s]) = (R)0;
      if (i >= nlev) {
        ra[s] = (R)0;
        rb[s] = (R)1;
        rc[s] = (R)0;
        rx[s] = (R)0;
      }
    }
  };
  auto wait_group = [&](int k) { mbar_wait(&bar[k / gch], 0u); };

  // ---------------------------------------------------------------- elimination: sd <- q, sx <- z (no division)
  R q1 = (R)1, q2 = (R)0, z1 = (R)0, cp = (R)0;
  R a[LB], b[LB], c[LB], x[LB];
  if (nk > 0) {
    wait_group(chunk_of(0));
    stage(chunk_of(0), a, b, c, x);
  }
  R qlast = (R)1, zlast = (R)0, qprev_last = (R)1, clast = (R)0;  // last processed level, for the interface
  for (int j = 0; j < nk; ++j) {
    const int k = chunk_of(j);
    R na[LB], nb[LB], nc[LB], nx[LB];
    if (j + 1 < nk) {
      const int kn = chunk_of(j + 1);
      if (kn / gch != k / gch) wait_group(kn);
      stage(kn, na, nb, nc, nx);
    }
    R o0[LB], o1[LB];
#pragma unroll
    for (int s = 0; s < LB; ++s) {
      R q, z;
      if (s % RN == 0) {  // state was renormalised: q[prev] == 1, q2 = q[prev-1]/q[prev]
        q = f_fma(-(a[s] * cp), q2, b[s]);
        z = f_fma(-a[s], z1, x[s]);
      } else {
        q = f_fma(b[s], q1, -((a[s] * cp) * q2));
        z = f_fma(-a[s], z1, x[s] * q1);
      }
      o0[s] = q;
      o1[s] = z;
      if (s == LB - 1) {
        qlast = q;
        zlast = z;
        qprev_last = (s % RN == 0) ? (R)1 : q1;
        clast = c[s];
      }
      if ((s + 1) % RN == 0) {  // renormalise: divide the state by q (the one reciprocal on the chain, once per RN levels)
        const R rq = t_rcp(q);
        q2 = (s % RN == 0) ? rq : q1 * rq;
        z1 = z * rq;
        q1 = (R)1;
      } else {
        q2 = (s % RN == 0) ? (R)1 : q1;
        q1 = q;
        z1 = z;
      }
      cp = c[s];
    }
#pragma unroll
    for (int s = 0; s < LB; ++s) {
      const size_t o = (size_t)lev_of(k, s) * 32 + lane;
      sd[o] = o0[s];
      sx[o] = o1[s];
    }
#pragma unroll
    for (int s = 0; s < LB; ++s) {
      a[s] = na[s];
      b[s] = nb[s];
      c[s] = nc[s];
      x[s] = nx[s];
    }
  }

  // ---------------------------------------------------------------- interface: x[m-1] + cT x[m] = yT, cB x[m-1] + x[m] = yB
  R xv = (R)0;  // the unknown just beyond this half, as the substitution below needs it
  if (other) {
    const R r = t_rcp(qlast);
    const R yI = zlast * r, cI = clast * (qprev_last * r);
    R* mine = xch + (UP ? 64 : 0);
    const R* theirs = xch + (UP ? 0 : 64);
    mine[lane] = yI;
    mine[32 + lane] = cI;
    asm volatile("bar.sync 1, 64;" ::: "memory");
    const R yO = theirs[lane], cO = theirs[32 + lane];
    xv = (yO - cO * yI) * t_rcp((R)1 - cO * cI);  // the neighbouring half's boundary unknown
  }

  // ---------------------------------------------------------------- substitution outwards, stored from registers
  // x[i] = y[i] - c'[i] x[next] with y = z/q, c' = c q[prev]/q (q[prev] == 1 where the state was renormalised);
  // only the last FMA is on the dependent chain.
  R qq[LB], zz[LB], uu[LB];
  auto stage_b = [&](int k, R (&rq)[LB], R (&rz)[LB], R (&ru)[LB]) {
#pragma unroll
    for (int s = 0; s < LB; ++s) {
      const int i = lev_of(k, s);
      const size_t o = (size_t)i * 32 + lane;
      rq[s] = sd[o];
      rz[s] = sx[o];
      if (MODE == 1) ru[s] = UP ? (i > 0 ? -sdu[o - 32] : (R)0) : -sdu[o];
      else ru[s] = UP ? sdl[o] : sdu[o];
      if (i == (UP ? 0 : nlev - 1) || i >= nlev) ru[s] = (R)0;
    }
  };
  if (nk > 0) stage_b(chunk_of(nk - 1), qq, zz, uu);
  for (int j = nk - 1; j >= 0; --j) {
    const int k = chunk_of(j);
    R nq[LB], nz[LB], nu[LB];
    if (j > 0) stage_b(chunk_of(j - 1), nq, nz, nu);
    R y[LB], cc[LB], o[LB];
#pragma unroll
    for (int s = 0; s < LB; ++s) {
      const R r = t_rcp(qq[s]);
      const R g = (s % RN == 0) ? r : qq[s % RN == 0 ? s : s - 1] * r;
      y[s] = zz[s] * r;
      cc[s] = uu[s] * g;
    }
#pragma unroll
    for (int s = LB - 1; s >= 0; --s) {
      xv = f_fma(-cc[s], xv, y[s]);
      o[s] = xv;
    }
    if (full_tile && (k + 1) * LB <= nlev) {  // warp-uniform: no branch around the individual stores
#pragma unroll
      for (int s = LB - 1; s >= 0; --s) __stcs(gx + (size_t)lev_of(k, s) * x_stride, o[s]);
    } else if (live) {
#pragma unroll
      for (int s = LB - 1; s >= 0; --s)
        if (lev_of(k, s) < nlev) __stcs(gx + (size_t)lev_of(k, s) * x_stride, o[s]);
    }
#pragma unroll
    for (int s = 0; s < LB; ++s) {
      qq[s] = nq[s];
      zz[s] = nz[s];
      uu[s] = nu[s];
    }
  }
}

template <typename R, int RN, int MODE>
__global__ void __launch_bounds__(64)
k_tridiag_small(const __grid_constant__ CUtensorMap mdl, const __grid_constant__ CUtensorMap md,
                const __grid_constant__ CUtensorMap mdu, const __grid_constant__ CUtensorMap mx, R* __restrict__ X, int64_t ncol,
                int nlev, int64_t x_stride, int GL) {
  constexpr int LB = kLB;
  static_assert(LB % RN == 0, "renormalisation interval must divide the chunk");
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nch = (nlev + LB - 1) / LB;
  const int ngrp = (nlev + GL - 1) / GL, gch = GL / LB;  // box groups per plane, chunks per group
  const int kt = (nch + 1) / 2;                          // chunks [0, kt) go to the top half, [kt, nch) to the bottom half
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem_raw);
  R* xch = reinterpret_cast<R*>(smem_raw + 128);       // interface exchange: [half][y | c][lane]
  R* P = reinterpret_cast<R*>(smem_raw + 128 + 1024);
  const size_t plane = (size_t)ngrp * GL * 32;
  // MODE 1 has no dl plane: md = the level coordinate z (its plane receives q), mdu = kd (its plane receives w)
  R* sdl = P;
  R* sd = MODE == 1 ? P : P + plane;
  R* sdu = sd + plane;
  R* sx = sdu + plane;
  const int c0 = blockIdx.x * 32;
  const int row0 = blockIdx.y * nlev;

  if (threadIdx.x == 0) {
    for (int g = 0; g < ngrp; ++g) mbar_init(&bar[g], 1);
    fence_mbar_init();
    const uint64_t pol = l2_policy_evict_first();
    for (int i = 0; i < ngrp; ++i) {
      const int g = (i & 1) ? ngrp - 1 - i / 2 : i / 2;  // outermost groups first: both halves start at once
      // a box past the last level of this matrix reads rows of the next one (or zeros past the array): neutralised at staging
      mbar_expect_tx(&bar[g], (MODE == 1 ? 3u : 4u) * (unsigned)GL * 32u * (unsigned)sizeof(R));
      const size_t o = (size_t)g * GL * 32;
      if (MODE == 0) tma_load_2d_hint(sdl + o, &mdl, c0, row0 + g * GL, &bar[g], pol);
      tma_load_2d_hint(sd + o, &md, c0, row0 + g * GL, &bar[g], pol);
      tma_load_2d_hint(sdu + o, &mdu, c0, row0 + g * GL, &bar[g], pol);
      tma_load_2d_hint(sx + o, &mx, c0, row0 + g * GL, &bar[g], pol);
    }
  }
  __syncthreads();  // barrier initialisation visible to both warps

  if (MODE == 1) {
    // the finite-difference step, fused: w[i] = kd[i] / (z[i+1] - z[i]) -- ekat::adj_diff<Forward> (reference
    // src/pack/ekat_pack_kokkos.hpp:139-152) followed by the caller's coefficient formula -- for every level of the
    // tile, in place over kd (one IEEE division each, so the implied matrix is bit-defined); w[n-1] = 0.
    for (int g = 0; g < ngrp; ++g) mbar_wait(&bar[g], 0u);
    for (int i = warp; i < nch * LB; i += 2) {
      const size_t o = (size_t)i * 32 + lane;
      const R dz = i < nlev - 1 ? x_sub(sd[o + 32], sd[o]) : (R)1;
      sdu[o] = i < nlev - 1 ? x_div(sdu[o], dz) : (R)0;
    }
    __syncthreads();
  }
  const bool full_tile = (int64_t)c0 + 32 <= ncol, live = (int64_t)c0 + lane < ncol;
  R* gx = X + (size_t)row0 * x_stride + c0 + lane;
  if (warp == 0) small_half<R, RN, false, MODE>(sdl, sd, sdu, sx, bar, xch, lane, nlev, nch, gch, kt, gx, x_stride, full_tile, live);
  else if (kt < nch) small_half<R, RN, true, MODE>(sdl, sd, sdu, sx, bar, xch, lane, nlev, nch, gch, kt, gx, x_stride, full_tile, live);
}

// Level-major operands: dl,d,du [nmat][nlev][a_stride], X [nmat][nlev][x_stride]. Returns EKAT_B200_ERR_UNSUPPORTED
// when the operands do not meet TMA's requirements or the batch is not one this kernel wins on (the caller then
// uses the pipelined kernels); `force` (tuning) skips that test as long as one tile fits.
// MODE 1: d = level coordinate z, du = kd (dl unused), one matrix and one right-hand side per column.
template <typename R, int MODE>
static int run_small(R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nmat, int64_t a_stride, int64_t x_stride, bool force,
                     cudaStream_t stream) {
  auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  if ((MODE == 0 && !al(dl)) || !al(d) || !al(du) || !al(X) || (a_stride * sizeof(R)) % 16 || (x_stride * sizeof(R)) % 16)
    return EKAT_B200_ERR_UNSUPPORTED;
  if (ncol >= (int64_t)1 << 31 || (int64_t)nmat * nlev >= (int64_t)1 << 31 || nmat > 65535) return EKAT_B200_ERR_UNSUPPORTED;
  const int nch = (nlev + kLB - 1) / kLB;
  const int gch = (nch + kSmallGroups - 1) / kSmallGroups, GL = gch * kLB;  // levels per TMA box (<= 256)
  if (GL > 256) return EKAT_B200_ERR_UNSUPPORTED;
  const int ngrp = (nlev + GL - 1) / GL;
  const size_t smem = 128 + 1024 + (MODE == 1 ? 3 : 4) * (size_t)ngrp * GL * 32 * sizeof(R);
  if (smem > kMaxSmemPerCta) return EKAT_B200_ERR_UNSUPPORTED;
  // Measured (profiles/r02_small_batches.txt): with >= 3 tiles per SM and at most two waves of tiles this kernel wins
  // (21,600 x 72 f64: 11.1 us against 13.0-14.1 us); longer columns (128 levels: 2-3 tiles per SM at best) and larger
  // batches stream better through the pipelined kernels, which keep 2 planes + rings per tile instead of 4 planes.
  const int per_sm = (int)((kMaxSmemPerCta + 1024) / (smem + 1024));
  const int64_t tiles = ((ncol + 31) / 32) * nmat;
  if (!force && (per_sm < 3 || nlev > 96 || tiles > (int64_t)2 * kNumSms * per_sm)) return EKAT_B200_ERR_UNSUPPORTED;
  CUtensorMap mdl, md, mdu, mx;
  const int64_t rows = (int64_t)nmat * nlev;
  if (!make_tensor_map_2d<R>(&md, d, ncol, rows, a_stride, 32, GL, 0) || !make_tensor_map_2d<R>(&mdu, du, ncol, rows, a_stride, 32, GL, 0) ||
      !make_tensor_map_2d<R>(&mx, X, ncol, rows, x_stride, 32, GL, 0))
    return EKAT_B200_ERR_UNSUPPORTED;
  if (MODE == 0) {
    if (!make_tensor_map_2d<R>(&mdl, dl, ncol, rows, a_stride, 32, GL, 0)) return EKAT_B200_ERR_UNSUPPORTED;
  } else {
    mdl = md;
  }
  auto k = k_tridiag_small<R, TmaCfg<R>::RN, MODE>;
  int rc = allow_smem(k, smem);
  if (rc) return rc;
  dim3 grid((unsigned)((ncol + 31) / 32), (unsigned)nmat);
  k<<<grid, 64, smem, stream>>>(mdl, md, mdu, mx, X, ncol, nlev, x_stride, GL);
  note_launch(MODE == 1 ? (sizeof(R) == 8 ? "diffusion_solve<f64>" : "diffusion_solve<f32>")
                        : (sizeof(R) == 8 ? "tridiag_small<f64,thomas>" : "tridiag_small<f32,thomas>"));
  return check_cuda(cudaGetLastError());
}

template <typename R>
int tridiag_small(R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nmat, int64_t a_stride, int64_t x_stride, bool force,
                  cudaStream_t stream) {
  return run_small<R, 0>(dl, d, du, X, ncol, nlev, nmat, a_stride, x_stride, force, stream);
}

// The fused "difference -> (dl, d, du) -> solve" step: z, kd [nlev][pitch] level-major, X
// [nlev][pitch] in/out. EKAT_B200_ERR_UNSUPPORTED when the operands do not meet TMA's requirements or a tile of
// 3 planes x nlev x 32 columns does not fit on chip.
template <typename R>
int diffusion_solve(const R* z, const R* kd, R* X, int64_t ncol, int nlev, int64_t pitch, cudaStream_t stream) {
  return run_small<R, 1>(nullptr, const_cast<R*>(z), const_cast<R*>(kd), X, ncol, nlev, 1, pitch, pitch, true, stream);
}

// The same stencil unfused: writes the matrix the fused kernel implies (for callers that want it, and as the two-pass
// baseline the fusion is measured against): w[i] = kd[i] / (z[i+1] - z[i]); dl[i] = -w[i-1]; du[i] = -w[i];
// d[i] = (1 + w[i]) + w[i-1]; w[-1] = w[n-1] = 0. One thread per column, rows coalesced.
template <typename R>
__global__ void __launch_bounds__(128)
k_diffusion_matrix(const R* __restrict__ z, const R* __restrict__ kd, R* __restrict__ dl, R* __restrict__ d, R* __restrict__ du,
                   int64_t ncol, int nlev, int64_t pitch) {
  const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= ncol) return;
  R wm = (R)0, zi = ld_stream(z + c);
  for (int i = 0; i < nlev; ++i) {
    const size_t o = (size_t)i * pitch + c;
    R wi = (R)0, zn = zi;
    if (i < nlev - 1) {
      zn = ld_stream(z + o + pitch);
      wi = x_div(ld_stream(kd + o), x_sub(zn, zi));
    }
    st_stream(dl + o, -wm);
    st_stream(du + o, -wi);
    st_stream(d + o, x_add(x_add((R)1, wi), wm));
    wm = wi;
    zi = zn;
  }
}
template <typename R>
int diffusion_matrix(const R* z, const R* kd, R* dl, R* d, R* du, int64_t ncol, int nlev, int64_t pitch, cudaStream_t stream) {
  if (ncol == 0) return EKAT_B200_OK;
  k_diffusion_matrix<R><<<(unsigned)((ncol + 127) / 128), 128, 0, stream>>>(z, kd, dl, d, du, ncol, nlev, pitch);
  note_launch("diffusion_matrix");
  return check_cuda(cudaGetLastError());
}

template int tridiag_small<double>(double*, double*, double*, double*, int64_t, int, int, int64_t, int64_t, bool, cudaStream_t);
template int tridiag_small<float>(float*, float*, float*, float*, int64_t, int, int, int64_t, int64_t, bool, cudaStream_t);
template int diffusion_solve<double>(const double*, const double*, double*, int64_t, int, int64_t, cudaStream_t);
template int diffusion_solve<float>(const float*, const float*, float*, int64_t, int, int64_t, cudaStream_t);
template int diffusion_matrix<double>(const double*, const double*, double*, double*, double*, int64_t, int, int64_t, cudaStream_t);
template int diffusion_matrix<float>(const float*, const float*, float*, float*, float*, int64_t, int, int64_t, cudaStream_t);

}  // namespace ekb
