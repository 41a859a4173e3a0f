// tridiag_tma.cu -- batched Thomas / bfb on the native level-major layout, TMA-fed, one warp per 32-lane tile.
//
// Replaces the per-column loops around ekat::tridiag::thomas (reference src/algorithm/ekat_tridiag.hpp:323-381)
// and ekat::tridiag::bfb (:292-320, :632-678) when one matrix belongs to one right-hand side (nA == nrhs).
//
// Why this shape (measured on B200, profiles/r01a_*): a column's forward sweep must stay on chip until its
// back-substitution (2 words per level, 3 for bfb), so shared memory admits only ~100 f64 columns of 128 levels
// per SM -- about three warps. The previous cp.async kernel spent those few warps on per-thread copy bookkeeping
// (129 instructions per level and 16 active lanes per warp) and on the ~80-cycle divide chain of every level.
// Here
//  * a CTA is ONE full warp that owns a tile of W = 32*CPT adjacent columns; three such CTAs share an SM;
//  * all operands arrive by 2D TMA tensor copies (cp.async.bulk.tensor, one instruction per LB levels x W
//    columns per array, issued by lane 0, completion on mbarriers): x and du land directly in the planes the
//    sweeps update in place, dl and d go through a small ring; no thread computes a global address;
//  * solutions are written in place into the plane and leave by TMA tensor stores, LB levels at a time;
//  * thomas uses the continuant ("ratio") form of the recurrence, which has no division on its dependent
//    chain:   q[i] = d[i] q[i-1] - dl[i] du[i-1] q[i-2],   z[i] = x[i] q[i-1] - dl[i] z[i-1]
//             c'[i] = du[i] q[i-1] / q[i],   y[i] = z[i] / q[i],   x[i] = y[i] - c'[i] x[i+1]
//    (q[i] is the i-th leading minor; c', y are the classical Thomas intermediates). 1/q[i] is off the chain, so
//    consecutive levels pipeline; every RN levels the state is renormalised to q = 1 (exactly the classical
//    step) so that minors cannot overflow: pivots may range over 2^+-(1022/RN) (f64) / 2^+-(126/RN) (f32).
//  * bfb keeps the reference's operation order with separately rounded * - / (bit-identical); its IEEE
//    division has the reciprocal part hoisted off the back-substitution chain (common.cuh).
// HBM traffic is the algorithmic minimum: read dl, d, du, x once, write x once.
#include <cuda.h>

#include "tridiag_tma.cuh"

namespace ekb {

// ------------------------------------------------------------------------------------------- tensor maps
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      return (EncodeTiledFn) nullptr;
    return (EncodeTiledFn)p;
  }();
  return fn;
}

// [nrows][pitch] array of R, of which the first ncol entries of a row are data; box = boxc columns x boxr rows.
// Columns beyond ncol read as zero and are not written (TMA bounds handling).
// Encoding a descriptor costs the host a few microseconds and a launch needs four or five of them -- more than the
// whole kernel takes at config 1 size (21,600 x 72: ~10 us of HBM time). A descriptor is a pure function of
// (address, element size, extents, pitch, box, swizzle); it holds no data and does not go stale when the memory
// behind the address is rewritten or re-allocated, so each host thread keeps the ones it built in a small
// direct-mapped cache. A physics driver calls the solvers every step on the same views: after the first step a call
// is argument checks + one kernel launch.
namespace {
struct TmapKey {
  const void* base;
  int64_t ncol, nrows, pitch;
  int boxc, boxr, swz, esize;
  bool operator==(const TmapKey& o) const {
    return base == o.base && ncol == o.ncol && nrows == o.nrows && pitch == o.pitch && boxc == o.boxc && boxr == o.boxr &&
           swz == o.swz && esize == o.esize;
  }
};
struct TmapEntry {
  TmapKey key;
  bool valid;
  alignas(64) CUtensorMap map;
};
constexpr int kTmapSlots = 256;
thread_local TmapEntry t_tmaps[kTmapSlots];
inline unsigned tmap_hash(const TmapKey& k) {
  uint64_t h = reinterpret_cast<uintptr_t>(k.base) * 0x9E3779B97F4A7C15ull;
  h ^= (uint64_t)k.ncol * 0xC2B2AE3D27D4EB4Full + (uint64_t)k.nrows * 0x165667B19E3779F9ull + (uint64_t)k.pitch * 0x27D4EB2F165667C5ull;
  h ^= (uint64_t)(k.boxc * 131 + k.boxr * 31 + k.swz * 7 + k.esize) * 0xD6E8FEB86659FD93ull;
  return (unsigned)(h >> 40) % kTmapSlots;
}
}  // namespace

template <typename R>
bool make_tensor_map_2d(CUtensorMap* m, const R* base, int64_t ncol, int64_t nrows, int64_t pitch, int boxc, int boxr, int swizzle64) {
  const TmapKey key{base, ncol, nrows, pitch, boxc, boxr, swizzle64, (int)sizeof(R)};
  TmapEntry& e = t_tmaps[tmap_hash(key)];
  if (e.valid && e.key == key) {
    *m = e.map;
    return true;
  }
  EncodeTiledFn enc = encode_fn();
  if (!enc) return false;
  const cuuint64_t dims[2] = {(cuuint64_t)ncol, (cuuint64_t)nrows};
  const cuuint64_t strides[1] = {(cuuint64_t)pitch * sizeof(R)};
  const cuuint32_t box[2] = {(cuuint32_t)boxc, (cuuint32_t)boxr};
  const cuuint32_t estr[2] = {1, 1};
  const CUtensorMapDataType dt = sizeof(R) == 8 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT64 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  if (enc(m, dt, 2, const_cast<R*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
          swizzle64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
    return false;
  e.key = key;
  e.map = *m;
  e.valid = true;
  return true;
}
template bool make_tensor_map_2d<double>(CUtensorMap*, const double*, int64_t, int64_t, int64_t, int, int, int);
template bool make_tensor_map_2d<float>(CUtensorMap*, const float*, int64_t, int64_t, int64_t, int, int, int);

// ------------------------------------------------------------------------------------------- the kernel
// grid = (column tiles, matrices); block = 32 threads. Lane l owns columns c0 + l*CPT .. +CPT-1 (CM: c0 + l + 32*j).
// Shared memory: [ring barriers NS][plane barriers nch] | planes [level][lane] | rings [stage][LB][lane].
//   FAST : P0 = du -> c',  P1 = x -> y -> solution;            rings dl, d
//   EXACT: P0 = d  -> d',  P1 = x -> x' -> solution,  P2 = du;  ring  dl
// IO (thomas only): a second warp, one thread of it, issues every copy -- a tensor copy costs its issuing thread ~130
// cycles and the sweeping warp of the one-warp form issued five per chunk of 8 levels, more than its whole arithmetic
// in f32. The sweeping warp releases a ring stage as soon as its operands are in registers (ebar) and announces a
// finished chunk of the solution plane (sbar[k], used once); the io thread refills and stores.
template <typename R, bool EXACT, int CPT, int RN, int NS, int LBT, bool CM, bool IO = false>
__global__ void __launch_bounds__(IO ? 64 : 32)
k_tridiag_tma(const __grid_constant__ CUtensorMap mdl, const __grid_constant__ CUtensorMap md,
              const __grid_constant__ CUtensorMap mdu, const __grid_constant__ CUtensorMap mx, R* __restrict__ X,
              int64_t ncol, int nlev, int64_t x_stride, int pd) {
  using V = Vec<R, CPT>;
  constexpr int LB = LBT, W = 32 * CPT;
  constexpr unsigned CHUNK_BYTES = LB * W * sizeof(R);
  static_assert(LB % RN == 0, "renormalisation interval must divide the chunk");
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  static_assert(!(IO && EXACT), "the io warp serves the thomas form only");
  const int lane = threadIdx.x & 31;
  const int nfull = nlev / LB, ntail = nlev - nfull * LB, nch = nfull + (ntail ? 1 : 0);

  // barriers: ring[NS] | plane[nch] | backward du ring[2*NS] (EXACT) or stage released[NS] + chunk solved[nch] (IO)
  uint64_t* rbar = reinterpret_cast<uint64_t*>(smem_raw);
  uint64_t* pbar = rbar + NS;
  uint64_t* bbar = pbar + nch;
  uint64_t* ebar = bbar;       // IO
  uint64_t* sbar = ebar + NS;  // IO
  const size_t off = ((size_t)(3 * NS + 2 * nch) * 8 + 1023) & ~(size_t)1023;  // chunks sit on 1 KB boundaries (swizzle atoms are 512 B)
  const size_t plane = (size_t)nch * LB * 32;  // V units
  V* P0 = reinterpret_cast<V*>(smem_raw + off);
  V* P1 = P0 + plane;
  V* RA = P1 + plane;                 // dl ring [NS][LB][32]
  V* RB = RA + (size_t)NS * LB * 32;  // FAST: d ring; EXACT: du ring. EXACT back-substitution: RA..RB = one du ring of 2*NS stages

  const int c0 = blockIdx.x * W;
  const int row0 = blockIdx.y * nlev;

  if (threadIdx.x == 0) {
    for (int s = 0; s < NS; ++s) mbar_init(&rbar[s], 1);
    for (int k = 0; k < nch; ++k) mbar_init(&pbar[k], 1);
    if (EXACT)
      for (int s = 0; s < 2 * NS; ++s) mbar_init(&bbar[s], 1);
    if (IO)
      for (int s = 0; s < NS + nch; ++s) mbar_init(&ebar[s], 1);
    fence_mbar_init();
  }
  if (IO) __syncthreads();
  else __syncwarp();

  // L2 policies: every operand is read once (evict-first) except bfb's du, which the back-substitution reads
  // again ~20 us later: it is kept in the 126 MB L2 (evict-last) instead of in a third shared-memory plane,
  // which is what lets three bfb CTAs share an SM.
  const uint64_t pol_stream = l2_policy_evict_first(), pol_keep = l2_policy_evict_last();

  // chunk k of a tensor: box coordinates (fastest, slowest) = (columns, levels) or, column-major, (levels, columns)
  auto load_chunk = [&](void* dst, const CUtensorMap* m, int k, uint64_t* bar, uint64_t pol) {
    if (CM) tma_load_2d_hint(dst, m, k * LB, c0, bar, pol);
    else tma_load_2d_hint(dst, m, c0, row0 + k * LB, bar, pol);
  };
  auto issue_ring = [&](int k) {
    const int s = k % NS;
    mbar_expect_tx(&rbar[s], 2u * CHUNK_BYTES);
    load_chunk(RA + (size_t)s * LB * 32, &mdl, k, &rbar[s], pol_stream);
    if (EXACT) load_chunk(RB + (size_t)s * LB * 32, &mdu, k, &rbar[s], pol_keep);
    else load_chunk(RB + (size_t)s * LB * 32, &md, k, &rbar[s], pol_stream);
  };
  auto issue_plane = [&](int k) {
    const size_t o = (size_t)k * LB * 32;
    mbar_expect_tx(&pbar[k], 2u * CHUNK_BYTES);
    load_chunk(P0 + o, EXACT ? &md : &mdu, k, &pbar[k], pol_stream);
    load_chunk(P1 + o, &mx, k, &pbar[k], pol_stream);
  };
  auto issue_back = [&](int j) {  // EXACT: du of chunk nch-1-j for the back-substitution
    const int s = j % (2 * NS);
    mbar_expect_tx(&bbar[s], CHUNK_BYTES);
    load_chunk(RA + (size_t)s * LB * 32, &mdu, nch - 1 - j, &bbar[s], pol_stream);
  };
  if (IO && threadIdx.x >= 32) {
    // ------------------------------------------------------------ io warp
    if (lane != 0) return;
    const int n0 = pd > NS ? pd : NS;
    for (int k = 0; k < n0 && k < nch; ++k) {
      if (k < NS) issue_ring(k);
      if (k < pd) issue_plane(k);
    }
    unsigned eph = 0;
    for (int k = 0; k < nch; ++k) {
      const int s = k % NS;
      mbar_wait(&ebar[s], (eph >> s) & 1u);
      eph ^= 1u << s;
      if (k + NS < nch) issue_ring(k + NS);
      if (k + pd < nch) issue_plane(k + pd);
    }
    for (int k = nch - 1; k >= 0; --k) {
      mbar_wait(&sbar[k], 0u);
      if (CM) {
        tma_store_2d(&mx, k * LB, c0, P1 + (size_t)k * LB * 32);
        tma_store_commit();
      } else if (k != nfull) {
        tma_store_2d(&mx, c0, row0 + k * LB, P1 + (size_t)k * LB * 32);
        tma_store_commit();
      }
    }
    tma_store_wait_read<0>();  // shared memory must outlive the stores' reads
    return;
  }
  if (!IO && lane == 0) {
    const int n0 = pd > NS ? pd : NS;
    for (int k = 0; k < n0 && k < nch; ++k) {
      if (k < NS) issue_ring(k);
      if (k < pd) issue_plane(k);
    }
  }

  // ---------------------------------------------------------------- forward sweep
  // FAST state: q1 = q[i-1], q2 = q[i-2], z1 = z[i-1], cp = du[i-1].  EXACT state: q1 = d'[i-1], z1 = x'[i-1], cp = du[i-1].
  R q1[CPT], q2[CPT], z1[CPT], cp[CPT];
#pragma unroll
  for (int j = 0; j < CPT; ++j) {
    q1[j] = (R)1;
    q2[j] = (R)0;
    z1[j] = (R)0;
    cp[j] = (R)0;
  }
  for (int k = 0; k < nch; ++k) {
    const int s = k % NS;
    mbar_wait(&rbar[s], (unsigned)(k / NS) & 1u);
    mbar_wait(&pbar[k], 0u);
    const V* ra = RA + (size_t)s * LB * 32;
    const V* rb = RB + (size_t)s * LB * 32;
    V* p0 = P0 + (size_t)k * LB * 32;
    V* p1 = P1 + (size_t)k * LB * 32;
    auto ld = [&](const V* chunk, int u) { return t_ld<R, CPT, LB, CM>(chunk, u, lane); };
    // stage the whole chunk in registers first: one shared-memory latency per chunk instead of one per level
    // (the in-place plane updates below would otherwise order every level's loads behind the previous stores)
    V a[LB], b[LB], c[LB], x[LB], o0[LB], o1[LB];
#pragma unroll
    for (int u = 0; u < LB; ++u) {
      a[u] = ld(ra, u);
      x[u] = ld(p1, u);
      b[u] = EXACT ? ld(p0, u) : ld(rb, u);
      c[u] = EXACT ? ld(rb, u) : ld(p0, u);
    }
    if (IO) {  // the stage's operands are in registers: it can be refilled while this chunk is swept
      __syncwarp();
      if (lane == 0) mbar_arrive(&ebar[s]);
    }
    if (EXACT) {
      R sd[CPT], sx[CPT], su[CPT];
#pragma unroll
      for (int j = 0; j < CPT; ++j) {
        sd[j] = q1[j];
        sx[j] = z1[j];
        su[j] = cp[j];
      }
      bool bad = false;
      bfb_fwd_chunk<R, CPT, LB, false>(a, b, c, x, o0, o1, q1, z1, cp, k == 0, bad);
      if (__any_sync(0xffffffffu, bad)) {  // rare operands somewhere in the warp: redo with the full-range division
#pragma unroll
        for (int j = 0; j < CPT; ++j) {
          q1[j] = sd[j];
          z1[j] = sx[j];
          cp[j] = su[j];
        }
        bfb_fwd_chunk<R, CPT, LB, true>(a, b, c, x, o0, o1, q1, z1, cp, k == 0, bad);
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
          R q, z;
          if (u % RN == 0) {  // state was renormalised: q[i-1] == 1, q2 = q[i-2]/q[i-1]
            q = f_fma(-(a[u].v[j] * cp[j]), q2[j], b[u].v[j]);
            z = f_fma(-a[u].v[j], z1[j], x[u].v[j]);
          } else {
            q = f_fma(b[u].v[j], q1[j], -((a[u].v[j] * cp[j]) * q2[j]));
            z = f_fma(-a[u].v[j], z1[j], x[u].v[j] * q1[j]);
          }
          const R rq = t_rcp(q);
          const R g = (u % RN == 0) ? rq : q1[j] * rq;  // q[i-1]/q[i] = 1/pivot
          const R y = z * rq;
          o0[u].v[j] = c[u].v[j] * g;
          o1[u].v[j] = y;
          if ((u + 1) % RN == 0) {  // renormalise: divide the state by q[i]
            q2[j] = g;
            z1[j] = y;
            q1[j] = (R)1;
          } else {
            q2[j] = (u % RN == 0) ? (R)1 : q1[j];
            q1[j] = q;
            z1[j] = z;
          }
          cp[j] = c[u].v[j];
        }
      }
    }
#pragma unroll
    for (int u = 0; u < LB; ++u) {
      t_st<R, CPT, LB, CM>(p0, u, lane, o0[u]);
      t_st<R, CPT, LB, CM>(p1, u, lane, o1[u]);
    }
    __syncwarp();  // every lane has read ring stage s
    if (!IO && lane == 0) {
      if (k + NS < nch) issue_ring(k + NS);
      if (k + pd < nch) issue_plane(k + pd);
    }
  }
  // du(nrow-1) is never used (:24-27): the coupling of the last level is zeroed so the back-substitution is uniform
  if (!EXACT) {
    V zero;
#pragma unroll
    for (int j = 0; j < CPT; ++j) zero.v[j] = (R)0;
    t_st<R, CPT, LB, CM>(P0 + (size_t)((nlev - 1) / LB) * LB * 32, (nlev - 1) % LB, lane, zero);
  } else if (lane == 0) {
    for (int j = 0; j < 2 * NS && j < nch; ++j) issue_back(j);  // both rings are free now: refill them with du from L2
  }

  // ---------------------------------------------------------------- back substitution
  R xv[CPT];
#pragma unroll
  for (int j = 0; j < CPT; ++j) xv[j] = (R)0;
  for (int jb = 0; jb < nch; ++jb) {
    const int k = nch - 1 - jb;
    V* q1p = P1 + (size_t)k * LB * 32;
    const V* q0p = P0 + (size_t)k * LB * 32;
    const V* bu = RA + (size_t)(jb % (2 * NS)) * LB * 32;
    auto ld = [&](const V* chunk, int u) { return t_ld<R, CPT, LB, CM>(chunk, u, lane); };
    if (EXACT) mbar_wait(&bbar[jb % (2 * NS)], (unsigned)(jb / (2 * NS)) & 1u);
    if (k == nfull) {
      // partial top chunk. Level-major: a tensor store would spill into the next matrix' rows -> plain streaming
      // stores. Column-major: the box is clipped at the end of the column, so it leaves by tensor store like the rest.
      const int64_t col = (int64_t)c0 + lane * CPT;
      for (int u = ntail - 1; u >= 0; --u) {
        const V y = ld(q1p, u), m0 = ld(q0p, u);
        V o;
        if (EXACT) {
          V ui = ld(bu, u);
          if (u == ntail - 1) {
#pragma unroll
            for (int j = 0; j < CPT; ++j) ui.v[j] = (R)0;
          }
#pragma unroll
          for (int j = 0; j < CPT; ++j) {
            xv[j] = x_div(x_sub(y.v[j], x_mul(ui.v[j], xv[j])), m0.v[j]);  // :317-319
            o.v[j] = xv[j];
          }
        } else {
#pragma unroll
          for (int j = 0; j < CPT; ++j) {
            xv[j] = f_fma(-m0.v[j], xv[j], y.v[j]);
            o.v[j] = xv[j];
          }
        }
        if (CM) {
          t_st<R, CPT, LB, CM>(q1p, u, lane, o);
        } else {
          R* g = X + (size_t)(row0 + k * LB + u) * x_stride + col;
#pragma unroll
          for (int j = 0; j < CPT; ++j)
            if (col + j < ncol) __stcs(g + j, o.v[j]);
        }
      }
      if (CM) fence_proxy_async();
    } else {
      V y[LB], m0[LB], ui[LB];
#pragma unroll
      for (int u = 0; u < LB; ++u) {
        y[u] = ld(q1p, u);
        m0[u] = ld(q0p, u);
        if (EXACT) ui[u] = ld(bu, u);
      }
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
        bfb_bwd_chunk<R, CPT, LB, false>(y, m0, ui, xv, bad);
        if (__any_sync(0xffffffffu, bad)) {
#pragma unroll
          for (int j = 0; j < CPT; ++j) xv[j] = sxv[j];
#pragma unroll
          for (int u = 0; u < LB; ++u) y[u] = ys[u];
          bfb_bwd_chunk<R, CPT, LB, true>(y, m0, ui, xv, bad);
        }
      } else {
#pragma unroll
        for (int u = LB - 1; u >= 0; --u) {
#pragma unroll
          for (int j = 0; j < CPT; ++j) {
            xv[j] = f_fma(-m0[u].v[j], xv[j], y[u].v[j]);
            y[u].v[j] = xv[j];
          }
        }
      }
#pragma unroll
      for (int u = 0; u < LB; ++u) t_st<R, CPT, LB, CM>(q1p, u, lane, y[u]);
      fence_proxy_async();  // generic-proxy writes of this chunk -> visible to the TMA store
    }
    __syncwarp();  // every lane has read its du stage / written its part of the chunk
    if (IO) {
      if (lane == 0) mbar_arrive(&sbar[k]);
    } else if (lane == 0) {
      if (CM) {
        tma_store_2d(&mx, k * LB, c0, P1 + (size_t)k * LB * 32);
        tma_store_commit();
      } else if (k != nfull) {
        tma_store_2d(&mx, c0, row0 + k * LB, P1 + (size_t)k * LB * 32);
        tma_store_commit();
      }
      if (EXACT && jb + 2 * NS < nch) issue_back(jb + 2 * NS);
    }
  }
  if (!IO && lane == 0) tma_store_wait_read<0>();  // shared memory must outlive the stores' reads
}

// ------------------------------------------------------------------------------------------- bfb, x' kept in L2
// bfb is bound by its division chain per level (ncu: the one warp of a tile is either issuing or waiting on a
// fixed-latency dependency 72% of the time), so its throughput is the number of tiles an SM can hold. This kernel
// keeps only the pivots d' of a tile in shared memory; the forward sweep writes x' back IN PLACE over X -- exactly
// what the reference's sweep does (ekat_tridiag.hpp:305-320) -- with an evict-last hint, and the back-substitution
// reads it (and du, as before) again through the ring. A tile's x' (32 KB) is re-read ~20 us after it was written
// and overwritten by the solution right after, so it lives and dies in the 126 MB L2: HBM still sees one read of
// dl, d, du, X and one write of X. Shared memory per tile drops from 72 KB to ~51 KB: four tiles per SM instead of
// three. The top chunk is never spilled: the back-substitution starts from it where it lies.
// Ring stage = [dl | du | x], LB levels each. Use u < nch: forward chunk u. Use u >= nch: (du, x') of chunk
// 2*nch-2-u for the back-substitution. A stage is refilled once the tensor store that reads it has done so.
// IO = 0: one warp per tile does everything. IO = 1: a second warp (one thread of it) issues every copy.
// IO = 2: as 1, and TWO sweeping warps take the chunks of the tile in turn: the dependent chain of a chunk is only
// ~45% of the time a single warp spends on it (ncu: the rest is waiting for the copy, 32 LDS, 16 STS, the proxy fence,
// the arrival, index arithmetic), so while one warp runs the chain of chunk k the other one stores the results of
// chunk k-1 and stages the operands of chunk k+1 in registers. The chain state (d', x', du of the previous level;
// x of the level above on the way back) crosses in shared memory under a pair of named barriers (bar.arrive by the
// warp that finished, bar.sync by the warp that continues).
template <typename R, int CPT, int NS, int LBT, bool CM, int IO>
__global__ void __launch_bounds__(IO == 2 ? 96 : IO ? 64 : 32)
k_tridiag_bfb_spill(const __grid_constant__ CUtensorMap mdl, const __grid_constant__ CUtensorMap md,
                    const __grid_constant__ CUtensorMap mdu, const __grid_constant__ CUtensorMap mx, R* __restrict__ X,
                    int64_t ncol, int nlev, int64_t x_stride, int pd) {
  using V = Vec<R, CPT>;
  constexpr int LB = LBT, W = 32 * CPT, CH = LB * 32;
  constexpr unsigned CHUNK_BYTES = LB * W * sizeof(R);
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  constexpr int NCW = IO == 2 ? 2 : 1;  // sweeping warps
  constexpr bool PP = IO == 2;
  const int lane = threadIdx.x & 31, cw = threadIdx.x >> 5;
  const bool io_warp = IO && cw == NCW;
  const int nfull = nlev / LB, ntail = nlev - nfull * LB, nch = nfull + (ntail ? 1 : 0);
  const int NU = 2 * nch - 1;  // ring uses

  uint64_t* rbar = reinterpret_cast<uint64_t*>(smem_raw);  // [NS] ring stage landed
  uint64_t* dbar = rbar + NS;                              // [2][NS] IO: the forward / backward sweep is done with the stage
  uint64_t* pbar = dbar + 2 * NS;                          // [nch] pivot-plane chunk landed
  constexpr size_t AL = CM ? 1024 : 128;  // swizzled chunks sit on 1 KB boundaries
  const size_t hs_off = ((size_t)(3 * NS + nch) * 8 + 15) & ~(size_t)15;
  V* HS = reinterpret_cast<V*>(smem_raw + hs_off);  // PP: chain state handed from one sweeping warp to the other [3][32]
  const size_t off = (hs_off + (PP ? 3 * 32 * sizeof(V) : 0) + AL - 1) & ~(AL - 1);
  V* P0 = reinterpret_cast<V*>(smem_raw + off);  // d -> d'
  V* RG = P0 + (size_t)nch * CH;                 // [NS][3][LB][32]
  auto hand_wait = [&]() { asm volatile("bar.sync %0, 64;" ::"r"(1 + cw) : "memory"); };       // state for this warp is there
  auto hand_over = [&]() { asm volatile("bar.arrive %0, 64;" ::"r"(2 - cw) : "memory"); };     // state for the other warp is there

  const int c0 = blockIdx.x * W;
  const int row0 = blockIdx.y * nlev;
  if (threadIdx.x == 0) {
    for (int s = 0; s < 3 * NS + nch; ++s) mbar_init(&rbar[s], 1);
    fence_mbar_init();
  }
  if (IO) __syncthreads();
  else __syncwarp();
  const uint64_t pol_stream = l2_policy_evict_first(), pol_keep = l2_policy_evict_last();

  auto load_chunk = [&](void* dst, const CUtensorMap* m, int k, uint64_t* bar, uint64_t pol) {
    if (CM) tma_load_2d_hint(dst, m, k * LB, c0, bar, pol);
    else tma_load_2d_hint(dst, m, c0, row0 + k * LB, bar, pol);
  };
  auto store_chunk = [&](int k, const void* src, uint64_t pol) {
    if (CM) tma_store_2d_hint(&mx, k * LB, c0, src, pol);
    else tma_store_2d_hint(&mx, c0, row0 + k * LB, src, pol);
    tma_store_commit();
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
      const int k = 2 * nch - 2 - u;
      mbar_expect_tx(&rbar[s], 2u * CHUNK_BYTES);
      load_chunk(st + CH, &mdu, k, &rbar[s], pol_stream);
      load_chunk(st + 2 * CH, &mx, k, &rbar[s], pol_stream);  // x', written by this CTA during the forward sweep
    }
  };
  auto issue_plane = [&](int k) {
    mbar_expect_tx(&pbar[k], CHUNK_BYTES);
    load_chunk(P0 + (size_t)k * CH, &md, k, &pbar[k], pol_stream);
  };
  auto ld = [&](const V* chunk, int u) { return t_ld<R, CPT, LB, CM>(chunk, u, lane); };
  if (io_warp) {
    // ------------------------------------------------------------ io warp (IO): one thread issues every copy.
    // A tensor copy costs its issuing thread ~130 cycles (measured); five per forward chunk and three per backward
    // chunk on the sweeping warp were a third of a tile's time. Here the sweeping warp only arrives on dbar[stage]
    // when a chunk's results are in shared memory (fenced for the async proxy); this thread stores the chunk, waits
    // until the store has read it, and refills the stage with the use NS ahead.
    if (lane != 0) return;
    for (int k = 0; k < pd && k < nch; ++k) issue_plane(k);
    for (int u = 0; u < NS && u < nch; ++u) issue_use(u);
    // Arrivals of the forward and of the backward sweep go to separate barriers: the top chunk's stage is finished
    // twice in a row (forward, then backward) and two arrivals on one barrier could flip its phase twice unseen.
    unsigned dph = 0;  // phase bit per barrier
    auto wait_done = [&](int s) {
      mbar_wait(&dbar[s], (dph >> s) & 1u);
      dph ^= 1u << s;
    };
    for (int k = 0; k < nch; ++k) {
      const int s = k % NS;
      wait_done(s);
      if (k + pd < nch) issue_plane(k + pd);
      if (k < nch - 1) {  // the top chunk stays where it is
        store_chunk(k, RG + (size_t)s * 3 * CH + 2 * CH, pol_keep);
        if (k + NS < nch) {
          tma_store_wait_read<0>();
          issue_use(k + NS);
        }
      }
    }
    tma_store_wait_all<0>();  // turn-around: every x' must have reached L2 before it is read back
    for (int u = nch; u < NU && u < nch + NS - 1; ++u) issue_use(u);
    for (int jb = 0; jb < nch; ++jb) {
      const int k = nch - 1 - jb, u0 = nch - 1 + jb, s = u0 % NS;
      wait_done(NS + s);
      if (CM || k != nfull) store_chunk(k, RG + (size_t)s * 3 * CH + 2 * CH, pol_stream);
      if (u0 + NS < NU) {
        tma_store_wait_read<0>();
        issue_use(u0 + NS);
      }
    }
    tma_store_wait_read<0>();  // shared memory must outlive the stores' reads
    return;
  }
  if (!IO && lane == 0) {
    for (int k = 0; k < pd && k < nch; ++k) issue_plane(k);
    for (int u = 0; u < NS && u < nch; ++u) issue_use(u);
  }

  // ---------------------------------------------------------------- forward sweep
  R dprev[CPT], xprev[CPT], uprev[CPT];
#pragma unroll
  for (int j = 0; j < CPT; ++j) {
    dprev[j] = (R)1;
    xprev[j] = (R)0;
    uprev[j] = (R)0;
  }
  for (int k = PP ? cw : 0; k < nch; k += NCW) {
    const int s = k % NS;
    mbar_wait(&rbar[s], (unsigned)(k / NS) & 1u);
    mbar_wait(&pbar[k], 0u);
    V* st = RG + (size_t)s * 3 * CH;
    V* p0 = P0 + (size_t)k * CH;
    V a[LB], b[LB], c[LB], x[LB], o0[LB], o1[LB];
#pragma unroll
    for (int u = 0; u < LB; ++u) {
      a[u] = ld(st, u);
      c[u] = ld(st + CH, u);
      x[u] = ld(st + 2 * CH, u);
      b[u] = ld(p0, u);
    }
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
    bfb_fwd_chunk<R, CPT, LB, false>(a, b, c, x, o0, o1, dprev, xprev, uprev, k == 0, bad);
    if (__any_sync(0xffffffffu, bad)) {  // rare operands somewhere in the warp: redo with the full-range division
#pragma unroll
      for (int j = 0; j < CPT; ++j) {
        dprev[j] = sd[j];
        xprev[j] = sx[j];
        uprev[j] = su[j];
      }
      bfb_fwd_chunk<R, CPT, LB, true>(a, b, c, x, o0, o1, dprev, xprev, uprev, k == 0, bad);
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
    for (int u = 0; u < LB; ++u) {
      t_st<R, CPT, LB, CM>(p0, u, lane, o0[u]);
      t_st<R, CPT, LB, CM>(st + 2 * CH, u, lane, o1[u]);
    }
    if (PP && k == nch - 1) hand_over();  // turn-around: the top chunk's d' and x' must be in shared memory first
    const bool spill = k < nch - 1;  // the top chunk stays where it is
    if (spill) fence_proxy_async();
    __syncwarp();
    if (IO) {
      if (lane == 0) mbar_arrive(&dbar[s]);
    } else if (lane == 0) {
      if (spill) store_chunk(k, st + 2 * CH, pol_keep);
      if (k >= 1 && k - 1 + NS < nch) {  // the stage of chunk k-1 is free once its store has read it
        if (spill) tma_store_wait_read<1>();
        else tma_store_wait_read<0>();
        issue_use(k - 1 + NS);
      }
      if (k + pd < nch) issue_plane(k + pd);
    }
  }
  // turn-around: every x' must have reached L2 before it is read back
  if (!IO && lane == 0) {
    tma_store_wait_all<0>();
    for (int u = nch; u < NU && u < nch + NS - 1; ++u) issue_use(u);
  }

  // ---------------------------------------------------------------- back substitution
  R xv[CPT];
#pragma unroll
  for (int j = 0; j < CPT; ++j) xv[j] = (R)0;
  // PP: the chunks keep alternating between the two warps across the turn-around (step nch + jb belongs to warp
  // (nch + jb) % 2). What a step reads was written by the other warp -- the pivots of chunk k always, x' of the top
  // chunk -- before a hand_over that precedes this step's hand_wait; only for jb <= 1 is that hand_over the one
  // this very step waits for, so those two steps wait before they load and all later ones load first.
  auto take_x = [&]() {
    hand_wait();
    const V h0 = HS[lane];
#pragma unroll
    for (int j = 0; j < CPT; ++j) xv[j] = h0.v[j];
  };
  for (int jb = PP ? ((nch + cw) & 1) : 0; jb < nch; jb += NCW) {
    const int k = nch - 1 - jb, u0 = nch - 1 + jb, s = u0 % NS;
    if (jb > 0) mbar_wait(&rbar[s], (unsigned)(u0 / NS) & 1u);
    V* st = RG + (size_t)s * 3 * CH;
    const V* bu = st + CH;
    V* xs = st + 2 * CH;
    const V* q0p = P0 + (size_t)k * CH;
    bool stored = true;
    if (PP && jb == 0) hand_wait();  // xv stays 0
    if (PP && jb == 1) take_x();
    if (k == nfull) {
      // partial top chunk (ntail levels)
      const int64_t col = (int64_t)c0 + lane * CPT;
      for (int u = ntail - 1; u >= 0; --u) {
        const V y = ld(xs, u), m0 = ld(q0p, u);
        V ui = ld(bu, u), o;
        if (u == ntail - 1) {  // du(nrow-1) is never used (:24-27)
#pragma unroll
          for (int j = 0; j < CPT; ++j) ui.v[j] = (R)0;
        }
#pragma unroll
        for (int j = 0; j < CPT; ++j) {
          xv[j] = x_div(x_sub(y.v[j], x_mul(ui.v[j], xv[j])), m0.v[j]);  // :317-319
          o.v[j] = xv[j];
        }
        if (CM) {
          t_st<R, CPT, LB, CM>(xs, u, lane, o);
        } else {  // a tensor store would spill into the next matrix' rows -> plain streaming stores
          R* g = X + (size_t)(row0 + k * LB + u) * x_stride + col;
#pragma unroll
          for (int j = 0; j < CPT; ++j)
            if (col + j < ncol) __stcs(g + j, o.v[j]);
        }
      }
      if (CM) fence_proxy_async();
      else stored = false;
    } else {
      V y[LB], m0[LB], ui[LB];
#pragma unroll
      for (int u = 0; u < LB; ++u) {
        y[u] = ld(xs, u);
        m0[u] = ld(q0p, u);
        ui[u] = ld(bu, u);
      }
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
      if (PP && jb < nch - 1) {
        V h0;
#pragma unroll
        for (int j = 0; j < CPT; ++j) h0.v[j] = xv[j];
        HS[lane] = h0;
        hand_over();
      }
#pragma unroll
      for (int u = 0; u < LB; ++u) t_st<R, CPT, LB, CM>(xs, u, lane, y[u]);
      fence_proxy_async();  // generic-proxy writes of this chunk -> visible to the TMA store
    }
    if (PP && k == nfull && jb < nch - 1) {  // after the ragged top chunk
      V h0;
#pragma unroll
      for (int j = 0; j < CPT; ++j) h0.v[j] = xv[j];
      HS[lane] = h0;
      hand_over();
    }
    __syncwarp();
    if (IO) {
      if (lane == 0) mbar_arrive(&dbar[NS + s]);
    } else if (lane == 0) {
      if (stored) store_chunk(k, xs, pol_stream);
      if (jb >= 1 && u0 - 1 + NS < NU) {  // the stage of the previous step is free once its store has read it
        if (stored) tma_store_wait_read<1>();
        else tma_store_wait_read<0>();
        issue_use(u0 - 1 + NS);
      }
    }
  }
  if (!IO && lane == 0) tma_store_wait_read<0>();  // shared memory must outlive the stores' reads
}

// ------------------------------------------------------------------------------------------- host side

template <typename R, bool EXACT, int NS, bool CM, bool IO = false>
static int launch_tma(R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nmat, int64_t a_stride, int64_t x_stride,
                      int pd, size_t smem, cudaStream_t stream, const char* name) {
  constexpr int CPT = TmaCfg<R>::CPT, LB = TmaGeom<R, CM>::LB, RN = EXACT ? 1 : TmaGeom<R, CM>::RN, W = 32 * CPT;
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
  auto k = k_tridiag_tma<R, EXACT, CPT, RN, NS, LB, CM, IO>;
  int rc = allow_smem(k, smem);
  if (rc) return rc;
  dim3 grid((unsigned)((ncol + W - 1) / W), (unsigned)nmat);
  k<<<grid, IO ? 64 : 32, smem, stream>>>(mdl, md, mdu, mx, X, ncol, nlev, x_stride, pd);
  note_launch(name);
  return check_cuda(cudaGetLastError());
}

template <typename R, int NS, bool CM, int IO>
static int launch_spill(R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nmat, int64_t a_stride, int64_t x_stride,
                        int pd, size_t smem, cudaStream_t stream, const char* name) {
  constexpr int CPT = TmaCfg<R>::CPT, LB = TmaGeom<R, CM>::LB, W = 32 * CPT;
  CUtensorMap mdl, md, mdu, mx;
  bool ok;
  const int swz = CM && LB * sizeof(R) == 64 ? 1 : 0;
  if (CM) {
    ok = make_tensor_map_2d<R>(&mdl, dl, nlev, ncol, a_stride, LB, W, swz) && make_tensor_map_2d<R>(&md, d, nlev, ncol, a_stride, LB, W, swz) &&
         make_tensor_map_2d<R>(&mdu, du, nlev, ncol, a_stride, LB, W, swz) && make_tensor_map_2d<R>(&mx, X, nlev, ncol, x_stride, LB, W, swz);
  } else {
    const int64_t rows = (int64_t)nmat * nlev;
    ok = make_tensor_map_2d<R>(&mdl, dl, ncol, rows, a_stride, W, LB, 0) && make_tensor_map_2d<R>(&md, d, ncol, rows, a_stride, W, LB, 0) &&
         make_tensor_map_2d<R>(&mdu, du, ncol, rows, a_stride, W, LB, 0) && make_tensor_map_2d<R>(&mx, X, ncol, rows, x_stride, W, LB, 0);
  }
  if (!ok) return EKAT_B200_ERR_UNSUPPORTED;
  auto k = k_tridiag_bfb_spill<R, CPT, NS, LB, CM, IO>;
  int rc = allow_smem(k, smem);
  if (rc) return rc;
  dim3 grid((unsigned)((ncol + W - 1) / W), (unsigned)nmat);
  k<<<grid, IO == 2 ? 96 : IO ? 64 : 32, smem, stream>>>(mdl, md, mdu, mx, X, ncol, nlev, x_stride, pd);
  note_launch(name);
  return check_cuda(cudaGetLastError());
}

// Returns EKAT_B200_ERR_UNSUPPORTED when the operands do not meet TMA's requirements (16-byte aligned bases and
// pitches, column fits on chip, sizes within tensor-map limits); the caller then uses the cp.async / tile kernels.
template <typename R, bool CM>
static int run_tma(bool exact, R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nmat, int64_t a_stride,
                   int64_t x_stride, cudaStream_t stream) {
  auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  if (!al(dl) || !al(d) || !al(du) || !al(X) || (a_stride * sizeof(R)) % 16 || (x_stride * sizeof(R)) % 16)
    return EKAT_B200_ERR_UNSUPPORTED;
  if (ncol >= (int64_t)1 << 31 || (int64_t)nmat * nlev >= (int64_t)1 << 31 || nmat > 65535) return EKAT_B200_ERR_UNSUPPORTED;
  if (CM && nmat != 1) return EKAT_B200_ERR_UNSUPPORTED;  // several right-hand sides interleave within a column there
  constexpr int CPT = TmaCfg<R>::CPT, W = 32 * CPT, LB = TmaGeom<R, CM>::LB;
  const Tuning tu = tuning();
  const int nch = (nlev + LB - 1) / LB;
  const size_t chunk = (size_t)LB * W * sizeof(R);
  const bool f64 = sizeof(R) == 8;
  // bfb from 96 levels up: input-only copy ring, two sweeping warps (tridiag_bfb_stream.cu). Shorter columns leave room
  // for five or more tiles of the one-sweeping-warp kernel below, which then wins (measured at 72 levels: 79% against
  // 71%; at 96: 74% against 76%; at 128: 66% against 75-82%); so it does for column-major f32 (12-level chunks).
  if (exact && (tu.variant == kVarAuto || tu.variant == kVarStreamOneWarp) && (nlev >= 96 || tu.variant == kVarStreamOneWarp) && !(CM && sizeof(R) == 4)) {
    const int rc = tridiag_bfb_stream<R, CM>(dl, d, du, X, ncol, nlev, nmat, a_stride, x_stride, stream);
    if (rc != EKAT_B200_ERR_UNSUPPORTED) return rc;
  }
  if (exact && tu.variant != kVarWarpTileBfb) {
    // bfb: pivots on chip, x' through L2 (k_tridiag_bfb_spill): as many tiles per SM as a 3- or 4-stage ring allows
    const size_t al = CM ? 1024 : 128;
    auto smem_for = [&](int stages) {
      const size_t hs = (((size_t)(3 * stages + nch) * 8 + 15) & ~(size_t)15) + 3 * 32 * sizeof(R) * CPT;  // barriers + hand-over state
      return ((hs + al - 1) & ~(al - 1)) + (size_t)nch * chunk + (size_t)stages * 3 * chunk;
    };
    auto tiles = [&](int stages) { return (int)((kMaxSmemPerCta + 1024) / (smem_for(stages) + 1024)); };
    // three stages (two chunks in flight behind the one being swept) are what the re-loads of the back-substitution
    // need; a fourth only if it costs no residency; two only if nothing else fits (measured: 2 stages 49-51%,
    // 3 stages 57-62%, 4 stages at one tile fewer 46%)
    int ns = tiles(4) == tiles(3) ? 4 : 3;
    if (smem_for(ns) > kMaxSmemPerCta) ns = 2;
    if (tu.cols_per_cta >= 2 && tu.cols_per_cta <= 4) ns = tu.cols_per_cta;
    if (smem_for(ns) <= kMaxSmemPerCta) {
      const int pd = tu.prefetch > 0 ? tu.prefetch : 3;
      const char* name = CM ? (f64 ? "tridiag_bfb_spill_cm<f64>" : "tridiag_bfb_spill_cm<f32>")
                            : (f64 ? "tridiag_bfb_spill<f64>" : "tridiag_bfb_spill<f32>");
      // kVarSpillOneWarp: the single-warp form (the sweeping warp issues its own copies), kVarSpillTwoWarps: two sweeping warps + io
      // warp; kept for A/B measurements. Default: one sweeping warp + io warp.
#define EKB_SPILL(NS_, IO_) launch_spill<R, NS_, CM, IO_>(dl, d, du, X, ncol, nlev, nmat, a_stride, x_stride, pd, smem_for(NS_), stream, name)
      if (tu.variant == kVarSpillOneWarp) return ns == 4 ? EKB_SPILL(4, 0) : ns == 3 ? EKB_SPILL(3, 0) : EKB_SPILL(2, 0);
      if (tu.variant == kVarSpillTwoWarps) return ns == 4 ? EKB_SPILL(4, 2) : ns == 3 ? EKB_SPILL(3, 2) : EKB_SPILL(2, 2);
      return ns == 4 ? EKB_SPILL(4, 1) : ns == 3 ? EKB_SPILL(3, 1) : EKB_SPILL(2, 1);
#undef EKB_SPILL
    }
  }
  const size_t bars = ((size_t)(12 + 2 * nch) * 8 + 1023) & ~(size_t)1023;
  const size_t planes = (size_t)2 * nch * chunk;
  const int nrings = 2;
  // ring depth: 4 stages, or 2 if that lets one more CTA share the SM (measured: residency wins over ring depth)
  auto per_sm = [&](int stages) { return (int)((kMaxSmemPerCta + 1024) / (bars + planes + (size_t)nrings * stages * chunk + 1024)); };
  int ns = per_sm(2) > per_sm(4) ? 2 : 4;
  if (tu.cols_per_cta == 2 || tu.cols_per_cta == 4) ns = tu.cols_per_cta;  // tuning override: ring stages
  const size_t smem = bars + planes + (size_t)nrings * ns * chunk;
  if (smem > kMaxSmemPerCta) return EKAT_B200_ERR_UNSUPPORTED;
  int pd = tu.prefetch > 0 ? tu.prefetch : (CM ? 4 : 3);  // plane chunks in flight ahead of the sweep
  if (pd < 1) pd = 1;
  const char* name = CM ? (exact ? (f64 ? "tridiag_tma_cm<f64,bfb>" : "tridiag_tma_cm<f32,bfb>")
                                 : (f64 ? "tridiag_tma_cm<f64,thomas>" : "tridiag_tma_cm<f32,thomas>"))
                        : (exact ? (f64 ? "tridiag_tma<f64,bfb>" : "tridiag_tma<f32,bfb>")
                                 : (f64 ? "tridiag_tma<f64,thomas>" : "tridiag_tma<f32,thomas>"));
  if (exact)
    return ns == 2 ? launch_tma<R, true, 2, CM>(dl, d, du, X, ncol, nlev, nmat, a_stride, x_stride, pd, smem, stream, name)
                   : launch_tma<R, true, 4, CM>(dl, d, du, X, ncol, nlev, nmat, a_stride, x_stride, pd, smem, stream, name);
  if (tu.variant == kVarWarpTileIo)  // io-warp form: measured slower here (f32 69% against 72%; f64 with one matrix per rhs 76% against
                         // 78%) -- unlike bfb, whose long chain leaves the sweeping warp no slack; kept for A/B measurements
    return ns == 2 ? launch_tma<R, false, 2, CM, true>(dl, d, du, X, ncol, nlev, nmat, a_stride, x_stride, pd, smem, stream, name)
                   : launch_tma<R, false, 4, CM, true>(dl, d, du, X, ncol, nlev, nmat, a_stride, x_stride, pd, smem, stream, name);
  return ns == 2 ? launch_tma<R, false, 2, CM>(dl, d, du, X, ncol, nlev, nmat, a_stride, x_stride, pd, smem, stream, name)
                 : launch_tma<R, false, 4, CM>(dl, d, du, X, ncol, nlev, nmat, a_stride, x_stride, pd, smem, stream, name);
}

// level-major operands: dl,d,du [nmat][nlev][a_stride], X [nmat][nlev][x_stride], columns fastest
template <typename R>
int tridiag_tma(bool exact, R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nmat, int64_t a_stride,
                int64_t x_stride, cudaStream_t stream) {
  return run_tma<R, false>(exact, dl, d, du, X, ncol, nlev, nmat, a_stride, x_stride, stream);
}
// the reference's column-major operands, one matrix and one right-hand side per column: dl,d,du [ncol][a_stride],
// X [ncol][x_stride], levels fastest
template <typename R>
int tridiag_tma_cm(bool exact, R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int64_t a_stride, int64_t x_stride,
                   cudaStream_t stream) {
  return run_tma<R, true>(exact, dl, d, du, X, ncol, nlev, 1, a_stride, x_stride, stream);
}

template int tridiag_tma<double>(bool, double*, double*, double*, double*, int64_t, int, int, int64_t, int64_t, cudaStream_t);
template int tridiag_tma<float>(bool, float*, float*, float*, float*, int64_t, int, int, int64_t, int64_t, cudaStream_t);
template int tridiag_tma_cm<double>(bool, double*, double*, double*, double*, int64_t, int, int64_t, int64_t, cudaStream_t);
template int tridiag_tma_cm<float>(bool, float*, float*, float*, float*, int64_t, int, int64_t, int64_t, cudaStream_t);

}  // namespace ekb
