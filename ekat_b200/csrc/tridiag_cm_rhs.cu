// tridiag_cm_rhs.cu -- Thomas / bfb on the REFERENCE layout with several right-hand sides per column:
// A[col][3][lev][nA], X[col][lev][nrhs] (tests/algorithm/tridiag_tests_performance.cpp:181-185), nA == 1 (one matrix
// per column shared by its right-hand sides: BASELINE config 4 as a reference caller holds it, and the reference perf
// driver's own default of 43 right-hand sides, :33-37) or nA == nrhs (one matrix per right-hand side).
//
// Replaces the per-column loops around ekat::tridiag::thomas / bfb on X[nrow][nrhs] (reference
// src/algorithm/ekat_tridiag.hpp:177-214, :229-290, :292-320). In this layout everything a column owns is
// contiguous, and the right-hand sides of a level are neighbours in memory. So:
//   * a CTA takes a tile of C whole columns; each of their rows (dl, d, du: nlev*nA words; X: nlev*nrhs words)
//     arrives by one 1D bulk TMA copy into shared memory and the solution leaves the same way -- HBM sees only
//     long contiguous bursts and exactly the algorithmic bytes;
//   * a solver thread is (column, right-hand side): the threads of a warp read nrhs consecutive words per column
//     (the pitch between columns is padded so that a warp's columns fall on different banks), and the
//     multipliers of a column are a broadcast;
//   * with one matrix per column, one lane of a dedicated factor warp per column runs the factorisation a chunk of 8
//     levels ahead of the solver threads (one mbarrier per chunk, arrive = release, wait = acquire), so the division chain is
//     computed once per column, not once per right-hand side, and overlaps the forward sweeps;
//   * a tile is small (as few columns as give ~64 solver threads), so four or more CTAs share an SM and one
//     tile's copies and chains overlap the others'.
// thomas: ratio-form factorisation (no division on the dependent chain; tridiag_tma_shared.cu) and one FMA per level
// on the solver chains -- agreement to the tolerance; bfb: the reference's operation order with separately rounded
// * - / -- bit-identical. EKAT_B200_CLOBBER_DIAGS: bfb leaves multipliers in dl(1..) and pivots in d as the
// reference does (:292-303); thomas then runs the bfb sweep and writes back d only (serial thomas, :216-227).
#include "thomas_two_sided.cuh"

namespace ekb {

__device__ __forceinline__ void tma_store_1d(void* gdst, const void* ssrc, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)), "r"(bytes) : "memory");
}

struct CmArgs {
  void *dl, *d, *du, *X;
  int64_t ncol, a_stride, x_stride;
  int nlev, nA, nrhs, C;
  int PP;   // elements between the three diagonal planes of a column in shared memory
  int APC;  // elements between the columns' diagonal blocks
  int XP;   // elements between the columns' right-hand-side blocks
  int clobber;  // bit 0: write dl back, bit 1: write d back
};

constexpr int kCmLB = 8;  // levels per hand-over from the factor lane to the solver threads

template <typename R, bool EXACT>
__global__ void __launch_bounds__(512)
k_tridiag_cm_rhs(const CmArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem_raw);  // the tile has landed
  uint64_t* cb = bar + 1;                                  // [nch] chunk k of every column of the tile is factorised
  const int nlev = a.nlev, nA = a.nA, nrhs = a.nrhs, C = a.C;
  const int nch = (nlev + kCmLB - 1) / kCmLB;
  R* SA = reinterpret_cast<R*>(smem_raw + ((8 * (size_t)(1 + nch) + 127) & ~(size_t)127));
  R* SX = SA + (size_t)C * a.APC;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t c0 = (int64_t)blockIdx.x * C;
  const int nc = (int)tmin<int64_t>(C, a.ncol - c0);
  const unsigned abytes = (unsigned)((size_t)nlev * nA * sizeof(R)), xbytes = (unsigned)((size_t)nlev * nrhs * sizeof(R));
  if (tid == 0) {
    mbar_init(bar, 1);
    if (nA == 1)
      for (int k = 0; k < nch; ++k) mbar_init(&cb[k], 1);
    fence_mbar_init();
    mbar_expect_tx(bar, (unsigned)nc * (3u * abytes + xbytes));
    for (int c = 0; c < nc; ++c) {
      R* sa = SA + (size_t)c * a.APC;
      tma_load_1d(sa, static_cast<const R*>(a.dl) + (c0 + c) * a.a_stride, abytes, bar);
      tma_load_1d(sa + a.PP, static_cast<const R*>(a.d) + (c0 + c) * a.a_stride, abytes, bar);
      tma_load_1d(sa + 2 * a.PP, static_cast<const R*>(a.du) + (c0 + c) * a.a_stride, abytes, bar);
      tma_load_1d(SX + (size_t)c * a.XP, static_cast<const R*>(a.X) + (c0 + c) * a.x_stride, xbytes, bar);
    }
  }
  __syncthreads();
  mbar_wait(bar, 0u);

  if (warp == 0) {
    // ================================================================ factor warp: lane = column (one matrix per column).
    // Lanes beyond the tile's columns repeat the last column (same reads, same writes, same values), so that the
    // warp stays converged for the hand-over: one mbarrier arrival per chunk of 8 levels releases that chunk of
    // every column to the solver threads (release / acquire: no separate fence).
    if (nA == 1) {
      R* pl = SA + (size_t)tmin(lane, nc - 1) * a.APC;
      R* pd = pl + a.PP;
      R* pu = pd + a.PP;
      auto hand_over = [&](int k) {
        __syncwarp();
        if (lane == 0) mbar_arrive(&cb[k]);
      };
      if (EXACT) {  // :292-303  dl(i) /= d(i-1); d(i) -= dl(i)*du(i-1)
        R dprev = pd[0], uprev = pu[0];
        for (int k = 0; k < nch; ++k) {
          const int i0 = k * kCmLB, n = tmin(kCmLB, nlev - i0);
          R va[kCmLB], vb[kCmLB], vc[kCmLB];
#pragma unroll
          for (int u = 0; u < kCmLB; ++u) {
            const int i = u < n ? i0 + u : i0;
            va[u] = pl[i];
            vb[u] = pd[i];
            vc[u] = pu[i];
          }
          // the reference's operations, the quotient speculatively (no branch on the chain); a chunk with operands out
          // of the speculative division's range is repeated with the compiler's full-range division
          const R sd = dprev, su = uprev;
          R om[kCmLB], od[kCmLB];
          bool bad = false;
          R rp = rcp_seed(dprev);  // ~ 1 / d'(i-1): reciprocal of the PREDICTED pivot, off the chain (common.cuh bfb_pivot_step)
#pragma unroll
          for (int u = 0; u < kCmLB; ++u) {
            bool f = false;
            R dn = dprev, rn = rp;
            const R m = bfb_factor_level(va[u], vb[u], uprev, dn, rn, f);
            bad = bad || (f && i0 + u > 0 && u < n);
            if (i0 + u > 0) {
              dprev = dn;
              rp = rn;
              uprev = vc[u];
            }
            om[u] = m;
            od[u] = dprev;
          }
          if (__any_sync(0xffffffffu, bad)) {
            dprev = sd;
            uprev = su;
#pragma unroll
            for (int u = 0; u < kCmLB; ++u) {
              const R m = x_div(va[u], dprev);
              if (i0 + u > 0) {
                dprev = x_sub(vb[u], x_mul(m, uprev));
                uprev = vc[u];
              }
              om[u] = m;
              od[u] = dprev;
            }
          }
#pragma unroll
          for (int u = 0; u < kCmLB; ++u) {
            if (u < n && i0 + u > 0) {
              pl[i0 + u] = om[u];
              pd[i0 + u] = od[u];
            }
          }
          hand_over(k);
        }
      } else {
        // thomas: ratio form (no division on the dependent chain), FROM BOTH ENDS -- the factor lane runs the downward
        // recurrence of the top half and the upward one of the bottom half interleaved (thomas_two_sided.cuh), one
        // hand-over per step = chunk t from the top and chunk nch-1-t from the bottom
        if (nlev % kCmLB == 0) two_sided_factor<R, true>(pl, pd, pu, 1, nlev, hand_over);
        else two_sided_factor<R, false>(pl, pd, pu, 1, nlev, hand_over);
      }
    }
  } else {
    // ================================================================ solver threads: (column, right-hand side)
    const int t = tid - 32, c = t / nrhs, j = t - c * nrhs;
    if (c < nc) {
      R* px = SX + (size_t)c * a.XP + j;
      const int ja = nA == 1 ? 0 : j;
      R* pl = SA + (size_t)c * a.APC + ja;
      R* pd = pl + a.PP;
      R* pu = pd + a.PP;
      const int sa = nA, sx = nrhs;  // element strides between levels
      if (nA == 1 && !EXACT) {
        // both sweeps from both ends, a step behind the factor lane of this column
        auto wait_step = [&](int t) { mbar_wait(&cb[t], 0u); };
        if (nlev % kCmLB == 0) two_sided_sweeps<R, true>(pl, pd, pu, 1, px, sx, nlev, wait_step);
        else two_sided_sweeps<R, false>(pl, pd, pu, 1, px, sx, nlev, wait_step);
      } else if (nA == 1) {
        // forward sweep a chunk behind the factor lane of this column
        R z = 0;
        for (int i0 = 0; i0 < nlev; i0 += kCmLB) {
          const int n = tmin(kCmLB, nlev - i0);
          mbar_wait(&cb[i0 / kCmLB], 0u);
          R vm[kCmLB], vr[kCmLB], vx[kCmLB];
#pragma unroll
          for (int u = 0; u < kCmLB; ++u) {
            const int i = u < n ? i0 + u : i0;
            vm[u] = pl[i];
            if (!EXACT) vr[u] = pd[i];
            vx[u] = px[i * sx];
          }
          if (EXACT && i0 == 0) vm[0] = (R)0;  // level 0 has no multiplier (dl(0) holds whatever the caller left there)
#pragma unroll
          for (int u = 0; u < kCmLB; ++u) {
            if (u < n) {
              if (EXACT) {
                z = (i0 + u == 0) ? vx[u] : x_sub(vx[u], x_mul(vm[u], z));  // X(i) -= dl(i)*X(i-1)   :315-316
                px[(i0 + u) * sx] = z;
              } else {
                z = f_fma(-vm[u], z, vx[u]);
                px[(i0 + u) * sx] = z * vr[u];
              }
            }
          }
        }
      } else {
        // one matrix per right-hand side: this thread factorises its own (:254-290 / :650-678), level by level
        if (EXACT) {
          R dprev = pd[0], uprev = pu[0], z = px[0];
          for (int i = 1; i < nlev; ++i) {
            const R m = x_div_pipelined(pl[i * sa], dprev);
            dprev = x_sub(pd[i * sa], x_mul(m, uprev));
            uprev = pu[i * sa];
            z = x_sub(px[i * sx], x_mul(m, z));
            pl[i * sa] = m;
            pd[i * sa] = dprev;
            px[i * sx] = z;
          }
        } else {
          constexpr int RN = TmaCfg<R>::RN;
          R q1 = 1, q2 = 0, cp = 0, gp = 0, z = 0;
          for (int i0 = 0; i0 < nlev; i0 += kCmLB) {
            const int n = tmin(kCmLB, nlev - i0);
            R va[kCmLB], vb[kCmLB], vc[kCmLB], vx[kCmLB];
#pragma unroll
            for (int u = 0; u < kCmLB; ++u) {
              const int i = u < n ? i0 + u : i0;
              va[u] = pl[i * sa];
              vb[u] = pd[i * sa];
              vc[u] = pu[i * sa];
              vx[u] = px[i * sx];
            }
            if (i0 == 0) va[0] = (R)0;
#pragma unroll
            for (int u = 0; u < kCmLB; ++u) {
              if (u < n) {
                R q;
                if (u % RN == 0) q = f_fma(-(va[u] * cp), q2, vb[u]);
                else q = f_fma(vb[u], q1, -((va[u] * cp) * q2));
                const R rq = t_rcp(q);
                const R g = (u % RN == 0) ? rq : q1 * rq;
                z = f_fma(-(va[u] * gp), z, vx[u]);
                px[(i0 + u) * sx] = z * g;
                pu[(i0 + u) * sa] = (i0 + u == nlev - 1) ? (R)0 : vc[u] * g;
                if ((u + 1) % RN == 0) {
                  q2 = g;
                  q1 = (R)1;
                } else {
                  q2 = (u % RN == 0) ? (R)1 : q1;
                  q1 = q;
                }
                cp = vc[u];
                gp = g;
              }
            }
          }
        }
      }
      // back substitution (the two-sided path has done its own)
      if (nA == 1 && !EXACT) {
      } else if (EXACT) {  // :317-319  X(n-1) /= d(n-1); X(i) = (X(i) - du(i)*X(i+1)) / d(i)
        R xv = x_div(px[(nlev - 1) * sx], pd[(nlev - 1) * sa]);
        px[(nlev - 1) * sx] = xv;
        int i = nlev - 2;
        for (; i >= 3; i -= 4) {  // four levels at a time: the reciprocals of their pivots are off the chain
          R num[4], dd[4], uu[4], rr[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            num[u] = px[(i - u) * sx];
            dd[u] = pd[(i - u) * sa];
            uu[u] = pu[(i - u) * sa];
            rr[u] = rcp_refined(dd[u]);
          }
          const R sxv = xv;
          R o[4];
          bool bad = false;
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            xv = div_spec(x_sub(num[u], x_mul(uu[u], xv)), dd[u], rr[u], bad);
            o[u] = xv;
          }
          if (bad) {
            xv = sxv;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              xv = x_div(x_sub(num[u], x_mul(uu[u], xv)), dd[u]);
              o[u] = xv;
            }
          }
#pragma unroll
          for (int u = 0; u < 4; ++u) px[(i - u) * sx] = o[u];
        }
        for (; i >= 0; --i) {
          xv = x_div(x_sub(px[i * sx], x_mul(pu[i * sa], xv)), pd[i * sa]);
          px[i * sx] = xv;
        }
      } else {
        R xv = px[(nlev - 1) * sx];
        int i = nlev - 2;
        for (; i >= 7; i -= 8) {
          R vy[8], vc[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            vy[u] = px[(i - u) * sx];
            vc[u] = pu[(i - u) * sa];
          }
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            xv = f_fma(-vc[u], xv, vy[u]);
            px[(i - u) * sx] = xv;
          }
        }
        for (; i >= 0; --i) {
          xv = f_fma(-pu[i * sa], xv, px[i * sx]);
          px[i * sx] = xv;
        }
      }
    }
  }
  fence_proxy_async();  // this thread's shared-memory writes -> visible to the bulk stores
  __syncthreads();
  if (tid == 0) {
    for (int c = 0; c < nc; ++c) {
      tma_store_1d(static_cast<R*>(a.X) + (c0 + c) * a.x_stride, SX + (size_t)c * a.XP, xbytes);
      if (a.clobber & 1) tma_store_1d(static_cast<R*>(a.dl) + (c0 + c) * a.a_stride, SA + (size_t)c * a.APC, abytes);
      if (a.clobber & 2) tma_store_1d(static_cast<R*>(a.d) + (c0 + c) * a.a_stride, SA + (size_t)c * a.APC + a.PP, abytes);
    }
    tma_store_commit();
    tma_store_wait_read<0>();
  }
}

// Returns EKAT_B200_ERR_UNSUPPORTED when the rows do not meet the bulk copy's requirements (16-byte aligned bases,
// strides and row lengths) or a column does not fit on chip; the caller then uses the generic tile kernel.
template <typename R>
int tridiag_cm_rhs(bool exact, R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nA, int nrhs, int64_t a_stride,
                   int64_t x_stride, unsigned flags, cudaStream_t stream) {
  auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  const size_t abytes = (size_t)nlev * nA * sizeof(R), xbytes = (size_t)nlev * nrhs * sizeof(R);
  if (!al(dl) || !al(d) || !al(du) || !al(X) || (a_stride * sizeof(R)) % 16 || (x_stride * sizeof(R)) % 16 || abytes % 16 || xbytes % 16)
    return EKAT_B200_ERR_UNSUPPORTED;
  constexpr int W16 = 16 / sizeof(R), NBK = 128 / sizeof(R);  // elements per 16 bytes / per sweep over the 32 banks
  auto up = [](int v, int m) { return (v + m - 1) / m * m; };
  auto pad_to = [&](int len, int want) {  // smallest pitch >= len, multiple of W16, congruent to `want` modulo NBK
    int p = up(len, W16);
    while (p % NBK != want % NBK) p += W16;
    return p;
  };
  CmArgs a;
  a.PP = up(nlev * nA, W16);
  // columns of a warp must fall on different banks: the multipliers of neighbouring columns one 16-byte unit apart
  // (a broadcast per column), their right-hand sides one after the other
  a.APC = pad_to(3 * a.PP, nA == 1 ? W16 : up(nrhs, W16));
  a.XP = pad_to(nlev * nrhs, up(nrhs, W16));
  const size_t per_col = (size_t)(a.APC + a.XP) * sizeof(R);
  const size_t head = (8 * (size_t)(1 + (nlev + kCmLB - 1) / kCmLB) + 127) & ~(size_t)127;
  if (per_col + head > kMaxSmemPerCta) return EKAT_B200_ERR_UNSUPPORTED;
  if ((size_t)3 * abytes + xbytes >= ((size_t)1 << 20)) return EKAT_B200_ERR_UNSUPPORTED;  // mbarrier tx-count range
  const Tuning tu = tuning();
  // columns per tile: the factor warp's lanes are columns and its chain is what a tile waits for, so a tile takes as
  // many columns as half an SM's shared memory holds (two CTAs per SM overlap their copies with each other's
  // chains; measured on B200 at 128 levels: 10 rhs 2 / 4 / 8 / 16 columns 3.7 / 3.9 / 4.3 / 3.6 TB/s) -- but no more
  // than spreads the batch over two CTAs per SM
  int C = (int)tmax<size_t>(1, (kMaxSmemPerCta / 2 - head) / per_col);
  C = (int)tmin<int64_t>(C, tmax<int64_t>(1, ncol / (2 * kNumSms)));
  if (tu.cols_per_cta > 0) C = tu.cols_per_cta;
  C = tmin(C, 32);
  while (C > 1 && ((size_t)C * nrhs > 480 || (size_t)C * per_col + head > kMaxSmemPerCta ||
                   (size_t)C * (3 * abytes + xbytes) >= ((size_t)1 << 20)))
    --C;
  if ((size_t)C * nrhs > 480) return EKAT_B200_ERR_UNSUPPORTED;
  const bool clobber = (flags & EKAT_B200_CLOBBER_DIAGS) != 0;
  a.clobber = !clobber ? 0 : (exact ? 3 : 2);
  if (clobber) exact = true;  // thomas that must leave the pivots in d runs the bfb sweep (within its tolerance by construction)
  a.dl = dl; a.d = d; a.du = du; a.X = X;
  a.ncol = ncol; a.a_stride = a_stride; a.x_stride = x_stride;
  a.nlev = nlev; a.nA = nA; a.nrhs = nrhs; a.C = C;
  const size_t nch = (size_t)(nlev + kCmLB - 1) / kCmLB;
  const size_t smem = ((8 * (1 + nch) + 127) & ~(size_t)127) + (size_t)C * per_col;
  const int threads = 32 + (C * nrhs + 31) / 32 * 32;
  const unsigned grid = (unsigned)((ncol + C - 1) / C);
  int rc;
  const bool f64 = sizeof(R) == 8;
  if (exact) {
    auto k = k_tridiag_cm_rhs<R, true>;
    if ((rc = allow_smem(k, smem))) return rc;
    k<<<grid, threads, smem, stream>>>(a);
    note_launch(f64 ? "tridiag_cm_rhs<f64,bfb>" : "tridiag_cm_rhs<f32,bfb>");
  } else {
    auto k = k_tridiag_cm_rhs<R, false>;
    if ((rc = allow_smem(k, smem))) return rc;
    k<<<grid, threads, smem, stream>>>(a);
    note_launch(f64 ? "tridiag_cm_rhs<f64,thomas>" : "tridiag_cm_rhs<f32,thomas>");
  }
  return check_cuda(cudaGetLastError());
}

template int tridiag_cm_rhs<double>(bool, double*, double*, double*, double*, int64_t, int, int, int, int64_t, int64_t, unsigned, cudaStream_t);
template int tridiag_cm_rhs<float>(bool, float*, float*, float*, float*, int64_t, int, int, int, int64_t, int64_t, unsigned, cudaStream_t);

}  // namespace ekb
