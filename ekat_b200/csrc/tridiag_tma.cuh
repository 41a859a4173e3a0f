// tridiag_tma.cuh -- pieces shared by the TMA-fed level-major tridiagonal kernels
// (tridiag_tma.cu: one matrix per right-hand side; tridiag_tma_shared.cu: one matrix, many right-hand sides).
#pragma once
#include <cuda.h>

#include "tridiag_lm_common.cuh"

namespace ekb {

// [nrows][pitch] array of R, of which the first ncol entries of a row are data; box = boxc columns x boxr rows.
// Columns beyond ncol / rows beyond nrows read as zero and are not written (TMA bounds handling). tridiag_tma.cu.
template <typename R>
bool make_tensor_map_2d(CUtensorMap* m, const R* base, int64_t ncol, int64_t nrows, int64_t pitch, int boxc, int boxr,
                        int swizzle64 = 0);

// Reciprocal for the ratio form: hardware seed (~2^-20) and one cubic Newton step (~2^-60 relative; eps is 2^-53).
__device__ __forceinline__ double t_rcp(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  double e = fma(-x, r, 1.0);
  e = fma(e, e, e);
  return fma(r, e, r);
}
__device__ __forceinline__ float t_rcp(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  const float e = fmaf(-x, r, 1.0f);
  return fmaf(r, e, r);
}

// columns per lane and renormalisation interval of the ratio-form recurrence, per precision
template <typename R>
struct TmaCfg;
template <>
struct TmaCfg<double> {
  static constexpr int CPT = 1, RN = 8;
};
template <>
struct TmaCfg<float> {
  static constexpr int CPT = 2, RN = 2;
};

constexpr int kLB = 8;  // levels per TMA box / mbarrier phase / unrolled chunk

// bfb sweeps over one chunk of LB levels held in registers. SLOW = false: speculative IEEE division (div_spec: no
// branch on the dependent chain, out-of-range operands only raise `bad`); SLOW = true: the compiler's full-range
// division. The kernel runs the speculative form and repeats the chunk in SLOW form if any lane raised the flag.
// Element (level u of a chunk, lane, j-th column of the lane) of a shared-memory chunk of LB levels x W columns.
//   level-major operands (CM = false): the TMA box is LB rows of W columns -> [u][lane][j], one vector access;
//   the reference's column-major operands (CM = true: a column's levels are contiguous in memory,
//   tests/algorithm/tridiag_tests_performance.cpp:181-185): the box is W rows (columns) of LB levels -> [j*32+lane][u].
//   A level step of the warp then strides by the row length, which TMA wants in multiples of 16 bytes -- never an
//   odd number of words, so some bank conflict is the price of needing no transpose pass. Rows of 64 bytes with the
//   TMA's 64-byte swizzle (16-byte unit index XOR bits 7-8 of the address, i.e. XOR (row/2)%4) spread a level over
//   8 of the 16 bank pairs (f64: 4 wavefronts instead of 2); unswizzled 48-byte rows do the same for f32.
template <int BYTES>
__device__ __forceinline__ int cm_elem(int row, int u, int LB) {  // element index of (row, level u) in a chunk
  if (BYTES * LB == 64) {  // swizzled 64-byte rows
    const int per16 = 16 / BYTES, unit = (u / per16) ^ ((row >> 1) & 3);
    return row * LB + unit * per16 + u % per16;
  }
  return row * LB + u;
}
template <typename R, int CPT, int LB, bool CM>
__device__ __forceinline__ Vec<R, CPT> t_ld(const Vec<R, CPT>* chunk, int u, int lane) {
  if (!CM) return chunk[u * 32 + lane];
  Vec<R, CPT> v;
  const R* e = reinterpret_cast<const R*>(chunk);
#pragma unroll
  for (int j = 0; j < CPT; ++j) v.v[j] = e[cm_elem<sizeof(R)>(j * 32 + lane, u, LB)];
  return v;
}
template <typename R, int CPT, int LB, bool CM>
__device__ __forceinline__ void t_st(Vec<R, CPT>* chunk, int u, int lane, const Vec<R, CPT>& v) {
  if (!CM) {
    chunk[u * 32 + lane] = v;
    return;
  }
  R* e = reinterpret_cast<R*>(chunk);
#pragma unroll
  for (int j = 0; j < CPT; ++j) e[cm_elem<sizeof(R)>(j * 32 + lane, u, LB)] = v.v[j];
}

template <typename R, int CPT, int LB, bool SLOW>
__device__ __forceinline__ void bfb_fwd_chunk(const Vec<R, CPT> (&a)[LB], const Vec<R, CPT> (&b)[LB], const Vec<R, CPT> (&c)[LB],
                                              const Vec<R, CPT> (&x)[LB], Vec<R, CPT> (&o0)[LB], Vec<R, CPT> (&o1)[LB],
                                              R (&dprev)[CPT], R (&xprev)[CPT], R (&uprev)[CPT], bool first, bool& bad) {
  R rp[CPT];  // ~ 1 / d'(i-1): reciprocal of the PREDICTED pivot, off the chain (common.cuh bfb_pivot_step)
  if (!SLOW) {
#pragma unroll
    for (int j = 0; j < CPT; ++j) rp[j] = rcp_seed(dprev[j]);
  }
#pragma unroll
  for (int u = 0; u < LB; ++u) {
#pragma unroll
    for (int j = 0; j < CPT; ++j) {
      const R av = (u == 0 && first) ? (R)0 : a[u].v[j];  // dl(0) is never used (ekat_tridiag.hpp:24-27)
      R m, dc;
      if (SLOW) {
        m = (u == 0 && first) ? (R)0 : x_div(av, dprev[j]);
        dc = x_sub(b[u].v[j], x_mul(m, uprev[j]));        // d(i) -= dl(i)*du(i-1)      :301
      } else {
        dc = dprev[j];
        m = bfb_factor_level(av, b[u].v[j], uprev[j], dc, rp[j], bad);
      }
      const R xc = x_sub(x[u].v[j], x_mul(m, xprev[j]));  // X(i) -= dl(i)*X(i-1)       :315-316
      o0[u].v[j] = dc;
      o1[u].v[j] = xc;
      dprev[j] = dc;
      xprev[j] = xc;
      uprev[j] = c[u].v[j];
    }
  }
}
// The same sweep that also hands out the multipliers (EKAT_B200_CLOBBER_DIAGS: the reference leaves them in dl(1..),
// ekat_tridiag.hpp:299-300; dl(0) keeps what the caller put there).
template <typename R, int CPT, int LB, bool SLOW>
__device__ __forceinline__ void bfb_fwd_chunk_m(const Vec<R, CPT> (&a)[LB], const Vec<R, CPT> (&b)[LB], const Vec<R, CPT> (&c)[LB],
                                                const Vec<R, CPT> (&x)[LB], Vec<R, CPT> (&o0)[LB], Vec<R, CPT> (&o1)[LB],
                                                Vec<R, CPT> (&om)[LB], R (&dprev)[CPT], R (&xprev)[CPT], R (&uprev)[CPT], bool first,
                                                bool& bad) {
  R rp[CPT];
  if (!SLOW) {
#pragma unroll
    for (int j = 0; j < CPT; ++j) rp[j] = rcp_seed(dprev[j]);
  }
#pragma unroll
  for (int u = 0; u < LB; ++u) {
#pragma unroll
    for (int j = 0; j < CPT; ++j) {
      const R av = (u == 0 && first) ? (R)0 : a[u].v[j];
      R m, dc;
      if (SLOW) {
        m = (u == 0 && first) ? (R)0 : x_div(av, dprev[j]);
        dc = x_sub(b[u].v[j], x_mul(m, uprev[j]));
      } else {
        dc = dprev[j];
        m = bfb_factor_level(av, b[u].v[j], uprev[j], dc, rp[j], bad);
      }
      om[u].v[j] = (u == 0 && first) ? a[u].v[j] : m;
      const R xc = x_sub(x[u].v[j], x_mul(m, xprev[j]));
      o0[u].v[j] = dc;
      o1[u].v[j] = xc;
      dprev[j] = dc;
      xprev[j] = xc;
      uprev[j] = c[u].v[j];
    }
  }
}
template <typename R, int CPT, int LB, bool SLOW>
__device__ __forceinline__ void bfb_bwd_chunk(Vec<R, CPT> (&y)[LB], const Vec<R, CPT> (&dd)[LB], const Vec<R, CPT> (&ui)[LB],
                                              R (&xv)[CPT], bool& bad) {
#pragma unroll
  for (int u = LB - 1; u >= 0; --u) {
#pragma unroll
    for (int j = 0; j < CPT; ++j) {
      const R num = x_sub(y[u].v[j], x_mul(ui[u].v[j], xv[j]));  // X(i-1) - du(i-1)*X(i)   :318-319
      if (SLOW) xv[j] = x_div(num, dd[u].v[j]);
      else xv[j] = div_spec(num, dd[u].v[j], rcp_refined(dd[u].v[j]), bad);  // reciprocal is off the x chain
      y[u].v[j] = xv[j];
    }
  }
}


// Chunk length and renormalisation interval per layout. Column-major rows are 64 bytes swizzled (f64) or 48 bytes (f32), see t_ld.
template <typename R, bool CM>
struct TmaGeom {
  static constexpr int LB = kLB, RN = TmaCfg<R>::RN;
};
template <>
struct TmaGeom<double, true> {
  static constexpr int LB = 8, RN = 8;  // 64-byte rows, swizzled
};
template <>
struct TmaGeom<float, true> {
  static constexpr int LB = 12, RN = 2;
};

// bfb, one matrix and one right-hand side per column, input-only copy ring (tridiag_bfb_stream.cu). Level-major
// operands [nmat][nlev][stride] (CM = false) or the reference's column-major [ncol][stride] (CM = true, nmat == 1).
// Returns EKAT_B200_ERR_UNSUPPORTED when a tile does not fit on chip.
template <typename R, bool CM>
int tridiag_bfb_stream(R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nmat, int64_t a_stride, int64_t x_stride,
                       cudaStream_t stream, int clob = 0);  // clob (level-major only): bit 0 multipliers -> dl, bit 1 pivots -> d

}  // namespace ekb
