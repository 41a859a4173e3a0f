// fused.cu -- "diffuse then remap" in one kernel (BASELINE config 4): shared-A, nrhs-RHS Thomas
// solve followed by LinInterp of every solution onto new levels, level-major operands.
//
// Fuses the two reference loops
//     thomas(team, dl, d, du, X)                      src/algorithm/ekat_tridiag.hpp:323-334
//     li.setup(...); li.lin_interp(...) per tracer     src/algorithm/ekat_lin_interp_impl.hpp:21-87
// so that the nrhs solutions never travel to HBM between them: the back-substituted columns
// stay in shared memory ([rhs][level][Tc]) and are the y1 operand of the interpolation, and the
// bracket search runs once per target level for all nrhs tracers (the reference's
// setup-once / interpolate-many contract, ekat_lin_interp.hpp:13-16).
// HBM traffic per column: (3 + nrhs) source level-arrays + x1 + x2 in, nrhs target arrays out
// ((3+10+2+10)*128*8 = 25,600 B at nrhs = 10).
//
// Tridiagonal arithmetic: the FMA/reciprocal Thomas of tridiag_levmajor.cu (tolerance-level).
// Interpolation arithmetic: the reference's 4 statements, unfused, in order (:153-156).
#include "common.cuh"

namespace ekb {

template <typename R>
__device__ __forceinline__ int bracket_strided(const R* __restrict__ x1, int stride, int km1, R v) {
  int lo = 0, n = km1;
  while (n > 0) {
    const int step = n >> 1, it = lo + step;
    const bool ge = v >= x1[(size_t)it * stride];
    lo = ge ? it + 1 : lo;
    n = ge ? n - step - 1 : step;
  }
  return lo > 0 ? lo - 1 : lo;
}

template <typename R, int PF>
__global__ void k_tridiag_lin_interp(const R* __restrict__ dl, const R* __restrict__ d, const R* __restrict__ du,
                                     const R* __restrict__ B, const R* __restrict__ x1, const R* __restrict__ x2,
                                     R* __restrict__ Y2, R* __restrict__ X, int64_t ncol, int nlev, int km2, int nrhs,
                                     int64_t pitch) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int Tc = blockDim.x, t = threadIdx.x, ty = threadIdx.y, Ty = blockDim.y;
  const size_t plane = (size_t)nlev * Tc;
  R* sa0 = reinterpret_cast<R*>(smem_raw);  // m = dl*r, later the staged x1 tile
  R* sa1 = sa0 + plane;                     // r = 1/pivot
  R* sa2 = sa1 + plane;                     // c = du*r
  R* sx = sa2 + plane;                      // [nrhs][nlev][Tc]: rhs -> y -> solution

  int64_t col = (int64_t)blockIdx.x * Tc + t;
  const bool live = col < ncol;
  if (!live) col = ncol - 1;

  if (ty == 0) {
    const R* pdl = dl + col;
    const R* pd = d + col;
    const R* pdu = du + col;
    R rdl[PF], rd[PF], rdu[PF];
#pragma unroll
    for (int u = 0; u < PF; ++u)
      if (u < nlev) {
        rdl[u] = ld_stream(pdl + (size_t)u * pitch);
        rd[u] = ld_stream(pd + (size_t)u * pitch);
        rdu[u] = ld_stream(pdu + (size_t)u * pitch);
      }
    R cprev = 0;
    for (int base = 0; base < nlev; base += PF) {
#pragma unroll
      for (int u = 0; u < PF; ++u) {
        const int i = base + u;
        if (i < nlev) {
          const R a = rdl[u], b = rd[u], c = rdu[u];
          const int nx = i + PF;
          if (nx < nlev) {
            rdl[u] = ld_stream(pdl + (size_t)nx * pitch);
            rd[u] = ld_stream(pd + (size_t)nx * pitch);
            rdu[u] = ld_stream(pdu + (size_t)nx * pitch);
          }
          const size_t si = (size_t)i * Tc + t;
          const R r = f_rcp(i > 0 ? f_fma(-a, cprev, b) : b);
          cprev = c * r;
          sa0[si] = a * r;
          sa1[si] = r;
          sa2[si] = cprev;
        }
      }
    }
  }
  __syncthreads();

  for (int j = ty; j < nrhs; j += Ty) {
    const R* pb = B + (size_t)j * nlev * pitch + col;
    R* sxj = sx + (size_t)j * plane;
    R rx[PF];
#pragma unroll
    for (int u = 0; u < PF; ++u)
      if (u < nlev) rx[u] = ld_stream(pb + (size_t)u * pitch);
    R xprev = 0;
    for (int base = 0; base < nlev; base += PF) {
#pragma unroll
      for (int u = 0; u < PF; ++u) {
        const int i = base + u;
        if (i < nlev) {
          const R x = rx[u];
          const int nx = i + PF;
          if (nx < nlev) rx[u] = ld_stream(pb + (size_t)nx * pitch);
          const size_t si = (size_t)i * Tc + t;
          const R xr = x * sa1[si];
          xprev = i > 0 ? f_fma(-sa0[si], xprev, xr) : xr;
          sxj[si] = xprev;
        }
      }
    }
    R xv = xprev;
    R* px = X ? X + (size_t)j * nlev * pitch + col : nullptr;
    if (px && live) st_stream(px + (size_t)(nlev - 1) * pitch, xv);
#pragma unroll 8
    for (int i = nlev - 2; i >= 0; --i) {
      const size_t si = (size_t)i * Tc + t;
      xv = f_fma(-sa2[si], xv, sxj[si]);
      sxj[si] = xv;
      if (px && live) st_stream(px + (size_t)i * pitch, xv);
    }
  }
  __syncthreads();
  if (km2 <= 0) return;

  // stage the source coordinates of the tile over the (now dead) multiplier plane
  {
    const int nthr = Tc * Ty, tid = ty * Tc + t;
    const int64_t c0 = (int64_t)blockIdx.x * Tc;
    for (int e = tid; e < nlev * Tc; e += nthr) {
      const int i = e / Tc, c = e - i * Tc;
      const int64_t gc = tmin<int64_t>(c0 + c, ncol - 1);
      sa0[e] = ld_stream(x1 + (size_t)i * pitch + gc);
    }
  }
  __syncthreads();

  const R* x1c = sa0 + t;
  for (int k2 = ty; k2 < km2; k2 += Ty) {
    const R xt = ld_stream(x2 + (size_t)k2 * pitch + col);
    const int k1 = bracket_strided<R>(x1c, Tc, nlev, xt);
    const int k1ph = k1 == nlev - 1 ? k1 - 1 : k1 + 1;
    const R xa = x1c[(size_t)k1 * Tc], xb = x1c[(size_t)k1ph * Tc];
    const R dxt = x_sub(xt, xa), dx = x_sub(xb, xa);
    for (int j = 0; j < nrhs; ++j) {
      const R* y1c = sx + (size_t)j * plane + t;
      const R ya = y1c[(size_t)k1 * Tc], yb = y1c[(size_t)k1ph * Tc];
      R v = x_sub(yb, ya);
      v = x_mul(v, dxt);
      v = x_div(v, dx);
      v = x_add(v, ya);
      if (live) st_stream(Y2 + ((size_t)j * km2 + k2) * pitch + col, v);
    }
  }
}

template <typename R>
int tridiag_lin_interp(const R* dl, const R* d, const R* du, const R* B, const R* x1, const R* x2, R* Y2, R* X,
                       int64_t ncol, int nlev, int km2, int nrhs, int64_t pitch, cudaStream_t stream) {
  const Tuning tu = tuning();
  int Tc = tu.cols_per_cta > 0 ? tu.cols_per_cta : 16;
  const size_t per_col = (size_t)(3 + nrhs) * nlev * sizeof(R);
  while (Tc > 1 && (size_t)Tc * per_col > kMaxSmemPerCta) Tc /= 2;
  if ((size_t)Tc * per_col > kMaxSmemPerCta) return EKAT_B200_ERR_TOO_LARGE;
  int Ty = nrhs;
  while (Tc * Ty > 1024) Ty = (Ty + 1) / 2;
  // a few more thread rows than rhs help the interpolation phase when nrhs is small
  while (Tc * Ty < 128 && Ty < 64) Ty *= 2;
  const size_t smem = (size_t)Tc * per_col;
  auto k = k_tridiag_lin_interp<R, 8>;
  int rc = allow_smem(k, smem);
  if (rc) return rc;
  dim3 grid((unsigned)((ncol + Tc - 1) / Tc)), block(Tc, Ty);
  k<<<grid, block, smem, stream>>>(dl, d, du, B, x1, x2, Y2, X, ncol, nlev, km2, nrhs, pitch);
  note_launch(sizeof(R) == 8 ? "tridiag_lin_interp<f64>" : "tridiag_lin_interp<f32>");
  return check_cuda(cudaGetLastError());
}

template int tridiag_lin_interp<double>(const double*, const double*, const double*, const double*, const double*,
                                        const double*, double*, double*, int64_t, int, int, int, int64_t, cudaStream_t);
template int tridiag_lin_interp<float>(const float*, const float*, const float*, const float*, const float*,
                                       const float*, float*, float*, int64_t, int, int, int, int64_t, cudaStream_t);

}  // namespace ekb
