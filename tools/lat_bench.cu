// Latency dissection of a one-warp-per-tile Thomas solve at small batch sizes (BASELINE config 1: 21,600 x 72 f64).
// Where do the ~10 us of a single 32-column tile go? Time stamps (globaltimer, ns) inside the kernel:
//   0 start | 1 barriers ready, loads issued | 2 first chunk landed | 3 all chunks landed | 4 forward sweep done
//   5 backward sweep done, stores issued | 6 stores complete
// and the kernel duration by CUDA events for grids of 1, 148, 444, 675 tiles.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lcuda -o lat_bench tools/lat_bench.cu && ./lat_bench
#include <cuda.h>
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <vector>

#include "../ekat_b200/csrc/common.cuh"

using namespace ekb;

static bool make_map(CUtensorMap* m, const double* base, int64_t ncol, int64_t nrows, int64_t pitch, int boxc, int boxr) {
  const cuuint64_t dims[2] = {(cuuint64_t)ncol, (cuuint64_t)nrows};
  const cuuint64_t strides[1] = {(cuuint64_t)pitch * 8};
  const cuuint32_t box[2] = {(cuuint32_t)boxc, (cuuint32_t)boxr};
  const cuuint32_t estr[2] = {1, 1};
  return cuTensorMapEncodeTiled(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, const_cast<double*>(base), dims, strides, box, estr,
                                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ double t_rcp(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  double e = fma(-x, r, 1.0);
  e = fma(e, e, e);
  return fma(r, e, r);
}

constexpr int LB = 8;

// MODE 0: TMA loads of all four planes up front (one mbarrier per chunk of LB levels), sweeps straight out of shared
//         memory, solution stored with plain coalesced st.global from the back-substitution.
// MODE 1: same loads, solution written to the plane and stored by one TMA tensor store per chunk.
// MODE 2: no TMA: every lane loads its own column with ld.global (coalesced rows), 8 levels of prefetch in registers.
template <int MODE>
__global__ void __launch_bounds__(32) k_small(const __grid_constant__ CUtensorMap mdl, const __grid_constant__ CUtensorMap md,
                                              const __grid_constant__ CUtensorMap mdu, const __grid_constant__ CUtensorMap mx,
                                              const double* __restrict__ gdl, const double* __restrict__ gd,
                                              const double* __restrict__ gdu, double* __restrict__ X, int64_t ncol, int nlev,
                                              int64_t pitch, unsigned long long* stamps) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int lane = threadIdx.x;
  const int nch = (nlev + LB - 1) / LB;
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem_raw);
  double* P = reinterpret_cast<double*>(smem_raw + 128);
  const size_t plane = (size_t)nch * LB * 32;
  double *sdl = P, *sd = P + plane, *sdu = P + 2 * plane, *sx = P + 3 * plane;
  const int c0 = blockIdx.x * 32;
  unsigned long long ts[7];
  ts[0] = gtime();
  if (MODE != 2) {
    if (lane == 0) {
      for (int k = 0; k < nch; ++k) mbar_init(&bar[k], 1);
      fence_mbar_init();
    }
    __syncwarp();
    if (lane == 0) {
      for (int k = 0; k < nch; ++k) {
        mbar_expect_tx(&bar[k], 4u * LB * 32 * 8);
        tma_load_2d(sdl + (size_t)k * LB * 32, &mdl, c0, k * LB, &bar[k]);
        tma_load_2d(sd + (size_t)k * LB * 32, &md, c0, k * LB, &bar[k]);
        tma_load_2d(sdu + (size_t)k * LB * 32, &mdu, c0, k * LB, &bar[k]);
        tma_load_2d(sx + (size_t)k * LB * 32, &mx, c0, k * LB, &bar[k]);
      }
    }
    ts[1] = gtime();
    mbar_wait(&bar[0], 0);
    ts[2] = gtime();
    // forward sweep (ratio form, renormalised every 8 levels), in place: sdu <- c', sx <- y
    double q1 = 1, q2 = 0, z1 = 0, cp = 0;
    for (int k = 0; k < nch; ++k) {
      mbar_wait(&bar[k], 0);
      if (k == nch - 1) ts[3] = gtime();
#pragma unroll
      for (int u = 0; u < LB; ++u) {
        const int i = k * LB + u;
        if (i < nlev) {
          const size_t s = (size_t)i * 32 + lane;
          double a = sdl[s];
          const double b = sd[s], c = sdu[s], x = sx[s];
          if (i == 0) a = 0;
          double q, z;
          if (u == 0) {
            q = fma(-(a * cp), q2, b);
            z = fma(-a, z1, x);
          } else {
            q = fma(b, q1, -((a * cp) * q2));
            z = fma(-a, z1, x * q1);
          }
          const double rq = t_rcp(q);
          const double g = (u == 0) ? rq : q1 * rq;
          const double y = z * rq;
          sdu[s] = (i == nlev - 1) ? 0.0 : c * g;
          sx[s] = y;
          if (u == LB - 1) {
            q2 = g;
            z1 = y;
            q1 = 1;
          } else {
            q2 = (u == 0) ? 1.0 : q1;
            q1 = q;
            z1 = z;
          }
          cp = c;
        }
      }
    }
    ts[4] = gtime();
    double xv = 0;
    const bool live = c0 + lane < ncol;
    for (int i = nlev - 1; i >= 0; --i) {
      const size_t s = (size_t)i * 32 + lane;
      xv = fma(-sdu[s], xv, sx[s]);
      if (MODE == 0) {
        if (live) __stcs(X + (size_t)i * pitch + c0 + lane, xv);
      } else {
        sx[s] = xv;
        if (i % LB == 0) {
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&mx, c0, i, sx + (size_t)i * 32);
            tma_store_commit();
          }
        }
      }
    }
    ts[5] = gtime();
    if (MODE == 1 && lane == 0) tma_store_wait_all<0>();
    ts[6] = gtime();
  } else {
    // plain loads: a lane walks its own column; rows are coalesced across the warp. Forward results to shared memory.
    const bool live = c0 + lane < ncol;
    const size_t col = live ? (size_t)c0 + lane : (size_t)ncol - 1;
    ts[1] = gtime();
    double q1 = 1, q2 = 0, z1 = 0, cp = 0;
    constexpr int PF = 8;
    double ra[PF], rb[PF], rc[PF], rx[PF];
#pragma unroll
    for (int u = 0; u < PF; ++u) {
      const size_t g = (size_t)u * pitch + col;
      ra[u] = __ldcs(gdl + g);
      rb[u] = __ldcs(gd + g);
      rc[u] = __ldcs(gdu + g);
      rx[u] = __ldcs(X + g);
    }
    ts[2] = gtime();
    for (int k = 0; k < nch; ++k) {
      double na[PF], nb[PF], nc[PF], nx[PF];
      if (k + 1 < nch) {
#pragma unroll
        for (int u = 0; u < PF; ++u) {
          const int i = (k + 1) * LB + u;
          const size_t g = (size_t)(i < nlev ? i : nlev - 1) * pitch + col;
          na[u] = __ldcs(gdl + g);
          nb[u] = __ldcs(gd + g);
          nc[u] = __ldcs(gdu + g);
          nx[u] = __ldcs(X + g);
        }
      }
      if (k == nch - 1) ts[3] = gtime();
#pragma unroll
      for (int u = 0; u < LB; ++u) {
        const int i = k * LB + u;
        if (i < nlev) {
          const size_t s = (size_t)i * 32 + lane;
          double a = ra[u];
          const double b = rb[u], c = rc[u], x = rx[u];
          if (i == 0) a = 0;
          double q, z;
          if (u == 0) {
            q = fma(-(a * cp), q2, b);
            z = fma(-a, z1, x);
          } else {
            q = fma(b, q1, -((a * cp) * q2));
            z = fma(-a, z1, x * q1);
          }
          const double rq = t_rcp(q);
          const double g = (u == 0) ? rq : q1 * rq;
          const double y = z * rq;
          sdu[s] = (i == nlev - 1) ? 0.0 : c * g;
          sx[s] = y;
          if (u == LB - 1) {
            q2 = g;
            z1 = y;
            q1 = 1;
          } else {
            q2 = (u == 0) ? 1.0 : q1;
            q1 = q;
            z1 = z;
          }
          cp = c;
        }
      }
#pragma unroll
      for (int u = 0; u < PF; ++u) {
        ra[u] = na[u];
        rb[u] = nb[u];
        rc[u] = nc[u];
        rx[u] = nx[u];
      }
    }
    ts[4] = gtime();
    double xv = 0;
    for (int i = nlev - 1; i >= 0; --i) {
      const size_t s = (size_t)i * 32 + lane;
      xv = fma(-sdu[s], xv, sx[s]);
      if (live) __stcs(X + (size_t)i * pitch + c0 + lane, xv);
    }
    ts[5] = ts[6] = gtime();
  }
  if (lane == 0 && stamps && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1)) {
    unsigned long long* o = stamps + (blockIdx.x == 0 ? 0 : 8);
    for (int i = 0; i < 7; ++i) o[i] = ts[i];
  }
}

// MODE 3 (separate kernel): the candidate small-batch kernel. Loads in groups of GL levels (few TMA issues), sweeps
// staged chunk-wise in registers with the next chunk's shared-memory loads issued before the current chunk's stores
// (no load ever waits behind a store to the same array), solution stored straight from registers.
template <int GLC>
__global__ void __launch_bounds__(32) k_small3(const __grid_constant__ CUtensorMap mdl, const __grid_constant__ CUtensorMap md,
                                               const __grid_constant__ CUtensorMap mdu, const __grid_constant__ CUtensorMap mx,
                                               double* __restrict__ X, int64_t ncol, int nlev, int64_t pitch,
                                               unsigned long long* stamps) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int lane = threadIdx.x;
  const int nch = (nlev + LB - 1) / LB;
  const int GL = GLC * LB, ngrp = (nch + GLC - 1) / GLC;
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem_raw);
  double* P = reinterpret_cast<double*>(smem_raw + 128);
  const size_t plane = (size_t)ngrp * GL * 32;
  double *sdl = P, *sd = P + plane, *sdu = P + 2 * plane, *sx = P + 3 * plane;
  const int c0 = blockIdx.x * 32;
  unsigned long long ts[7];
  ts[0] = gtime();
  if (lane == 0) {
    for (int g = 0; g < ngrp; ++g) mbar_init(&bar[g], 1);
    fence_mbar_init();
    for (int g = 0; g < ngrp; ++g) {
      mbar_expect_tx(&bar[g], 4u * GL * 32 * 8);
      tma_load_2d(sdl + (size_t)g * GL * 32, &mdl, c0, g * GL, &bar[g]);
      tma_load_2d(sd + (size_t)g * GL * 32, &md, c0, g * GL, &bar[g]);
      tma_load_2d(sdu + (size_t)g * GL * 32, &mdu, c0, g * GL, &bar[g]);
      tma_load_2d(sx + (size_t)g * GL * 32, &mx, c0, g * GL, &bar[g]);
    }
  }
  __syncwarp();
  ts[1] = gtime();
  mbar_wait(&bar[0], 0);
  ts[2] = gtime();
  double q1 = 1, q2 = 0, z1 = 0, cp = 0;
  double a[LB], b[LB], c[LB], x[LB];
  auto stage = [&](int k, double (&ra)[LB], double (&rb)[LB], double (&rc)[LB], double (&rx)[LB]) {
    const size_t s0 = (size_t)k * LB * 32 + lane;
#pragma unroll
    for (int u = 0; u < LB; ++u) {
      ra[u] = sdl[s0 + u * 32];
      rb[u] = sd[s0 + u * 32];
      rc[u] = sdu[s0 + u * 32];
      rx[u] = sx[s0 + u * 32];
    }
  };
  stage(0, a, b, c, x);
  a[0] = 0;
  for (int k = 0; k < nch; ++k) {
    double na[LB], nb[LB], nc[LB], nx[LB];
    if (k + 1 < nch) {
      if ((k + 1) % GLC == 0) mbar_wait(&bar[(k + 1) / GLC], 0);
      stage(k + 1, na, nb, nc, nx);
    }
    if (k == nch - 1) ts[3] = gtime();
    double o0[LB], o1[LB];
#pragma unroll
    for (int u = 0; u < LB; ++u) {
      double q, z;
      if (u == 0) {
        q = fma(-(a[u] * cp), q2, b[u]);
        z = fma(-a[u], z1, x[u]);
      } else {
        q = fma(b[u], q1, -((a[u] * cp) * q2));
        z = fma(-a[u], z1, x[u] * q1);
      }
      const double rq = t_rcp(q);
      const double g = (u == 0) ? rq : q1 * rq;
      const double y = z * rq;
      o0[u] = (k * LB + u == nlev - 1) ? 0.0 : c[u] * g;
      o1[u] = y;
      if (u == LB - 1) {
        q2 = g;
        z1 = y;
        q1 = 1;
      } else {
        q2 = (u == 0) ? 1.0 : q1;
        q1 = q;
        z1 = z;
      }
      cp = c[u];
    }
    const size_t s0 = (size_t)k * LB * 32 + lane;
#pragma unroll
    for (int u = 0; u < LB; ++u) {
      sdu[s0 + u * 32] = o0[u];
      sx[s0 + u * 32] = o1[u];
    }
#pragma unroll
    for (int u = 0; u < LB; ++u) {
      a[u] = na[u];
      b[u] = nb[u];
      c[u] = nc[u];
      x[u] = nx[u];
    }
  }
  ts[4] = gtime();
  double xv = 0;
  const bool live = c0 + lane < ncol;
  double y[LB], cc[LB];
  auto stage_b = [&](int k, double (&ry)[LB], double (&rc)[LB]) {
    const size_t s0 = (size_t)k * LB * 32 + lane;
#pragma unroll
    for (int u = 0; u < LB; ++u) {
      ry[u] = sx[s0 + u * 32];
      rc[u] = sdu[s0 + u * 32];
    }
  };
  stage_b(nch - 1, y, cc);
  for (int k = nch - 1; k >= 0; --k) {
    double ny[LB], ncc[LB];
    if (k > 0) stage_b(k - 1, ny, ncc);
#pragma unroll
    for (int u = LB - 1; u >= 0; --u) {
      const int i = k * LB + u;
      if (i < nlev) {
        xv = fma(-cc[u], xv, y[u]);
        if (live) __stcs(X + (size_t)i * pitch + c0 + lane, xv);
      }
    }
#pragma unroll
    for (int u = 0; u < LB; ++u) {
      y[u] = ny[u];
      cc[u] = ncc[u];
    }
  }
  ts[5] = ts[6] = gtime();
  if (lane == 0 && stamps && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1)) {
    unsigned long long* o = stamps + (blockIdx.x == 0 ? 0 : 8);
    for (int i = 0; i < 7; ++i) o[i] = ts[i];
  }
}

template <int GLC>
__global__ void __launch_bounds__(32) k_small4(const __grid_constant__ CUtensorMap mdl, const __grid_constant__ CUtensorMap md,
                                               const __grid_constant__ CUtensorMap mdu, const __grid_constant__ CUtensorMap mx,
                                               double* __restrict__ X, int64_t ncol, int nlev, int64_t pitch,
                                               unsigned long long* stamps) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int lane = threadIdx.x;
  const int nch = (nlev + LB - 1) / LB;
  const int GL = GLC * LB, ngrp = (nch + GLC - 1) / GLC;
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem_raw);
  double* P = reinterpret_cast<double*>(smem_raw + 128);
  const size_t plane = (size_t)ngrp * GL * 32;
  double *sdl = P, *sd = P + plane, *sdu = P + 2 * plane, *sx = P + 3 * plane;
  const int c0 = blockIdx.x * 32;
  unsigned long long ts[7];
  ts[0] = gtime();
  if (lane == 0) {
    for (int g = 0; g < ngrp; ++g) mbar_init(&bar[g], 1);
    fence_mbar_init();
    for (int g = 0; g < ngrp; ++g) {
      mbar_expect_tx(&bar[g], 4u * GL * 32 * 8);
      tma_load_2d(sdl + (size_t)g * GL * 32, &mdl, c0, g * GL, &bar[g]);
      tma_load_2d(sd + (size_t)g * GL * 32, &md, c0, g * GL, &bar[g]);
      tma_load_2d(sdu + (size_t)g * GL * 32, &mdu, c0, g * GL, &bar[g]);
      tma_load_2d(sx + (size_t)g * GL * 32, &mx, c0, g * GL, &bar[g]);
    }
  }
  __syncwarp();
  ts[1] = gtime();
  mbar_wait(&bar[0], 0);
  ts[2] = gtime();
  double q1 = 1, q2 = 0, z1 = 0, cp = 0;
  double a[LB], b[LB], c[LB], x[LB];
  auto stage = [&](int k, double (&ra)[LB], double (&rb)[LB], double (&rc)[LB], double (&rx)[LB]) {
    const size_t s0 = (size_t)k * LB * 32 + lane;
#pragma unroll
    for (int u = 0; u < LB; ++u) {
      ra[u] = sdl[s0 + u * 32];
      rb[u] = sd[s0 + u * 32];
      rc[u] = sdu[s0 + u * 32];
      rx[u] = sx[s0 + u * 32];
    }
  };
  stage(0, a, b, c, x);
  a[0] = 0;
  for (int k = 0; k < nch; ++k) {
    double na[LB], nb[LB], nc[LB], nx[LB];
    if (k + 1 < nch) {
      if ((k + 1) % GLC == 0) mbar_wait(&bar[(k + 1) / GLC], 0);
      stage(k + 1, na, nb, nc, nx);
    }
    if (k == nch - 1) ts[3] = gtime();
    double o0[LB], o1[LB];
#pragma unroll
    for (int u = 0; u < LB; ++u) {
      double q, z;
      if (u == 0) {
        q = fma(-(a[u] * cp), q2, b[u]);
        z = fma(-a[u], z1, x[u]);
      } else {
        q = fma(b[u], q1, -((a[u] * cp) * q2));
        z = fma(-a[u], z1, x[u] * q1);
      }
      o0[u] = q;
      o1[u] = z;
      if (u == LB - 1) {
        const double rq = t_rcp(q);
        q2 = q1 * rq;
        z1 = z * rq;
        q1 = 1;
      } else {
        q2 = (u == 0) ? 1.0 : q1;
        q1 = q;
        z1 = z;
      }
      cp = c[u];
    }
    const size_t s0 = (size_t)k * LB * 32 + lane;
#pragma unroll
    for (int u = 0; u < LB; ++u) {
      sd[s0 + u * 32] = o0[u];
      sx[s0 + u * 32] = o1[u];
    }
#pragma unroll
    for (int u = 0; u < LB; ++u) {
      a[u] = na[u];
      b[u] = nb[u];
      c[u] = nc[u];
      x[u] = nx[u];
    }
  }
  ts[4] = gtime();
  double xv = 0;
  double qq[LB], zz[LB], uu[LB];
  auto stage_b = [&](int k, double (&rq)[LB], double (&rz)[LB], double (&ru)[LB]) {
    const size_t s0 = (size_t)k * LB * 32 + lane;
#pragma unroll
    for (int u = 0; u < LB; ++u) {
      rq[u] = sd[s0 + u * 32];
      rz[u] = sx[s0 + u * 32];
      ru[u] = sdu[s0 + u * 32];
    }
  };
  stage_b(nch - 1, qq, zz, uu);
  double* gx = X + c0 + lane;
  for (int k = nch - 1; k >= 0; --k) {
    double nq[LB], nz[LB], nu[LB];
    if (k > 0) stage_b(k - 1, nq, nz, nu);
    double y[LB], cc[LB], o[LB];
#pragma unroll
    for (int u = 0; u < LB; ++u) {
      const double r = t_rcp(qq[u]);
      const double g = (u == 0) ? r : qq[u == 0 ? 0 : u - 1] * r;
      y[u] = zz[u] * r;
      cc[u] = (k * LB + u == nlev - 1) ? 0.0 : uu[u] * g;
    }
#pragma unroll
    for (int u = LB - 1; u >= 0; --u) {
      xv = fma(-cc[u], xv, y[u]);
      o[u] = xv;
    }
    double* g = gx + (size_t)k * LB * pitch;
#pragma unroll
    for (int u = LB - 1; u >= 0; --u) __stcs(g + (size_t)u * pitch, o[u]);
#pragma unroll
    for (int u = 0; u < LB; ++u) {
      qq[u] = nq[u];
      zz[u] = nz[u];
      uu[u] = nu[u];
    }
  }
  ts[5] = ts[6] = gtime();
  if (lane == 0 && stamps && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1)) {
    unsigned long long* o = stamps + (blockIdx.x == 0 ? 0 : 8);
    for (int i = 0; i < 7; ++i) o[i] = ts[i];
  }
}

__global__ void k_empty(int* p) {
  if (p && threadIdx.x == 1000) *p = 1;
}
__global__ void k_spin(long long cycles) {
  const long long t0 = clock64();
  while (clock64() - t0 < cycles) {
  }
}

template <class F>
static float timed(F launch, int reps = 20) {
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  float tot = 0;
  for (int i = 0; i < reps + 3; ++i) {
    k_spin<<<1, 1>>>(200000);
    cudaEventRecord(a);
    launch();
    cudaEventRecord(b);
    cudaDeviceSynchronize();
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    if (i >= 3) tot += ms;
  }
  return tot / reps * 1e3f;
}

int main() {
  const int nlev = 72;
  const int64_t ncol = 21600, pitch = 21600;
  double *dl, *d, *du, *X, *B;
  const size_t n = (size_t)nlev * pitch;
  cudaMalloc(&dl, n * 8);
  cudaMalloc(&d, n * 8);
  cudaMalloc(&du, n * 8);
  cudaMalloc(&X, n * 8);
  cudaMalloc(&B, n * 8);
  std::vector<double> h(n);
  for (size_t i = 0; i < n; ++i) h[i] = 1.0 + (double)(i % 17) * 0.01;
  cudaMemcpy(dl, h.data(), n * 8, cudaMemcpyHostToDevice);
  cudaMemcpy(du, h.data(), n * 8, cudaMemcpyHostToDevice);
  cudaMemcpy(B, h.data(), n * 8, cudaMemcpyHostToDevice);
  for (size_t i = 0; i < n; ++i) h[i] = 4.0 + (double)(i % 13) * 0.01;
  cudaMemcpy(d, h.data(), n * 8, cudaMemcpyHostToDevice);
  cudaMemcpy(X, B, n * 8, cudaMemcpyDeviceToDevice);
  CUtensorMap mdl, md, mdu, mx;
  if (!make_map(&mdl, dl, ncol, nlev, pitch, 32, LB) || !make_map(&md, d, ncol, nlev, pitch, 32, LB) ||
      !make_map(&mdu, du, ncol, nlev, pitch, 32, LB) || !make_map(&mx, X, ncol, nlev, pitch, 32, LB)) {
    printf("tensor map failed\n");
    return 1;
  }
  unsigned long long* stamps;
  cudaMalloc(&stamps, 16 * 8);
  const size_t smem = 128 + 4 * (size_t)((nlev + LB - 1) / LB) * LB * 32 * 8;
  cudaFuncSetAttribute(k_small<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaFuncSetAttribute(k_small<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaFuncSetAttribute(k_small<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  printf("empty kernel <<<1,32>>>: %.2f us   <<<675,32>>>: %.2f us   <<<675,128>>>: %.2f us\n", timed([&] { k_empty<<<1, 32>>>(nullptr); }),
         timed([&] { k_empty<<<675, 32>>>(nullptr); }), timed([&] { k_empty<<<675, 128>>>(nullptr); }));
  for (int mode = 0; mode < 3; ++mode) {
    for (int tiles : {1, 148, 444, 675}) {
      auto launch = [&] {
        if (mode == 0) k_small<0><<<tiles, 32, smem>>>(mdl, md, mdu, mx, dl, d, du, X, ncol, nlev, pitch, stamps);
        if (mode == 1) k_small<1><<<tiles, 32, smem>>>(mdl, md, mdu, mx, dl, d, du, X, ncol, nlev, pitch, stamps);
        if (mode == 2) k_small<2><<<tiles, 32, smem>>>(mdl, md, mdu, mx, dl, d, du, X, ncol, nlev, pitch, stamps);
      };
      const float us = timed(launch);
      unsigned long long hs[16];
      cudaMemcpy(hs, stamps, sizeof(hs), cudaMemcpyDeviceToHost);
      printf("mode %d tiles %3d: %6.2f us | first CTA ns:", mode, tiles, us);
      for (int i = 1; i < 7; ++i) printf(" %llu", hs[i] - hs[0]);
      printf(" | last CTA start +%lld ns, its stamps:", (long long)(hs[8] - hs[0]));
      for (int i = 1; i < 7; ++i) printf(" %llu", hs[8 + i] - hs[8]);
      printf("\n");
    }
  }
  {
    constexpr int GLC = 3;
    CUtensorMap gdl, gd, gdu, gx;
    if (!make_map(&gdl, dl, ncol, nlev, pitch, 32, GLC * LB) || !make_map(&gd, d, ncol, nlev, pitch, 32, GLC * LB) ||
        !make_map(&gdu, du, ncol, nlev, pitch, 32, GLC * LB) || !make_map(&gx, X, ncol, nlev, pitch, 32, GLC * LB)) {
      printf("tensor map failed\n");
      return 1;
    }
    const int nch = (nlev + LB - 1) / LB, ngrp = (nch + GLC - 1) / GLC;
    const size_t smem3 = 128 + 4 * (size_t)ngrp * GLC * LB * 32 * 8;
    cudaFuncSetAttribute(k_small3<GLC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem3);
    for (int tiles : {1, 148, 444, 675}) {
      auto launch = [&] { k_small3<GLC><<<tiles, 32, smem3>>>(gdl, gd, gdu, gx, X, ncol, nlev, pitch, stamps); };
      const float us = timed(launch);
      unsigned long long hs[16];
      cudaMemcpy(hs, stamps, sizeof(hs), cudaMemcpyDeviceToHost);
      printf("mode 3 tiles %3d: %6.2f us | first CTA ns:", tiles, us);
      for (int i = 1; i < 7; ++i) printf(" %llu", hs[i] - hs[0]);
      printf(" | last CTA start +%lld ns, its stamps:", (long long)(hs[8] - hs[0]));
      for (int i = 1; i < 7; ++i) printf(" %llu", hs[8 + i] - hs[8]);
      // back to back: 50 launches between one event pair
      cudaEvent_t a, b;
      cudaEventCreate(&a);
      cudaEventCreate(&b);
      for (int i = 0; i < 5; ++i) launch();
      k_spin<<<1, 1>>>(200000);
      cudaEventRecord(a);
      for (int i = 0; i < 50; ++i) launch();
      cudaEventRecord(b);
      cudaDeviceSynchronize();
      float ms;
      cudaEventElapsedTime(&ms, a, b);
      printf(" | back-to-back %.2f us/launch\n", ms / 50 * 1e3f);
    }
    cudaFuncSetAttribute(k_small4<GLC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem3);
    for (int tiles : {1, 148, 444, 675}) {
      auto launch = [&] { k_small4<GLC><<<tiles, 32, smem3>>>(gdl, gd, gdu, gx, X, ncol, nlev, pitch, stamps); };
      const float us = timed(launch);
      unsigned long long hs[16];
      cudaMemcpy(hs, stamps, sizeof(hs), cudaMemcpyDeviceToHost);
      printf("mode 4 tiles %3d: %6.2f us | first CTA ns:", tiles, us);
      for (int i = 1; i < 7; ++i) printf(" %llu", hs[i] - hs[0]);
      printf(" | last CTA start +%lld ns, its stamps:", (long long)(hs[8] - hs[0]));
      for (int i = 1; i < 7; ++i) printf(" %llu", hs[8 + i] - hs[8]);
      // back to back: 50 launches between one event pair
      cudaEvent_t a, b;
      cudaEventCreate(&a);
      cudaEventCreate(&b);
      for (int i = 0; i < 5; ++i) launch();
      k_spin<<<1, 1>>>(200000);
      cudaEventRecord(a);
      for (int i = 0; i < 50; ++i) launch();
      cudaEventRecord(b);
      cudaDeviceSynchronize();
      float ms;
      cudaEventElapsedTime(&ms, a, b);
      printf(" | back-to-back %.2f us/launch\n", ms / 50 * 1e3f);
    }
    cudaMemcpy(X, B, n * 8, cudaMemcpyDeviceToDevice);
    k_small3<GLC><<<675, 32, smem3>>>(gdl, gd, gdu, gx, X, ncol, nlev, pitch, stamps);
    std::vector<double> r3(n), r0(n);
    cudaMemcpy(r3.data(), X, n * 8, cudaMemcpyDeviceToHost);
    cudaMemcpy(X, B, n * 8, cudaMemcpyDeviceToDevice);
    k_small<0><<<675, 32, smem>>>(mdl, md, mdu, mx, dl, d, du, X, ncol, nlev, pitch, stamps);
    cudaMemcpy(r0.data(), X, n * 8, cudaMemcpyDeviceToHost);
    double md_ = 0;
    for (size_t i = 0; i < n; ++i) md_ = fmax(md_, fabs(r3[i] - r0[i]));
    printf("mode 3 vs mode 0 max abs diff %.3e (x[0]=%.6f)\n", md_, r3[0]);
    auto e0 = [&] { k_empty<<<675, 32>>>(nullptr); };
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    k_spin<<<1, 1>>>(200000);
    cudaEventRecord(a);
    for (int i = 0; i < 50; ++i) e0();
    cudaEventRecord(b);
    cudaDeviceSynchronize();
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    printf("empty kernel back-to-back: %.2f us/launch\n", ms / 50 * 1e3f);
  }
  cudaError_t e = cudaDeviceSynchronize();
  printf("status: %s\n", cudaGetErrorString(e));
  return 0;
}
