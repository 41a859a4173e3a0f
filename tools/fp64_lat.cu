// fp64 latency / per-warp issue interval on sm_100a: what bounds the bfb chains.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o /tmp/fp64_lat tools/fp64_lat.cu && /tmp/fp64_lat
// One CTA of W warps on one SM; per warp: (a) a chain of dependent DFMAs, (b) 8 independent DFMA chains interleaved,
// (c) a chain of MUFU.RCP64H + 1 DFMA, (d) dependent FFMA chain, (e) dependent DMUL/DADD pairs. Cycles per instruction
// by clock64 around 4096 x 8 instructions.
#include <cuda_runtime.h>

#include <cstdio>

__device__ __forceinline__ double rcp64h(double x) {
  double r;
  asm volatile("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  return r;
}

template <int MODE>
__global__ void k(double* out, long long* cyc, double a, double b) {
  double x[8];
  for (int i = 0; i < 8; ++i) x[i] = a + threadIdx.x * 1e-3 + i;
  float f = (float)a;
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < 4096; ++it) {
    if (MODE == 0) {
#pragma unroll
      for (int i = 0; i < 8; ++i) x[0] = fma(x[0], b, a);
    } else if (MODE == 1) {
#pragma unroll
      for (int i = 0; i < 8; ++i) x[i] = fma(x[i], b, a);
    } else if (MODE == 2) {
#pragma unroll
      for (int i = 0; i < 4; ++i) x[0] = fma(rcp64h(x[0]), b, a);
    } else if (MODE == 3) {
#pragma unroll
      for (int i = 0; i < 8; ++i) f = fmaf(f, (float)b, (float)a);
    } else if (MODE == 4) {
#pragma unroll
      for (int i = 0; i < 4; ++i) x[0] = __dadd_rn(__dmul_rn(x[0], b), a);
    } else if (MODE == 5) {  // two independent chains
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        x[0] = fma(x[0], b, a);
        x[1] = fma(x[1], b, a);
      }
    }
  }
  const long long t1 = clock64();
  double s = f;
  for (int i = 0; i < 8; ++i) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

int main() {
  double* out;
  long long* cyc;
  cudaMalloc(&out, 1 << 20);
  cudaMallocManaged(&cyc, 1024);
  const char* names[] = {"dependent DFMA", "8 independent DFMA chains", "MUFU.RCP64H + DFMA (per pair)", "dependent FFMA",
                         "dependent DMUL + DADD (per pair)", "2 independent DFMA chains"};
  const int per_iter[] = {8, 8, 4, 8, 4, 8};
  for (int W : {1, 4, 8, 16}) {
    printf("--- %d warp(s) on one SM\n", W);
    for (int m = 0; m < 6; ++m) {
      for (int rep = 0; rep < 2; ++rep) {
        switch (m) {
          case 0: k<0><<<1, 32 * W>>>(out, cyc, 1.0000001, 0.9999999); break;
          case 1: k<1><<<1, 32 * W>>>(out, cyc, 1.0000001, 0.9999999); break;
          case 2: k<2><<<1, 32 * W>>>(out, cyc, 1.0000001, 0.9999999); break;
          case 3: k<3><<<1, 32 * W>>>(out, cyc, 1.0000001, 0.9999999); break;
          case 4: k<4><<<1, 32 * W>>>(out, cyc, 1.0000001, 0.9999999); break;
          case 5: k<5><<<1, 32 * W>>>(out, cyc, 1.0000001, 0.9999999); break;
        }
        cudaDeviceSynchronize();
      }
      printf("%-36s %7.2f cycles per instruction (per warp)\n", names[m], (double)cyc[0] / (4096.0 * per_iter[m]));
    }
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
