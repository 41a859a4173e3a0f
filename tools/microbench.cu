// Micro-benchmarks that size the design: fp64/fp32 dependent-chain latency and per-SM throughput of the
// operations on the tridiag critical path (FMA, IEEE divide, fast reciprocal), measured with clock64().
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o microbench tools/microbench.cu && ./microbench
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ double f_rcp(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  double e = fma(-x, r, 1.0);
  e = fma(e, e, e);
  r = fma(r, e, r);
  e = fma(-x, r, 1.0);
  return fma(r, e, r);
}

template <int OP, typename T>
__global__ void chain(T* out, T a, T b, int n, long long* cyc) {
  T x = a + threadIdx.x * (T)1e-3;
  long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < n; ++i) {
#pragma unroll
    for (int u = 0; u < 16; ++u) {
      if (OP == 0) x = fma(x, b, a);                 // dependent FMA
      if (OP == 1) x = a / x + b;                    // IEEE divide (+ add to keep it bounded)
      if (OP == 2) x = (T)f_rcp((double)x) + b;      // fast reciprocal
      if (OP == 3) x = __dmul_rn(x, b) - a;          // mul + sub unfused
    }
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = x;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int OP, typename T>
__global__ void thru(T* out, T a, T b, int n, long long* cyc) {
  T x[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) x[k] = a + (threadIdx.x + k) * (T)1e-3;
  long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < n; ++i) {
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        if (OP == 0) x[k] = fma(x[k], b, a);
        if (OP == 1) x[k] = a / x[k] + b;
        if (OP == 2) x[k] = (T)f_rcp((double)x[k]) + b;
      }
  }
  long long t1 = clock64();
  T s = 0;
  for (int k = 0; k < 8; ++k) s += x[k];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <typename F>
double run(F launch) {
  long long* d;
  cudaMalloc(&d, 8);
  launch(d);
  launch(d);
  cudaDeviceSynchronize();
  long long h;
  cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
  cudaFree(d);
  return (double)h;
}

int main() {
  double* o;
  float* of;
  cudaMalloc(&o, 1 << 24);
  cudaMalloc(&of, 1 << 24);
  const int n = 2000;
  const char* names[] = {"fma", "ieee div (+add)", "fast rcp (+add)", "mul,sub unfused"};
  for (int op = 0; op < 4; ++op) {
    double c;
    if (op == 0) c = run([&](long long* d) { chain<0, double><<<1, 32>>>(o, 1.0000001, 0.999, n, d); });
    if (op == 1) c = run([&](long long* d) { chain<1, double><<<1, 32>>>(o, 1.0000001, 0.999, n, d); });
    if (op == 2) c = run([&](long long* d) { chain<2, double><<<1, 32>>>(o, 1.0000001, 0.999, n, d); });
    if (op == 3) c = run([&](long long* d) { chain<3, double><<<1, 32>>>(o, 1.0000001, 0.999, n, d); });
    printf("f64 latency  %-18s %.1f cycles/op (1 warp, dependent chain)\n", names[op], c / (n * 16.0));
  }
  for (int op = 0; op < 2; ++op) {
    double c;
    if (op == 0) c = run([&](long long* d) { chain<0, float><<<1, 32>>>(of, 1.0000001f, 0.999f, n, d); });
    if (op == 1) c = run([&](long long* d) { chain<1, float><<<1, 32>>>(of, 1.0000001f, 0.999f, n, d); });
    printf("f32 latency  %-18s %.1f cycles/op\n", names[op], c / (n * 16.0));
  }
  for (int warps : {1, 2, 4, 8, 16, 32}) {
    for (int op = 0; op < 3; ++op) {
      double c;
      const int thr = 32 * warps;
      if (op == 0) c = run([&](long long* d) { thru<0, double><<<148, thr>>>(o, 1.0000001, 0.999, n, d); });
      if (op == 1) c = run([&](long long* d) { thru<1, double><<<148, thr>>>(o, 1.0000001, 0.999, n, d); });
      if (op == 2) c = run([&](long long* d) { thru<2, double><<<148, thr>>>(o, 1.0000001, 0.999, n, d); });
      const double ops = (double)n * 4 * 8 * thr;  // per SM
      printf("f64 thruput  %-18s warps/SM=%2d  %.2f ops/cycle/SM  (%.1f cycles per warp-op)\n", names[op], warps, ops / c,
             c / (n * 4.0 * 8 * warps) );
    }
  }
  int clk;
  cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  printf("clock rate attr %d kHz\n", clk);
  return 0;
}
