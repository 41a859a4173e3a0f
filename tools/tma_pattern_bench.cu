// What HBM bandwidth can the level-major tile access pattern reach at all?  Pure data movement with the same
// shape as the tridiag kernels: one-warp CTAs, each streaming a W-column tile of NIN [nlev][ncol] arrays through
// shared memory by 2D TMA boxes (LB levels x ROWB bytes) and writing one array back by TMA stores.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/tmab tools/tma_pattern_bench.cu && /tmp/tmab [ncol] [nlev]
#include <cstdio>
#include <cstdlib>
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn enc() {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
  return (EncodeTiledFn)p;
}
static CUtensorMap mk(double* base, int64_t ncol, int64_t nrows, int boxc, int boxr, int promo) {
  CUtensorMap m;
  cuuint64_t dims[2] = {(cuuint64_t)ncol, (cuuint64_t)nrows};
  cuuint64_t strides[1] = {(cuuint64_t)ncol * 8};
  cuuint32_t box[2] = {(cuuint32_t)boxc, (cuuint32_t)boxr};
  cuuint32_t es[2] = {1, 1};
  CUresult r = enc()(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, (CUtensorMapL2promotion)promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); exit(1); }
  return m;
}

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, unsigned c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(b)), "r"(c)); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, unsigned n) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(b)), "r"(n) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* b, unsigned ph) {
  asm volatile("{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n" ::"r"(s32(b)), "r"(ph) : "memory");
}
__device__ __forceinline__ void ld2d(void* dst, const void* m, int c, int r, uint64_t* b) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(s32(dst)), "l"(m), "r"(c), "r"(r), "r"(s32(b)) : "memory");
}
__device__ __forceinline__ void st2d(const void* m, int c, int r, const void* src) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];" ::"l"(m), "r"(c), "r"(r), "r"(s32(src)) : "memory");
}

// MODE 0: chunk-synchronous ring of S stages holding all NIN arrays (load k+S after chunk k is stored)
template <int NIN>
__global__ void __launch_bounds__(32) k_copy(const __grid_constant__ CUtensorMap min_, const __grid_constant__ CUtensorMap mout,
                                             int nlev, int W, int LB, int S) {
  extern __shared__ __align__(128) unsigned char sm[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm);
  unsigned char* buf = sm + 128;
  const unsigned chunk = (unsigned)LB * W * 8;
  const int nch = nlev / LB, c0 = blockIdx.x * W;
  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) mbar_init(&bar[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    for (int k = 0; k < S && k < nch; ++k) {
      mbar_expect(&bar[k], NIN * chunk);
      for (int a = 0; a < NIN; ++a) ld2d(buf + ((size_t)k * NIN + a) * chunk, &min_, c0, a * nlev + k * LB, &bar[k]);
    }
  }
  __syncwarp();
  for (int k = 0; k < nch; ++k) {
    const int s = k % S;
    mbar_wait(&bar[s], (k / S) & 1);
    if (threadIdx.x == 0) {
      st2d(&mout, c0, k * LB, buf + ((size_t)s * NIN) * chunk);
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      if (k + S < nch) {
        mbar_expect(&bar[s], NIN * chunk);
        for (int a = 0; a < NIN; ++a) ld2d(buf + ((size_t)s * NIN + a) * chunk, &min_, c0, a * nlev + (k + S) * LB, &bar[s]);
      }
    }
    __syncwarp();
  }
}

int main(int argc, char** argv) {
  const int64_t ncol = argc > 1 ? atoll(argv[1]) : 345600;
  const int nlev = argc > 2 ? atoi(argv[2]) : 128;
  constexpr int NIN = 4;
  double *in, *out;
  cudaMalloc(&in, (size_t)NIN * nlev * ncol * 8);
  cudaMalloc(&out, (size_t)nlev * ncol * 8);
  cudaMemset(in, 0, (size_t)NIN * nlev * ncol * 8);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const double bytes = (double)(NIN + 1) * nlev * ncol * 8;
  struct Cfg { int W, LB, S, pad_kb, promo; };
  const Cfg cfgs[] = {{32, 8, 2, 75, 2},  {32, 8, 4, 75, 2},  {32, 8, 8, 75, 2},  {32, 8, 4, 0, 2},   {32, 8, 8, 0, 2},
                      {32, 16, 4, 75, 2}, {64, 8, 4, 75, 2},  {64, 8, 4, 110, 2}, {64, 8, 6, 110, 2}, {128, 8, 4, 110, 2},
                      {32, 8, 4, 75, 0},  {32, 8, 4, 75, 1},  {32, 8, 4, 75, 3},  {32, 4, 8, 75, 2},  {32, 32, 2, 75, 2}};
  for (const Cfg& c : cfgs) {
    CUtensorMap mi = mk(in, ncol, (int64_t)NIN * nlev, c.W, c.LB, c.promo), mo = mk(out, ncol, nlev, c.W, c.LB, c.promo);
    size_t smem = 128 + (size_t)c.S * NIN * c.LB * c.W * 8;
    if (smem < (size_t)c.pad_kb * 1024) smem = (size_t)c.pad_kb * 1024;
    if (smem > 227 * 1024) { printf("skip\n"); continue; }
    cudaFuncSetAttribute(k_copy<NIN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const unsigned grid = (unsigned)((ncol + c.W - 1) / c.W);
    for (int i = 0; i < 3; ++i) k_copy<NIN><<<grid, 32, smem>>>(mi, mo, nlev, c.W, c.LB, c.S);
    cudaEventRecord(e0);
    for (int i = 0; i < 10; ++i) k_copy<NIN><<<grid, 32, smem>>>(mi, mo, nlev, c.W, c.LB, c.S);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    ms /= 10;
    printf("W=%3d (%4d B rows) LB=%2d S=%d smem=%3zu KB (%d CTA/SM) promo=%d : %.4f ms  %.0f GB/s  err=%s\n", c.W, c.W * 8, c.LB, c.S,
           smem / 1024, (int)(233472 / (smem + 1024)), c.promo, ms, bytes / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
