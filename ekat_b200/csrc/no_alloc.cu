// no_alloc.cu -- ekat_b200_lin_interp_static_map_{f64,f32}: LinInterp::setup followed by LinInterp::lin_interp with the
// index map (LinInterp::m_indx_map, the one piece of scratch memory this library's kernels need from their caller) in a
// file-scope device array instead of a caller-owned allocation. No counterpart in the reference: added
// for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
// harnesses that forbid it). The entry point calls no cudaMalloc / cudaFree / allocating library routine.
//
// The array is part of the library's device image: EVERY GPU that loads libekat_b200.so holds kStaticMapBytes of it
// until the process exits. All calls (all streams, all host threads) on one device share it, and it keeps the last
// call's contents:
//   * nothing in it has to start at zero -- the setup kernel writes all ncol*km2 entries the apply kernel then reads,
//     in stream order, so a stale map can never be read;
//   * two calls that overlap on the GPU (different streams, no dependency between them) RACE on it: serialise them.
// In host code `g_static_indx_map` is only a stand-in for the device symbol, not a device address; the address comes from
// cudaGetSymbolAddress, once per device, with that device current (static_map_address below).
#include <mutex>

#include "common.cuh"

namespace ekb {

template <typename R>
int lin_interp(int mode, const R* x1, const R* x2, const R* y1, R* y2, int* indx, const int* indx_in, int64_t ncol,
               int km1, int km2, int64_t pitch1, int64_t pitch2, int nfields, int64_t y1_fstride, int64_t y2_fstride,
               cudaStream_t stream);  // lin_interp.cu: mode 1 = setup (writes indx), 2 = apply (reads indx_in)

constexpr int64_t kStaticMapInts = EKAT_B200_STATIC_MAP_MAX_ENTRIES;  // 4 Mi entries = 16 MiB per device
static __device__ __align__(256) int g_static_indx_map[kStaticMapInts];

constexpr int kMaxDevices = 64;
static std::mutex g_addr_mutex;
static int* g_static_map_addr[kMaxDevices] = {};

// Device address of g_static_indx_map on the CURRENT device; looked up once per device.
static int static_map_address(int** out) {
  int dev = 0;
  int rc = check_cuda(cudaGetDevice(&dev));
  if (rc != EKAT_B200_OK) return rc;
  if (dev < 0 || dev >= kMaxDevices) return EKAT_B200_ERR_UNSUPPORTED;
  std::lock_guard<std::mutex> lock(g_addr_mutex);
  if (!g_static_map_addr[dev]) {
    void* p = nullptr;
    rc = check_cuda(cudaGetSymbolAddress(&p, g_static_indx_map));
    if (rc != EKAT_B200_OK) return rc;
    if (!p) return EKAT_B200_ERR_CUDA;
    g_static_map_addr[dev] = static_cast<int*>(p);
  }
  *out = g_static_map_addr[dev];
  return EKAT_B200_OK;
}

template <typename R>
static int lin_interp_static_map(const R* x1, const R* x2, const R* y1, R* y2, int64_t ncol, int km1, int km2,
                                 int64_t pitch1, int64_t pitch2, int nfields, int64_t y1_fstride, int64_t y2_fstride,
                                 cudaStream_t stream) {
  // the same argument checks as ekat_b200_lin_interp_* (capi.cu li_entry)
  if (ncol < 0 || km1 < 0 || km2 < 0 || nfields < 0) return EKAT_B200_ERR_INVALID;
  if (ncol == 0 || km2 == 0 || nfields == 0) return EKAT_B200_OK;
  if (km1 < 2) return EKAT_B200_ERR_INVALID;
  if (pitch1 < km1 || pitch2 < km2) return EKAT_B200_ERR_INVALID;
  if (!x1 || !x2 || !y1 || !y2) return EKAT_B200_ERR_INVALID;
  if (nfields > 1 && (y1_fstride < ncol * pitch1 || y2_fstride < ncol * pitch2)) return EKAT_B200_ERR_INVALID;
  if (ncol > kStaticMapInts / km2) {  // ncol * km2 > kStaticMapInts, without overflow
    set_error_text("ekat_b200_lin_interp_static_map: ncol*km2 exceeds EKAT_B200_STATIC_MAP_MAX_ENTRIES (4194304); "
                   "use ekat_b200_lin_interp_* with a caller-owned map");
    return EKAT_B200_ERR_TOO_LARGE;
  }
  int* map = nullptr;
  int rc = static_map_address(&map);
  if (rc != EKAT_B200_OK) return rc;
  rc = lin_interp<R>(1, x1, x2, nullptr, nullptr, map, nullptr, ncol, km1, km2, pitch1, pitch2, 0, 0, 0, stream);
  if (rc != EKAT_B200_OK) return rc;
  return lin_interp<R>(2, x1, x2, y1, y2, nullptr, map, ncol, km1, km2, pitch1, pitch2, nfields, y1_fstride, y2_fstride,
                       stream);
}

}  // namespace ekb

extern "C" {

int ekat_b200_lin_interp_static_map_f64(const double* x1, const double* x2, const double* y1, double* y2, int64_t ncol,
                                        int km1, int km2, int64_t pitch1, int64_t pitch2, int nfields,
                                        int64_t y1_fstride, int64_t y2_fstride, ekat_b200_stream_t s) {
  return ekb::lin_interp_static_map<double>(x1, x2, y1, y2, ncol, km1, km2, pitch1, pitch2, nfields, y1_fstride,
                                            y2_fstride, reinterpret_cast<cudaStream_t>(s));
}
int ekat_b200_lin_interp_static_map_f32(const float* x1, const float* x2, const float* y1, float* y2, int64_t ncol,
                                        int km1, int km2, int64_t pitch1, int64_t pitch2, int nfields,
                                        int64_t y1_fstride, int64_t y2_fstride, ekat_b200_stream_t s) {
  return ekb::lin_interp_static_map<float>(x1, x2, y1, y2, ncol, km1, km2, pitch1, pitch2, nfields, y1_fstride,
                                           y2_fstride, reinterpret_cast<cudaStream_t>(s));
}

}  // extern "C"
