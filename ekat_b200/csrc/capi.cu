// capi.cu -- the extern "C" boundary of libekat_b200.so (include/ekat_b200/ekat_b200.h):
// argument validation (the reference's asserts become error codes), layout dispatch,
// per-thread error/tuning state and the process-wide launch counter.
#include <atomic>
#include <cstring>
#include <mutex>

#include "common.cuh"

namespace ekb {

// kernels implemented in the other translation units
template <typename R>
int tridiag_levmajor(bool exact, R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nA, int nrhs,
                     int64_t a_stride, int64_t x_stride, unsigned flags, cudaStream_t stream);
template <typename R>
int tridiag_tile(int algo, R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nA, int nrhs, int64_t a_stride,
                 int64_t x_stride, int layout, unsigned flags, cudaStream_t stream);
template <typename R>
int tridiag_tma_cm(bool exact, R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int64_t a_stride, int64_t x_stride,
                   cudaStream_t stream);
template <typename R>
int tridiag_cm_rhs(bool exact, R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nA, int nrhs, int64_t a_stride,
                   int64_t x_stride, unsigned flags, cudaStream_t stream);
template <typename R>
int tridiag_cr_warp(R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nrhs, int64_t a_stride, int64_t x_stride, int layout,
                    cudaStream_t stream);
template <typename R>
int tridiag_cr_tma(R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int64_t a_stride, int64_t x_stride, cudaStream_t stream);
template <typename R>
int tridiag_tma_shared_cm(bool exact, R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int64_t a_stride, int64_t x_stride,
                          cudaStream_t stream);
template <typename R>
int lin_interp(int mode, const R* x1, const R* x2, const R* y1, R* y2, int* indx, const int* indx_in, int64_t ncol,
               int km1, int km2, int64_t pitch1, int64_t pitch2, int nfields, int64_t y1_fstride, int64_t y2_fstride,
               cudaStream_t stream);
template <typename R>
int lin_interp_lm(const R* x1, const R* x2, const R* y1, R* y2, int64_t ncol, int km1, int km2, int64_t pitch, int nfields,
                  cudaStream_t stream);
template <typename R>
int tridiag_lin_interp(const R* dl, const R* d, const R* du, const R* B, const R* x1, const R* x2, R* Y2, R* X,
                       int64_t ncol, int nlev, int km2, int nrhs, int64_t pitch, cudaStream_t stream);
template <typename R>
int fill_tridiag(R* dl, R* d, R* du, R* X, int64_t ncol, int64_t col0, int nlev, int nA, int nrhs, int64_t a_stride,
                 int64_t x_stride, int layout, int seed_mod, cudaStream_t st);
template <typename R>
int fill_lin_interp(R* x1, R* y1, R* x2, int64_t ncol, int64_t col0, int km1, int km2, int64_t pitch1,
                    int64_t pitch2, int layout, uint64_t seed, cudaStream_t st);
template <typename R>
int col_to_lev(const R* src, R* dst, int64_t ncol, int n, int64_t src_stride, int64_t pitch, cudaStream_t st);
template <typename R>
int lev_to_col(const R* src, R* dst, int64_t ncol, int n, int64_t pitch, int64_t dst_stride, cudaStream_t st);
template <typename R>
int adj_diff(const R* in, R* out, int64_t ncol, int n, int64_t pitch_in, int64_t pitch_out, int col_major, int forward,
             cudaStream_t st);
template <typename R>
int diffusion_solve(const R* z, const R* kd, R* X, int64_t ncol, int nlev, int64_t pitch, cudaStream_t stream);
template <typename R>
int diffusion_matrix(const R* z, const R* kd, R* dl, R* d, R* du, int64_t ncol, int nlev, int64_t pitch, cudaStream_t stream);
template <typename R>
int shift(const R* in, R* out, int64_t ncol, int n, int64_t pitch_in, int64_t pitch_out, int col_major, int s, R fill,
          cudaStream_t st);
template <typename R>
int index_and_shift(const R* a, const int* idx, R* out, R* out_shift, int64_t ncol, int n, int m, int64_t pitch_a,
                    int64_t pitch_idx, int64_t pitch_out, int col_major, int s, cudaStream_t st);
template <typename R>
int transpose3(const R* src, R* dst, int64_t ni, int nk, int nj, int64_t c_stride, int64_t f_pitch, int c2f, cudaStream_t st);
template <typename R>
int repack(const R* src, R* dst, int64_t nrows, int n, int64_t src_stride, int64_t dst_pitch, int zero_pad, cudaStream_t st);
template <typename R>
int view_reduction(const R* in, R* out, int64_t ncol, int begin, int end, int64_t pitch, int col_major, int packn,
                   cudaStream_t st);
template <typename R>
int checksum(const R* a, int64_t nrows, int64_t ncol, int64_t pitch, uint64_t* out, cudaStream_t st);
template <typename R>
int residual(const R* dl, const R* d, const R* du, const R* B, const R* X, int64_t ncol, int nlev, int nA, int nrhs,
             int64_t pitch, double* out, cudaStream_t st);

int debug_trace(long long* out, int n);

template <typename R>
int selftest_div(uint64_t n, uint64_t seed, int mode, uint64_t* mismatches, cudaStream_t st);

// ------------------------------------------------------------------ per-thread / process state
static std::atomic<uint64_t> g_launches{0};
static thread_local char t_err[256] = "";
static thread_local const char* t_kernel = "";
static thread_local Tuning t_tuning = {0, 0, 0};

void note_launch(const char* name) {
  g_launches.fetch_add(1, std::memory_order_relaxed);
  t_kernel = name;
}
int check_cuda(cudaError_t e) {
  if (e == cudaSuccess) return EKAT_B200_OK;
  std::strncpy(t_err, cudaGetErrorString(e), sizeof(t_err) - 1);
  t_err[sizeof(t_err) - 1] = 0;
  return EKAT_B200_ERR_CUDA;
}
void set_error_text(const char* text) {
  std::strncpy(t_err, text, sizeof(t_err) - 1);
  t_err[sizeof(t_err) - 1] = 0;
}
Tuning tuning() { return t_tuning; }

// (kernel, device) -> largest dynamic shared memory size already granted by cudaFuncSetAttribute. The attribute
// belongs to the process, so the table does too (a per-thread table would let one thread lower what another needs).
struct SmemGrant {
  const void* kernel;
  int device;
  size_t bytes;
};
constexpr int kSmemSlots = 128;
static SmemGrant t_grants[kSmemSlots] = {};
static std::mutex g_grant_mutex;
static inline SmemGrant& grant_slot(const void* kernel, int device) {
  const uint64_t h = (reinterpret_cast<uintptr_t>(kernel) >> 4) * 0x9E3779B97F4A7C15ull + (uint64_t)device;
  return t_grants[(h >> 32) % kSmemSlots];
}
bool smem_attr_needed(const void* kernel, size_t bytes) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return true;
  std::lock_guard<std::mutex> lock(g_grant_mutex);
  const SmemGrant& g = grant_slot(kernel, dev);
  return !(g.kernel == kernel && g.device == dev && g.bytes >= bytes);
}
void smem_attr_granted(const void* kernel, size_t bytes) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return;
  std::lock_guard<std::mutex> lock(g_grant_mutex);
  SmemGrant& g = grant_slot(kernel, dev);
  g = {kernel, dev, bytes};
}

enum { SOLVER_THOMAS = 0, SOLVER_BFB = 1, SOLVER_CR = 2 };

template <typename R>
static int tridiag_entry(int solver, R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nA, int nrhs,
                         int64_t a_stride, int64_t x_stride, int layout, unsigned flags, ekat_b200_stream_t stream) {
  if (ncol < 0 || nlev < 1 || nrhs < 1) return EKAT_B200_ERR_INVALID;
  if (nA != 1 && nA != nrhs) return EKAT_B200_ERR_INVALID;  // ekat_tridiag.hpp:18-31, :658-661
  if (layout != EKAT_B200_LAYOUT_COL_MAJOR && layout != EKAT_B200_LAYOUT_LEV_MAJOR) return EKAT_B200_ERR_INVALID;
  if (flags & ~EKAT_B200_CLOBBER_DIAGS) return EKAT_B200_ERR_INVALID;
  if (ncol == 0) return EKAT_B200_OK;
  if (!dl || !d || !du || !X) return EKAT_B200_ERR_INVALID;
  if (layout == EKAT_B200_LAYOUT_LEV_MAJOR) {
    if (a_stride < ncol || x_stride < ncol) return EKAT_B200_ERR_INVALID;
  } else {
    if (a_stride < (int64_t)nlev * nA || x_stride < (int64_t)nlev * nrhs) return EKAT_B200_ERR_INVALID;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (solver == SOLVER_CR) {
    const int var = tuning().variant;
    if (layout == EKAT_B200_LAYOUT_LEV_MAJOR && nA == 1 && nrhs == 1 && !flags && (var == kVarAuto || var == kVarCrPerLevel)) {
      const int rc = tridiag_cr_tma<R>(dl, d, du, X, ncol, nlev, a_stride, x_stride, st);  // many columns (tridiag_cr_tma.cu)
      if (rc != EKAT_B200_ERR_UNSUPPORTED) return rc;
    }
    if (nA == 1 && !flags && var == kVarAuto) {
      // few columns, the reference layout, several right-hand sides: a warp per (column, rhs group), rows in registers,
      // neighbours by shuffle (tridiag_cr_warp.cu)
      const int rc = tridiag_cr_warp<R>(dl, d, du, X, ncol, nlev, nrhs, a_stride, x_stride, layout, st);
      if (rc != EKAT_B200_ERR_UNSUPPORTED) return rc;
    }
    return tridiag_tile<R>(2, dl, d, du, X, ncol, nlev, nA, nrhs, a_stride, x_stride, layout, flags, st);
  }
  if (layout == EKAT_B200_LAYOUT_LEV_MAJOR && tuning().variant != kVarForceTile)
    return tridiag_levmajor<R>(solver == SOLVER_BFB, dl, d, du, X, ncol, nlev, nA, nrhs, a_stride, x_stride, flags, st);
  if (layout == EKAT_B200_LAYOUT_COL_MAJOR && nA == 1 && nrhs == 1 && !flags && var_allows_warp_tile(tuning().variant)) {
    // reference layout, one right-hand side: the TMA kernel reads the columns where they lie (tridiag_tma.cu)
    if (solver == SOLVER_THOMAS && tuning().variant == kVarAuto) {  // f64: factor and solve chains on different warps
      const int rc = tridiag_tma_shared_cm<R>(false, dl, d, du, X, ncol, nlev, a_stride, x_stride, st);
      if (rc != EKAT_B200_ERR_UNSUPPORTED) return rc;
    }
    const int rc = tridiag_tma_cm<R>(solver == SOLVER_BFB, dl, d, du, X, ncol, nlev, a_stride, x_stride, st);
    if (rc != EKAT_B200_ERR_UNSUPPORTED) return rc;
  }
  if (layout == EKAT_B200_LAYOUT_COL_MAJOR && (nrhs > 1 || flags) && tuning().variant == kVarAuto) {
    // reference layout, several right-hand sides (or the in-place side effect): bulk-copied column tiles, one thread
    // per (column, right-hand side) (tridiag_cm_rhs.cu)
    const int rc = tridiag_cm_rhs<R>(solver == SOLVER_BFB, dl, d, du, X, ncol, nlev, nA, nrhs, a_stride, x_stride, flags, st);
    if (rc != EKAT_B200_ERR_UNSUPPORTED) return rc;
  }
  return tridiag_tile<R>(solver == SOLVER_BFB ? 1 : 0, dl, d, du, X, ncol, nlev, nA, nrhs, a_stride, x_stride, layout,
                         flags, st);
}

template <typename R>
static int li_entry(int mode, const R* x1, const R* x2, const R* y1, R* y2, int* indx, const int* indx_in,
                    int64_t ncol, int km1, int km2, int64_t pitch1, int64_t pitch2, int nfields, int64_t y1_fstride,
                    int64_t y2_fstride, ekat_b200_stream_t stream) {
  if (ncol < 0 || km1 < 0 || km2 < 0 || nfields < 0) return EKAT_B200_ERR_INVALID;
  if (ncol == 0 || km2 == 0) return EKAT_B200_OK;  // lin_interp_test.cpp:47-92: empty problem is a no-op
  if (km1 < 2) return EKAT_B200_ERR_INVALID;        // k1ph neighbour (impl :119-131) needs two sources
  if (pitch1 < km1 || pitch2 < km2) return EKAT_B200_ERR_INVALID;
  if (!x1 || !x2) return EKAT_B200_ERR_INVALID;
  if (mode == 1 && !indx) return EKAT_B200_ERR_INVALID;
  if (mode == 2 && !indx_in) return EKAT_B200_ERR_INVALID;
  if (mode != 1 && nfields > 0) {
    if (!y1 || !y2) return EKAT_B200_ERR_INVALID;
    if (nfields > 1 && (y1_fstride < ncol * pitch1 || y2_fstride < ncol * pitch2)) return EKAT_B200_ERR_INVALID;
  }
  return lin_interp<R>(mode, x1, x2, y1, y2, indx, indx_in, ncol, km1, km2, pitch1, pitch2, nfields, y1_fstride,
                       y2_fstride, reinterpret_cast<cudaStream_t>(stream));
}

}  // namespace ekb

using namespace ekb;

#define ST(s) reinterpret_cast<cudaStream_t>(s)

extern "C" {

#define DEF_TRIDIAG(name, T, solver)                                                                          \
  int name(T* dl, T* d, T* du, T* X, int64_t ncol, int nlev, int nA, int nrhs, int64_t a_stride,             \
           int64_t x_stride, int layout, unsigned flags, ekat_b200_stream_t stream) {                         \
    return tridiag_entry<T>(solver, dl, d, du, X, ncol, nlev, nA, nrhs, a_stride, x_stride, layout, flags,    \
                            stream);                                                                          \
  }
DEF_TRIDIAG(ekat_b200_tridiag_thomas_f64, double, SOLVER_THOMAS)
DEF_TRIDIAG(ekat_b200_tridiag_thomas_f32, float, SOLVER_THOMAS)
DEF_TRIDIAG(ekat_b200_tridiag_bfb_f64, double, SOLVER_BFB)
DEF_TRIDIAG(ekat_b200_tridiag_bfb_f32, float, SOLVER_BFB)
DEF_TRIDIAG(ekat_b200_tridiag_cr_f64, double, SOLVER_CR)
DEF_TRIDIAG(ekat_b200_tridiag_cr_f32, float, SOLVER_CR)

#define DEF_LI(T, sfx)                                                                                          \
  int ekat_b200_lin_interp_setup_##sfx(const T* x1, const T* x2, int* indx_map, int64_t ncol, int km1, int km2, \
                                       int64_t pitch1, int64_t pitch2, ekat_b200_stream_t s) {                  \
    return li_entry<T>(1, x1, x2, nullptr, nullptr, indx_map, nullptr, ncol, km1, km2, pitch1, pitch2, 0, 0, 0, \
                       s);                                                                                      \
  }                                                                                                             \
  int ekat_b200_lin_interp_apply_##sfx(const T* x1, const T* x2, const T* y1, T* y2, const int* indx_map,       \
                                       int64_t ncol, int km1, int km2, int64_t pitch1, int64_t pitch2,          \
                                       int nfields, int64_t y1_fstride, int64_t y2_fstride,                     \
                                       ekat_b200_stream_t s) {                                                  \
    return li_entry<T>(2, x1, x2, y1, y2, nullptr, indx_map, ncol, km1, km2, pitch1, pitch2, nfields,           \
                       y1_fstride, y2_fstride, s);                                                              \
  }                                                                                                             \
  int ekat_b200_lin_interp_##sfx(const T* x1, const T* x2, const T* y1, T* y2, int* indx_map, int64_t ncol,     \
                                 int km1, int km2, int64_t pitch1, int64_t pitch2, int nfields,                 \
                                 int64_t y1_fstride, int64_t y2_fstride, ekat_b200_stream_t s) {                \
    return li_entry<T>(0, x1, x2, y1, y2, indx_map, nullptr, ncol, km1, km2, pitch1, pitch2, nfields,           \
                       y1_fstride, y2_fstride, s);                                                              \
  }
DEF_LI(double, f64)
DEF_LI(float, f32)

#define DEF_FUSED(T, sfx)                                                                                     \
  int ekat_b200_tridiag_lin_interp_##sfx(const T* dl, const T* d, const T* du, const T* B, const T* x1,       \
                                         const T* x2, T* Y2, T* X, int64_t ncol, int nlev, int km2, int nrhs, \
                                         int64_t pitch, ekat_b200_stream_t stream) {                          \
    if (ncol < 0 || nlev < 2 || km2 < 0 || nrhs < 1) return EKAT_B200_ERR_INVALID;                            \
    if (ncol == 0) return EKAT_B200_OK;                                                                       \
    if (pitch < ncol || !dl || !d || !du || !B || !x1 || (km2 > 0 && (!x2 || !Y2)))                           \
      return EKAT_B200_ERR_INVALID;                                                                           \
    return tridiag_lin_interp<T>(dl, d, du, B, x1, x2, Y2, X, ncol, nlev, km2, nrhs, pitch, ST(stream));      \
  }
DEF_FUSED(double, f64)
DEF_FUSED(float, f32)

#define DEF_LI_LM(T, sfx)                                                                                      \
  int ekat_b200_lin_interp_lm_##sfx(const T* x1, const T* x2, const T* y1, T* y2, int64_t ncol, int km1,       \
                                    int km2, int64_t pitch, int nfields, ekat_b200_stream_t stream) {          \
    if (ncol < 0 || km1 < 0 || km2 < 0 || nfields < 0) return EKAT_B200_ERR_INVALID;                           \
    if (ncol == 0 || km2 == 0 || nfields == 0) return EKAT_B200_OK;                                            \
    if (km1 < 2 || pitch < ncol || !x1 || !x2 || !y1 || !y2) return EKAT_B200_ERR_INVALID;                     \
    return lin_interp_lm<T>(x1, x2, y1, y2, ncol, km1, km2, pitch, nfields, ST(stream));                       \
  }
DEF_LI_LM(double, f64)
DEF_LI_LM(float, f32)

#define DEF_UTIL(T, sfx)                                                                                       \
  int ekat_b200_col_to_lev_##sfx(const T* src, T* dst, int64_t ncol, int n, int64_t src_stride, int64_t pitch, \
                                 ekat_b200_stream_t s) {                                                       \
    if (ncol < 0 || n < 0 || src_stride < n || pitch < ncol) return EKAT_B200_ERR_INVALID;                     \
    if (ncol && n && (!src || !dst)) return EKAT_B200_ERR_INVALID;                                             \
    return col_to_lev<T>(src, dst, ncol, n, src_stride, pitch, ST(s));                                         \
  }                                                                                                            \
  int ekat_b200_lev_to_col_##sfx(const T* src, T* dst, int64_t ncol, int n, int64_t pitch, int64_t dst_stride, \
                                 ekat_b200_stream_t s) {                                                       \
    if (ncol < 0 || n < 0 || dst_stride < n || pitch < ncol) return EKAT_B200_ERR_INVALID;                     \
    if (ncol && n && (!src || !dst)) return EKAT_B200_ERR_INVALID;                                             \
    return lev_to_col<T>(src, dst, ncol, n, pitch, dst_stride, ST(s));                                         \
  }                                                                                                            \
  int ekat_b200_adj_diff_##sfx(const T* in, T* out, int64_t ncol, int n, int64_t pitch_in, int64_t pitch_out,  \
                               ekat_b200_layout layout, int forward, ekat_b200_stream_t s) {                  \
    if (ncol < 0 || n < 0) return EKAT_B200_ERR_INVALID;                                                       \
    if (layout != EKAT_B200_LAYOUT_COL_MAJOR && layout != EKAT_B200_LAYOUT_LEV_MAJOR) return EKAT_B200_ERR_INVALID; \
    const int64_t need = layout == EKAT_B200_LAYOUT_COL_MAJOR ? (int64_t)n : ncol;                             \
    if (pitch_in < need || pitch_out < need) return EKAT_B200_ERR_INVALID;                                     \
    if (ncol && n && (!in || !out || in == out)) return EKAT_B200_ERR_INVALID;                                 \
    return adj_diff<T>(in, out, ncol, n, pitch_in, pitch_out, layout == EKAT_B200_LAYOUT_COL_MAJOR, forward,   \
                       ST(s));                                                                                 \
  }                                                                                                            \
  int ekat_b200_diffusion_solve_##sfx(const T* z, const T* kd, T* X, int64_t ncol, int nlev, int64_t pitch,    \
                                      ekat_b200_stream_t s) {                                                  \
    if (ncol < 0 || nlev < 1 || pitch < ncol) return EKAT_B200_ERR_INVALID;                                    \
    if (ncol == 0) return EKAT_B200_OK;                                                                        \
    if (!z || !kd || !X) return EKAT_B200_ERR_INVALID;                                                         \
    return diffusion_solve<T>(z, kd, X, ncol, nlev, pitch, ST(s));                                             \
  }                                                                                                            \
  int ekat_b200_diffusion_matrix_##sfx(const T* z, const T* kd, T* dl, T* d, T* du, int64_t ncol, int nlev,    \
                                       int64_t pitch, ekat_b200_stream_t s) {                                  \
    if (ncol < 0 || nlev < 1 || pitch < ncol) return EKAT_B200_ERR_INVALID;                                    \
    if (ncol && (!z || !kd || !dl || !d || !du)) return EKAT_B200_ERR_INVALID;                                 \
    return diffusion_matrix<T>(z, kd, dl, d, du, ncol, nlev, pitch, ST(s));                                    \
  }                                                                                                            \
  int ekat_b200_shift_##sfx(const T* in, T* out, int64_t ncol, int n, int64_t pitch_in, int64_t pitch_out,     \
                            ekat_b200_layout layout, int shift_by, T fill, ekat_b200_stream_t s) {             \
    if (ncol < 0 || n < 0) return EKAT_B200_ERR_INVALID;                                                       \
    if (layout != EKAT_B200_LAYOUT_COL_MAJOR && layout != EKAT_B200_LAYOUT_LEV_MAJOR) return EKAT_B200_ERR_INVALID; \
    const int64_t need = layout == EKAT_B200_LAYOUT_COL_MAJOR ? (int64_t)n : ncol;                             \
    if (pitch_in < need || pitch_out < need) return EKAT_B200_ERR_INVALID;                                     \
    if (ncol && n && (!in || !out || in == out)) return EKAT_B200_ERR_INVALID;                                 \
    return shift<T>(in, out, ncol, n, pitch_in, pitch_out, layout == EKAT_B200_LAYOUT_COL_MAJOR, shift_by,     \
                    fill, ST(s));                                                                              \
  }                                                                                                            \
  int ekat_b200_index_and_shift_##sfx(const T* a, const int* idx, T* out, T* out_shift, int64_t ncol, int n,   \
                                      int m, int64_t pitch_a, int64_t pitch_idx, int64_t pitch_out,            \
                                      ekat_b200_layout layout, int shift_by, ekat_b200_stream_t s) {          \
    if (ncol < 0 || n < 0 || m < 0) return EKAT_B200_ERR_INVALID;                                              \
    if (layout != EKAT_B200_LAYOUT_COL_MAJOR && layout != EKAT_B200_LAYOUT_LEV_MAJOR) return EKAT_B200_ERR_INVALID; \
    const bool cm = layout == EKAT_B200_LAYOUT_COL_MAJOR;                                                      \
    if (pitch_a < (cm ? (int64_t)n : ncol) || pitch_idx < (cm ? (int64_t)m : ncol) ||                          \
        pitch_out < (cm ? (int64_t)m : ncol))                                                                  \
      return EKAT_B200_ERR_INVALID;                                                                            \
    if (ncol && m && (!a || !idx || !out || !out_shift)) return EKAT_B200_ERR_INVALID;                         \
    return index_and_shift<T>(a, idx, out, out_shift, ncol, n, m, pitch_a, pitch_idx, pitch_out, cm, shift_by, \
                              ST(s));                                                                          \
  }                                                                                                            \
  int ekat_b200_transpose3_##sfx(const T* src, T* dst, int64_t ni, int nk, int nj, int64_t c_stride,           \
                                 int64_t f_pitch, int c2f, ekat_b200_stream_t s) {                             \
    if (ni < 0 || nk < 0 || nj < 0 || c_stride < (int64_t)nk * nj || f_pitch < ni) return EKAT_B200_ERR_INVALID; \
    if (ni && nk && nj && (!src || !dst || src == dst)) return EKAT_B200_ERR_INVALID;                          \
    return transpose3<T>(src, dst, ni, nk, nj, c_stride, f_pitch, c2f, ST(s));                                 \
  }                                                                                                            \
  int ekat_b200_repack_##sfx(const T* src, T* dst, int64_t nrows, int n, int64_t src_stride, int64_t dst_pitch, \
                             int zero_pad, ekat_b200_stream_t s) {                                             \
    if (nrows < 0 || n < 0 || src_stride < n || dst_pitch < n) return EKAT_B200_ERR_INVALID;                   \
    if (nrows && n && (!src || !dst || src == dst)) return EKAT_B200_ERR_INVALID;                              \
    return repack<T>(src, dst, nrows, n, src_stride, dst_pitch, zero_pad, ST(s));                              \
  }                                                                                                            \
  int ekat_b200_view_reduction_##sfx(const T* in, T* out, int64_t ncol, int begin, int end, int64_t pitch,     \
                                     ekat_b200_layout layout, int pack_n, ekat_b200_stream_t s) {             \
    if (ncol < 0 || begin < 0 || end < begin || pack_n < 0) return EKAT_B200_ERR_INVALID;                      \
    if (layout != EKAT_B200_LAYOUT_COL_MAJOR && layout != EKAT_B200_LAYOUT_LEV_MAJOR) return EKAT_B200_ERR_INVALID; \
    const int64_t need = pack_n > 1 ? ((int64_t)(end + pack_n - 1) / pack_n) * pack_n : end;                   \
    if (layout == EKAT_B200_LAYOUT_COL_MAJOR ? pitch < need : pitch < ncol) return EKAT_B200_ERR_INVALID;      \
    if (ncol && (!in || !out)) return EKAT_B200_ERR_INVALID;                                                   \
    return view_reduction<T>(in, out, ncol, begin, end, pitch, layout == EKAT_B200_LAYOUT_COL_MAJOR, pack_n,   \
                             ST(s));                                                                           \
  }                                                                                                            \
  int ekat_b200_checksum_##sfx(const T* a, int64_t nrows, int64_t ncol, int64_t pitch, uint64_t* out,          \
                               ekat_b200_stream_t s) {                                                         \
    if (nrows < 0 || ncol < 0 || pitch < ncol || !out) return EKAT_B200_ERR_INVALID;                           \
    if (nrows && ncol && !a) return EKAT_B200_ERR_INVALID;                                                     \
    return checksum<T>(a, nrows, ncol, pitch, out, ST(s));                                                     \
  }                                                                                                            \
  int ekat_b200_tridiag_residual_##sfx(const T* dl, const T* d, const T* du, const T* B, const T* X,           \
                                       int64_t ncol, int nlev, int nA, int nrhs, int64_t pitch, double* out,   \
                                       ekat_b200_stream_t s) {                                                 \
    if (ncol < 0 || nlev < 1 || nrhs < 1 || (nA != 1 && nA != nrhs) || pitch < ncol || !out)                   \
      return EKAT_B200_ERR_INVALID;                                                                            \
    if (ncol && (!dl || !d || !du || !B || !X)) return EKAT_B200_ERR_INVALID;                                  \
    return residual<T>(dl, d, du, B, X, ncol, nlev, nA, nrhs, pitch, out, ST(s));                              \
  }                                                                                                            \
  int ekat_b200_fill_tridiag_##sfx(T* dl, T* d, T* du, T* X, int64_t ncol, int64_t col0, int nlev, int nA,     \
                                   int nrhs, int64_t a_stride, int64_t x_stride, int layout, int seed_mod,     \
                                   ekat_b200_stream_t s) {                                                     \
    if (ncol < 0 || nlev < 1 || nA < 1 || nrhs < 1 || !dl || !d || !du) return EKAT_B200_ERR_INVALID;          \
    if (layout == EKAT_B200_LAYOUT_LEV_MAJOR ? (a_stride < ncol || (X && x_stride < ncol))                     \
                                             : (a_stride < (int64_t)nlev * nA ||                               \
                                                (X && x_stride < (int64_t)nlev * nrhs)))                       \
      return EKAT_B200_ERR_INVALID;                                                                            \
    return fill_tridiag<T>(dl, d, du, X, ncol, col0, nlev, nA, nrhs, a_stride, x_stride, layout, seed_mod,     \
                           ST(s));                                                                             \
  }                                                                                                            \
  int ekat_b200_fill_lin_interp_##sfx(T* x1, T* y1, T* x2, int64_t ncol, int64_t col0, int km1, int km2,       \
                                      int64_t pitch1, int64_t pitch2, int layout, uint64_t seed,               \
                                      ekat_b200_stream_t s) {                                                  \
    if (ncol < 0 || km1 < 1 || km2 < 0 || !x1) return EKAT_B200_ERR_INVALID;                                   \
    if (layout == EKAT_B200_LAYOUT_LEV_MAJOR ? (pitch1 < ncol || (x2 && pitch2 < ncol))                        \
                                             : (pitch1 < km1 || (x2 && pitch2 < km2)))                         \
      return EKAT_B200_ERR_INVALID;                                                                            \
    return fill_lin_interp<T>(x1, y1, x2, ncol, col0, km1, km2, pitch1, pitch2, layout, seed, ST(s));          \
  }
DEF_UTIL(double, f64)
DEF_UTIL(float, f32)

int ekat_b200_selftest_div_f64(uint64_t n, uint64_t seed, int mode, uint64_t* mismatches, ekat_b200_stream_t s) {
  return mismatches ? selftest_div<double>(n, seed, mode, mismatches, ST(s)) : EKAT_B200_ERR_INVALID;
}
int ekat_b200_selftest_div_f32(uint64_t n, uint64_t seed, int mode, uint64_t* mismatches, ekat_b200_stream_t s) {
  return mismatches ? selftest_div<float>(n, seed, mode, mismatches, ST(s)) : EKAT_B200_ERR_INVALID;
}

int ekat_b200_device_alloc(void** ptr, size_t bytes) {
  if (!ptr) return EKAT_B200_ERR_INVALID;
  *ptr = nullptr;
  if (bytes == 0) return EKAT_B200_OK;
  return check_cuda(cudaMalloc(ptr, bytes));
}
int ekat_b200_device_free(void* ptr) { return ptr ? check_cuda(cudaFree(ptr)) : EKAT_B200_OK; }

int ekat_b200_debug_trace(int64_t* out, int n) {
  if (!out || n <= 0) return EKAT_B200_ERR_INVALID;
  return debug_trace(reinterpret_cast<long long*>(out), n);
}

int ekat_b200_version(void) { return EKAT_B200_VERSION; }
const char* ekat_b200_last_cuda_error(void) { return ekb::t_err; }
uint64_t ekat_b200_launch_count(void) { return ekb::g_launches.load(std::memory_order_relaxed); }
const char* ekat_b200_last_kernel(void) { return ekb::t_kernel; }
void ekat_b200_set_tuning(int cols_per_cta, int prefetch_levels, int variant) {
  ekb::t_tuning = {cols_per_cta, prefetch_levels, variant};
}

}  // extern "C"
