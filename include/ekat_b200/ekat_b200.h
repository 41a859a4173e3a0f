/* ekat_b200.h -- C ABI of libekat_b200.so: B200 (sm_100a) kernels for EKAT's
 * Algorithm-package hot path (batched tridiagonal solvers + LinInterp).
 *
 * The reference has no FFI: the path is a set of header-only *team-level device
 * functions* that a caller invokes once per column from inside its own Kokkos
 * team kernel. A hand-written CUDA kernel cannot be
 * inlined into someone else's kernel, so the drop-in unit is the whole batch
 * idiom
 *     parallel_for(TeamPolicy(ncol, ...), [=](team){ slice column; solver(team, ...); })
 * (tests/algorithm/tridiag_tests_performance.cpp:204-226,
 *  tests/algorithm/lin_interp_test.cpp:141-154), and each entry point below
 * replaces one such loop. The C++ adapters in ekat_tridiag.hpp /
 * ekat_lin_interp.hpp keep the reference's names on top of this ABI.
 *
 * Conventions
 *  - every data pointer is a DEVICE pointer unless the name says "host";
 *  - asynchronous on `stream` (a cudaStream_t; 0 = legacy default stream);
 *  - no allocation and no synchronisation inside; no global DATA state -- the
 *    only process-wide state is bookkeeping: the launch counter, per-thread
 *    last-error / last-kernel strings, per-thread tuning overrides, a
 *    per-thread cache of encoded TMA descriptors keyed by (pointer, shape,
 *    pitch, box), and a debug trace buffer used by one instrumented variant;
 *  - returns EKAT_B200_OK or a negative ekat_b200_status; never throws, never
 *    aborts (the reference's debug `assert`s and EKAT_KERNEL_REQUIRE aborts,
 *    src/algorithm/ekat_tridiag.hpp:184-185,328-330 and
 *    src/pack/ekat_pack_kokkos.hpp:266-267, become EKAT_B200_ERR_INVALID);
 *  - zero-size batches (ncol == 0, or km2 == 0 for LinInterp) are a no-op
 *    (tests/algorithm/lin_interp_test.cpp:47-92);
 *  - diagonal dominance / sorted x1 are the caller's contract, as in the
 *    reference (src/algorithm/ekat_tridiag.hpp:12-17); nothing is checked.
 *
 * There is no CPU fallback anywhere in this library.
 */
#ifndef EKAT_B200_H
#define EKAT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EKAT_B200_VERSION 100 /* 0.1.0 */

typedef struct CUstream_st* ekat_b200_stream_t; /* == cudaStream_t */

typedef enum {
  EKAT_B200_OK = 0,
  EKAT_B200_ERR_INVALID = -1,     /* bad shape / null pointer / misaligned pitch */
  EKAT_B200_ERR_UNSUPPORTED = -2, /* valid in the reference, not built here */
  EKAT_B200_ERR_CUDA = -3,        /* a CUDA runtime call failed; see ekat_b200_last_cuda_error */
  EKAT_B200_ERR_TOO_LARGE = -4    /* one column does not fit the kernel's on-chip staging */
} ekat_b200_status;

/* ------------------------------------------------------------------------
 * Data layouts for the batched tridiagonal solvers.
 *
 * EKAT_B200_LAYOUT_COL_MAJOR -- the reference's: every column's levels are
 *   contiguous (LayoutRight views sliced per column,
 *   tests/algorithm/tridiag_tests_performance.cpp:90-124,181-185).
 *     dl/d/du element (col c, level i, matrix p) at  base[c*a_stride + i*nA   + p]
 *     X       element (col c, level i, rhs j)    at  X   [c*x_stride + i*nrhs + j]
 *   a_stride >= nlev*nA, x_stride >= nlev*nrhs (elements). For the perf-test
 *   array A[ncol][3][nlev][nA]: dl=A, d=A+nlev*nA, du=A+2*nlev*nA, a_stride=3*nlev*nA.
 *
 * EKAT_B200_LAYOUT_LEV_MAJOR -- native: columns are the fastest index, so one
 *   level step of a warp is one coalesced access.
 *     dl/d/du element (c, i, p) at  base[(p*nlev + i)*a_stride + c]
 *     X       element (c, i, j) at  X   [(j*nlev + i)*x_stride + c]
 *   a_stride, x_stride >= ncol = column pitch in elements.
 * ---------------------------------------------------------------------- */
typedef enum { EKAT_B200_LAYOUT_COL_MAJOR = 0, EKAT_B200_LAYOUT_LEV_MAJOR = 1 } ekat_b200_layout;

/* flags */
#define EKAT_B200_CLOBBER_DIAGS 1u /* also write the factorisation back into dl/d, as the reference's
                                      in-place solvers do (ekat_tridiag.hpp:44-47). Default: dl,d,du are
                                      read-only, which saves 2 of 7 words of HBM traffic per level. */

/* ------------------------------------------------------------------------
 * Batched tridiagonal solvers: X <- A^-1 X for ncol independent columns.
 *   nA   = 1 (one matrix per column, shared by its nrhs right-hand sides) or
 *          nrhs (one matrix per right-hand side)       -- problem formats 1-3,
 *          src/algorithm/ekat_tridiag.hpp:18-31; bfb's "nA==1 && nrhs>1" rule :658-661.
 *   dl(0) and du(nlev-1) are never used (:24-27) but must be readable.
 *
 * thomas : replaces the loop around ekat::tridiag::thomas (team :323-334 and
 *          serial :336-381). Fused-multiply-add and reciprocal form allowed;
 *          agrees with the reference to 1e-13 (f64) / 1e-5 (f32) relative.
 * bfb    : replaces the loop around ekat::tridiag::bfb (:632-678). Fixed
 *          operation order, separately rounded * and -, IEEE division:
 *          bit-identical to the reference built with -ffp-contract=off.
 * cr     : replaces the loop around ekat::tridiag::cr (:385-630). Same update
 *          formulas and order as the reference, unfused: bit-identical to the
 *          reference's cyclic reduction built with -ffp-contract=off.
 * ---------------------------------------------------------------------- */
#define EKAT_B200_DECL_TRIDIAG(name, T)                                                              \
  int name(T* dl, T* d, T* du, T* X, int64_t ncol, int nlev, int nA, int nrhs, int64_t a_stride,     \
           int64_t x_stride, int layout, unsigned flags, ekat_b200_stream_t stream)
EKAT_B200_DECL_TRIDIAG(ekat_b200_tridiag_thomas_f64, double);
EKAT_B200_DECL_TRIDIAG(ekat_b200_tridiag_thomas_f32, float);
EKAT_B200_DECL_TRIDIAG(ekat_b200_tridiag_bfb_f64, double);
EKAT_B200_DECL_TRIDIAG(ekat_b200_tridiag_bfb_f32, float);
EKAT_B200_DECL_TRIDIAG(ekat_b200_tridiag_cr_f64, double);
EKAT_B200_DECL_TRIDIAG(ekat_b200_tridiag_cr_f32, float);

/* ------------------------------------------------------------------------
 * LinInterp: 1D vertical linear interpolation of y1(x1) onto x2, per column.
 * Layout is the reference's (src/algorithm/ekat_lin_interp_impl.hpp:12-19,
 * scalarized Pack views, src/pack/ekat_pack_kokkos.hpp:166-279):
 *     x1, y1 : [ncol][pitch1]   only the first km1 entries of a row are data
 *     x2, y2 : [ncol][pitch2]   only the first km2 entries of a row are data
 *   pitch = npack(km)*N elements for a Pack<T,N> view (N = 1: pitch = km).
 *   Padding lanes are neither read as data nor written.
 *   nfields fields sharing (x1, x2): field f of y1 at y1 + f*y1_fstride, of y2
 *   at y2 + f*y2_fstride (elements).
 *   indx_map: [ncol][km2] ints, the bracketing source index of every target
 *   (LinInterp::m_indx_map, ekat_lin_interp.hpp:156).
 *
 * setup  : replaces the loop around LinInterp::setup  (impl :21-46,160-186).
 * apply  : replaces the loop around LinInterp::lin_interp (impl :48-158) with
 *          a map produced by setup.
 * fused  : setup + lin_interp in one pass; the map stays on chip unless
 *          indx_map != NULL.
 * Results are bit-identical to the reference built with -ffp-contract=off
 * (the 4-statement formula :153-156 in that order, IEEE division).
 * km1 >= 2 is required when km2 > 0 (the reference reads x1[k1 +- 1]).
 * ---------------------------------------------------------------------- */
#define EKAT_B200_DECL_LI(T, sfx)                                                                        \
  int ekat_b200_lin_interp_setup_##sfx(const T* x1, const T* x2, int* indx_map, int64_t ncol, int km1,   \
                                       int km2, int64_t pitch1, int64_t pitch2, ekat_b200_stream_t s);   \
  int ekat_b200_lin_interp_apply_##sfx(const T* x1, const T* x2, const T* y1, T* y2, const int* indx_map,\
                                       int64_t ncol, int km1, int km2, int64_t pitch1, int64_t pitch2,   \
                                       int nfields, int64_t y1_fstride, int64_t y2_fstride,              \
                                       ekat_b200_stream_t s);                                            \
  int ekat_b200_lin_interp_##sfx(const T* x1, const T* x2, const T* y1, T* y2, int* indx_map,            \
                                 int64_t ncol, int km1, int km2, int64_t pitch1, int64_t pitch2,         \
                                 int nfields, int64_t y1_fstride, int64_t y2_fstride, ekat_b200_stream_t s)
EKAT_B200_DECL_LI(double, f64);
EKAT_B200_DECL_LI(float, f32);

/* ------------------------------------------------------------------------
 * Fused "diffuse then remap" (BASELINE config 4): for every column solve the
 * shared-A, nrhs-RHS system (thomas arithmetic) and interpolate each solution
 * from source coordinates x1[nlev] onto x2[km2] without the solutions leaving
 * the SM. All arrays LEV_MAJOR with one column pitch `pitch`:
 *   dl,d,du: [nlev][pitch]  B: [nrhs][nlev][pitch]  x1: [nlev][pitch]
 *   x2: [km2][pitch]        Y2: [nrhs][km2][pitch]
 * X (nullable): [nrhs][nlev][pitch] also receives the un-remapped solutions.
 * ---------------------------------------------------------------------- */
int ekat_b200_tridiag_lin_interp_f64(const double* dl, const double* d, const double* du, const double* B,
                                     const double* x1, const double* x2, double* Y2, double* X, int64_t ncol,
                                     int nlev, int km2, int nrhs, int64_t pitch, ekat_b200_stream_t stream);
int ekat_b200_tridiag_lin_interp_f32(const float* dl, const float* d, const float* du, const float* B,
                                     const float* x1, const float* x2, float* Y2, float* X, int64_t ncol,
                                     int nlev, int km2, int nrhs, int64_t pitch, ekat_b200_stream_t stream);

/* LinInterp on LEVEL-MAJOR operands (the tridiagonal kernels' native layout), fused setup + lin_interp for
 * nfields fields sharing the coordinates (ekat_lin_interp.hpp:13-16: setup once, interpolate many):
 *   x1: [km1][pitch]   x2: [km2][pitch]   y1: [nfields][km1][pitch]   y2: [nfields][km2][pitch]
 * Bit-identical to ekat::LinInterp. y2 may alias y1 when km2 == km1. Needs 16-byte aligned x1, y1 and pitch and
 * km1 <= 256 (one TMA box per 32-column tile): EKAT_B200_ERR_UNSUPPORTED otherwise -- transpose with
 * ekat_b200_lev_to_col_* and use ekat_b200_lin_interp_* then. */
int ekat_b200_lin_interp_lm_f64(const double* x1, const double* x2, const double* y1, double* y2, int64_t ncol,
                                int km1, int km2, int64_t pitch, int nfields, ekat_b200_stream_t stream);
int ekat_b200_lin_interp_lm_f32(const float* x1, const float* x2, const float* y1, float* y2, int64_t ncol, int km1,
                                int km2, int64_t pitch, int nfields, ekat_b200_stream_t stream);

/* LinInterp::setup + LinInterp::lin_interp WITHOUT a caller-owned index map and without any allocation.
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it). Arguments, layout and
 * results are those of ekat_b200_lin_interp_* above (bit-identical y2); the index map lives in a file-scope device
 * array of EKAT_B200_STATIC_MAP_MAX_ENTRIES ints (16 MiB, held by every GPU that loads the library until exit):
 *   - ncol*km2 > EKAT_B200_STATIC_MAP_MAX_ENTRIES returns EKAT_B200_ERR_TOO_LARGE (text in
 *     ekat_b200_last_cuda_error()); use ekat_b200_lin_interp_* with your own map then;
 *   - every call on a device shares the array, which keeps the previous call's contents (each call overwrites all
 *     entries it reads, nothing needs clearing); calls that OVERLAP on the GPU -- different streams with no
 *     dependency between them -- race on it and must be serialised by the caller;
 *   - calls no cudaMalloc / cudaFree / allocating library routine (ekat_b200/csrc/no_alloc.cu). */
#define EKAT_B200_STATIC_MAP_MAX_ENTRIES 4194304
int ekat_b200_lin_interp_static_map_f64(const double* x1, const double* x2, const double* y1, double* y2, int64_t ncol,
                                        int km1, int km2, int64_t pitch1, int64_t pitch2, int nfields,
                                        int64_t y1_fstride, int64_t y2_fstride, ekat_b200_stream_t stream);
int ekat_b200_lin_interp_static_map_f32(const float* x1, const float* x2, const float* y1, float* y2, int64_t ncol,
                                        int km1, int km2, int64_t pitch1, int64_t pitch2, int nfields,
                                        int64_t y1_fstride, int64_t y2_fstride, ekat_b200_stream_t stream);

/* ------------------------------------------------------------------------
 * Layout marshalling: [ncol][n] <-> [n][pitch] device
 * transposes at copy speed, so callers holding reference-layout views can use
 * the LEV_MAJOR kernels. src element (c,k) at src[c*src_stride + k].
 * ---------------------------------------------------------------------- */
int ekat_b200_col_to_lev_f64(const double* src, double* dst, int64_t ncol, int n, int64_t src_stride,
                             int64_t pitch, ekat_b200_stream_t stream);
int ekat_b200_col_to_lev_f32(const float* src, float* dst, int64_t ncol, int n, int64_t src_stride,
                             int64_t pitch, ekat_b200_stream_t stream);
int ekat_b200_lev_to_col_f64(const double* src, double* dst, int64_t ncol, int n, int64_t pitch,
                             int64_t dst_stride, ekat_b200_stream_t stream);
int ekat_b200_lev_to_col_f32(const float* src, float* dst, int64_t ncol, int n, int64_t pitch,
                             int64_t dst_stride, ekat_b200_stream_t stream);

/* ------------------------------------------------------------------------
 * ekat::ReductionUtils<ExeSpace,Serialize>::view_reduction for every column
 * (reference src/algorithm/ekat_reduction_utils.hpp:
 * 67-144, 156-179; call form tests/algorithm/reduction_tests.cpp:117-126):
 * out[c] = sum of column c over the SCALAR range [begin, end).
 *   layout COL_MAJOR: in[c*pitch + k] (the reference's [ncol][npack*N] rows);
 *   layout LEV_MAJOR: in[k*pitch + c].
 *   pack_n == 0: Serialize = true  -- element after element, in order;
 *   pack_n >= 1: Serialize = false over Pack<Real,pack_n> -- leading partial
 *                pack, per-lane sums of the full packs added in lane order,
 *                trailing partial pack, exactly as the reference orders them.
 * Either way the order is fixed: bit-identical to the reference's host build.
 * Scalars outside what that order reads (pack padding) are never touched.
 * COL_MAJOR needs pitch >= npack(end)*pack_n; LEV_MAJOR that many rows.
 * ---------------------------------------------------------------------- */
int ekat_b200_view_reduction_f64(const double* in, double* out, int64_t ncol, int begin, int end, int64_t pitch,
                                 ekat_b200_layout layout, int pack_n, ekat_b200_stream_t stream);
int ekat_b200_view_reduction_f32(const float* in, float* out, int64_t ncol, int begin, int end, int64_t pitch,
                                 ekat_b200_layout layout, int pack_n, ekat_b200_stream_t stream);

/* ------------------------------------------------------------------------
 * ekat::adj_diff<Forward> at every index of every column (SURVEY.md section
 * 8f-3; reference src/pack/ekat_pack_kokkos.hpp:100-164; call form
 * tests/pack/pack_kokkos_utils.cpp:1055-1080) -- the finite-difference step
 * in front of a tridiagonal fill:
 *   forward != 0: out[k] = (k < n-1 ? in[k+1] : 0) - in[k]
 *   forward == 0: out[k] = in[k] - (k > 0 ? in[k-1] : 0)
 * over the n scalars of the scalarized column (for Pack<Real,N> views n =
 * npack*N: padding lanes take part, as in the reference). in != out.
 *   layout COL_MAJOR: in[c*pitch_in + k], out[c*pitch_out + k];
 *   layout LEV_MAJOR: in[k*pitch_in + c], out[k*pitch_out + c].
 * ---------------------------------------------------------------------- */
int ekat_b200_adj_diff_f64(const double* in, double* out, int64_t ncol, int n, int64_t pitch_in, int64_t pitch_out,
                           ekat_b200_layout layout, int forward, ekat_b200_stream_t stream);
int ekat_b200_adj_diff_f32(const float* in, float* out, int64_t ncol, int n, int64_t pitch_in, int64_t pitch_out,
                           ekat_b200_layout layout, int forward, ekat_b200_stream_t stream);

/* ------------------------------------------------------------------------
 * ekat::shift_right / ekat::shift_left along every column (SURVEY.md section
 * 8f-3; reference src/pack/ekat_pack.hpp:521-555: applied pack after pack over
 * a scalarized column, the ScalarType overloads supplying the end value):
 *   out[k] = (0 <= k - shift_by < n) ? in[k - shift_by] : fill
 * shift_by = +1 is shift_right, -1 shift_left; any |shift_by| is accepted.
 * in != out. Layouts as for adj_diff.
 *
 * ekat::index_and_shift<Shift> for every index of every column (reference
 * src/pack/ekat_pack_kokkos.hpp:71-98): m indices idx per column into the n
 * scalars of a,
 *   out[j] = a[idx[j]],  out_shift[j] = a[idx[j] + shift_by].
 * Where idx[j] or idx[j] + shift_by lies outside [0, n): out[j] is left as the
 * caller had it and out_shift[j] = quiet NaN -- the reference's debug-build
 * behaviour (:85-93; its release build reads out of bounds).
 *   COL_MAJOR: a[c*pitch_a + k], idx[c*pitch_idx + j], out*[c*pitch_out + j];
 *   LEV_MAJOR: a[k*pitch_a + c], idx[j*pitch_idx + c], out*[j*pitch_out + c].
 * ---------------------------------------------------------------------- */
int ekat_b200_shift_f64(const double* in, double* out, int64_t ncol, int n, int64_t pitch_in, int64_t pitch_out,
                        ekat_b200_layout layout, int shift_by, double fill, ekat_b200_stream_t stream);
int ekat_b200_shift_f32(const float* in, float* out, int64_t ncol, int n, int64_t pitch_in, int64_t pitch_out,
                        ekat_b200_layout layout, int shift_by, float fill, ekat_b200_stream_t stream);
int ekat_b200_index_and_shift_f64(const double* a, const int* idx, double* out, double* out_shift, int64_t ncol,
                                  int n, int m, int64_t pitch_a, int64_t pitch_idx, int64_t pitch_out,
                                  ekat_b200_layout layout, int shift_by, ekat_b200_stream_t stream);
int ekat_b200_index_and_shift_f32(const float* a, const int* idx, float* out, float* out_shift, int64_t ncol, int n,
                                  int m, int64_t pitch_a, int64_t pitch_idx, int64_t pitch_out,
                                  ekat_b200_layout layout, int shift_by, ekat_b200_stream_t stream);

/* ------------------------------------------------------------------------
 * Fused "difference -> (dl, d, du) -> solve" (SURVEY.md section 8f-3): the
 * finite-difference step that column physics runs right before a tridiagonal
 * fill -- ekat::adj_diff<Forward> of the level coordinate (reference
 * src/pack/ekat_pack_kokkos.hpp:139-152) -- the caller's coefficient formula
 * and the Thomas solve in ONE kernel, so that the three diagonals are never
 * written to or read from HBM: 4 words per column-level cross HBM (z, kd, X
 * in, X out) instead of 10 (adj_diff + fill: read 2, write 3; solve: read 4,
 * write 1). The stencil is the implicit vertical-diffusion step (SHOC-style):
 *   w[i]  = kd[i] / (z[i+1] - z[i])        i < nlev-1;   w[-1] = w[nlev-1] = 0
 *   dl[i] = -w[i-1],  du[i] = -w[i],  d[i] = (1 + w[i]) + w[i-1]
 *   X <- A^-1 X                            (thomas arithmetic, tolerance level)
 * z, kd, X: [nlev][pitch] level-major (kd[i] = dt * diffusivity at the
 * interface between levels i and i+1; kd[nlev-1] is not used). The diagonals
 * are formed with separately rounded IEEE operations, so the implied matrix is
 * bit-defined: ekat_b200_diffusion_matrix_* writes exactly that matrix (the
 * two-pass form, and what a caller needs if it wants the matrix itself).
 * diffusion_solve needs 16-byte aligned operands and pitch and 3 planes of
 * nlev x 32 columns on chip (nlev <= 300 in double): EKAT_B200_ERR_UNSUPPORTED
 * otherwise -- use ekat_b200_diffusion_matrix_* + ekat_b200_tridiag_thomas_*.
 * ---------------------------------------------------------------------- */
int ekat_b200_diffusion_solve_f64(const double* z, const double* kd, double* X, int64_t ncol, int nlev,
                                  int64_t pitch, ekat_b200_stream_t stream);
int ekat_b200_diffusion_solve_f32(const float* z, const float* kd, float* X, int64_t ncol, int nlev, int64_t pitch,
                                  ekat_b200_stream_t stream);
int ekat_b200_diffusion_matrix_f64(const double* z, const double* kd, double* dl, double* d, double* du,
                                   int64_t ncol, int nlev, int64_t pitch, ekat_b200_stream_t stream);
int ekat_b200_diffusion_matrix_f32(const float* z, const float* kd, float* dl, float* d, float* du, int64_t ncol,
                                   int nlev, int64_t pitch, ekat_b200_stream_t stream);

/* ------------------------------------------------------------------------
 * Rank-3 layout marshalling: ekat::transpose<c2f|f2c>
 * for [ni][nk][nj] arrays (reference src/kokkos/ekat_math_utils.hpp:124-147),
 * on the device, with pitches:
 *   C side       element (i,k,j) at c[i*c_stride + k*nj + j]      (c_stride >= nk*nj)
 *   Fortran side element (i,k,j) at f[(j*nk + k)*f_pitch + i]     (f_pitch  >= ni)
 * c2f != 0: src is the C side, dst the Fortran side; c2f == 0: the reverse.
 * With (i,k,j) = (column, level, rhs) the C side is the reference's batch
 * X[col][lev][rhs] (tests/algorithm/tridiag_tests_performance.cpp:181-185) and
 * the Fortran side the native X[rhs][lev][col]; nj == 1 is the rank-2 case.
 *
 * ekat_b200_repack: the row repacking of host_to_device / device_to_host
 * (reference src/pack/ekat_pack_kokkos.hpp:332-1035): [nrows][n] scalars with
 * row stride src_stride -> rows of dst_pitch >= n scalars (= npack*N for a
 * view of Pack<T,N>); zero_pad != 0 also writes zeros into the padding lanes,
 * as a freshly allocated Kokkos view has them.
 * ---------------------------------------------------------------------- */
int ekat_b200_transpose3_f64(const double* src, double* dst, int64_t ni, int nk, int nj, int64_t c_stride,
                             int64_t f_pitch, int c2f, ekat_b200_stream_t stream);
int ekat_b200_transpose3_f32(const float* src, float* dst, int64_t ni, int nk, int nj, int64_t c_stride,
                             int64_t f_pitch, int c2f, ekat_b200_stream_t stream);
int ekat_b200_repack_f64(const double* src, double* dst, int64_t nrows, int n, int64_t src_stride, int64_t dst_pitch,
                         int zero_pad, ekat_b200_stream_t stream);
int ekat_b200_repack_f32(const float* src, float* dst, int64_t nrows, int n, int64_t src_stride, int64_t dst_pitch,
                         int zero_pad, ekat_b200_stream_t stream);

/* ------------------------------------------------------------------------
 * Verification reductions for the column-sharded multi-GPU driver (SURVEY.md
 * section 8e). Results are DEVICE scalars so they can go straight into one
 * NCCL all-reduce (ekat::Comm::all_reduce, src/core/ekat_comm.hpp:152-162).
 *  checksum : *out += wrap-around sum of the raw bit patterns (f32 zero-
 *             extended) of the [nrows][ncol] block of a pitched array --
 *             associative, hence identical for any column partition.
 *  residual : *out = max(*out, max_c max|B - A X| / max|B|), the reference's
 *             acceptance norm (tests/algorithm/tridiag_tests.hpp:42-106),
 *             LEV_MAJOR operands, shared A (nA = 1) or nA = nrhs.
 * ---------------------------------------------------------------------- */
int ekat_b200_checksum_f64(const double* a, int64_t nrows, int64_t ncol, int64_t pitch, uint64_t* out,
                           ekat_b200_stream_t stream);
int ekat_b200_checksum_f32(const float* a, int64_t nrows, int64_t ncol, int64_t pitch, uint64_t* out,
                           ekat_b200_stream_t stream);
int ekat_b200_tridiag_residual_f64(const double* dl, const double* d, const double* du, const double* B,
                                   const double* X, int64_t ncol, int nlev, int nA, int nrhs, int64_t pitch,
                                   double* out, ekat_b200_stream_t stream);
int ekat_b200_tridiag_residual_f32(const float* dl, const float* d, const float* du, const float* B,
                                   const float* X, int64_t ncol, int nlev, int nA, int nrhs, int64_t pitch,
                                   double* out, ekat_b200_stream_t stream);

/* ------------------------------------------------------------------------
 * Synthetic inputs, generated on the device, bit-identical to the reference's
 * deterministic fills (tests/algorithm/tridiag_tests.hpp:11-40 as driven at
 * tests/algorithm/tridiag_tests_performance.cpp:190-196) with
 * seed = (col0 + c) mod seed_mod, and to SURVEY.md section 8d's counter-based
 * LinInterp coordinates. X may be NULL. Used by bench.py and the tests only.
 * ---------------------------------------------------------------------- */
int ekat_b200_fill_tridiag_f64(double* dl, double* d, double* du, double* X, int64_t ncol, int64_t col0,
                               int nlev, int nA, int nrhs, int64_t a_stride, int64_t x_stride, int layout,
                               int seed_mod, ekat_b200_stream_t stream);
int ekat_b200_fill_tridiag_f32(float* dl, float* d, float* du, float* X, int64_t ncol, int64_t col0, int nlev,
                               int nA, int nrhs, int64_t a_stride, int64_t x_stride, int layout, int seed_mod,
                               ekat_b200_stream_t stream);
int ekat_b200_fill_lin_interp_f64(double* x1, double* y1, double* x2, int64_t ncol, int64_t col0, int km1,
                                  int km2, int64_t pitch1, int64_t pitch2, int layout, uint64_t seed,
                                  ekat_b200_stream_t stream);
int ekat_b200_fill_lin_interp_f32(float* x1, float* y1, float* x2, int64_t ncol, int64_t col0, int km1,
                                  int km2, int64_t pitch1, int64_t pitch2, int layout, uint64_t seed,
                                  ekat_b200_stream_t stream);

/* ------------------------------------------------------------------------
 * Host-memory entry points: identical argument meaning, but every data pointer
 * is a HOST pointer (pinned memory recommended) -- for callers that today run
 * the reference on host views (tests/algorithm/tridiag_tests_performance.cpp:
 * 260-303). Columns are cut into chunks that flow through a multi-stream
 * upload -> solve -> download pipeline; device staging belongs to an explicit
 * pipeline object (device_bytes split over nslots <= 8 slots, one stream each).
 * Synchronous by default: results are on the host when the call returns.
 * dl/d/du are not modified. solver: 0 thomas, 1 bfb, 2 cr.
 * Deferred mode (ekat_b200_pipeline_set_deferred(p, 1)): a host entry point
 * returns as soon as its chunks are enqueued, so that consecutive calls on one
 * pipeline overlap -- the upload of the next call's first chunks runs under the
 * download of the previous call's last ones (the slots keep rotating across
 * calls; stream order protects their reuse). The host buffers of every
 * enqueued call belong to the pipeline until ekat_b200_pipeline_sync returns.
 * On any error an entry point first drains all of the pipeline's streams (no
 * copy still touches the caller's buffers after an error return) and reports
 * the first error.
 * ---------------------------------------------------------------------- */
typedef struct ekat_b200_pipeline ekat_b200_pipeline;
int ekat_b200_pipeline_create(ekat_b200_pipeline** out, size_t device_bytes, int nslots);
int ekat_b200_pipeline_destroy(ekat_b200_pipeline* p);
int ekat_b200_pipeline_set_deferred(ekat_b200_pipeline* p, int on);
int ekat_b200_pipeline_sync(ekat_b200_pipeline* p); /* waits for every enqueued chunk; first CUDA error, if any */
/* totals since creation: bytes uploaded, bytes downloaded, chunks processed */
int ekat_b200_pipeline_stats(ekat_b200_pipeline* p, uint64_t* h2d_bytes, uint64_t* d2h_bytes, uint64_t* chunks);
int ekat_b200_tridiag_host_f64(ekat_b200_pipeline* p, int solver, const double* dl, const double* d,
                               const double* du, double* X, int64_t ncol, int nlev, int nA, int nrhs,
                               int64_t a_stride, int64_t x_stride, int layout);
int ekat_b200_tridiag_host_f32(ekat_b200_pipeline* p, int solver, const float* dl, const float* d, const float* du,
                               float* X, int64_t ncol, int nlev, int nA, int nrhs, int64_t a_stride,
                               int64_t x_stride, int layout);
int ekat_b200_lin_interp_host_f64(ekat_b200_pipeline* p, const double* x1, const double* x2, const double* y1,
                                  double* y2, int64_t ncol, int km1, int km2, int64_t pitch1, int64_t pitch2,
                                  int nfields, int64_t y1_fstride, int64_t y2_fstride);
int ekat_b200_lin_interp_host_f32(ekat_b200_pipeline* p, const float* x1, const float* x2, const float* y1,
                                  float* y2, int64_t ncol, int km1, int km2, int64_t pitch1, int64_t pitch2,
                                  int nfields, int64_t y1_fstride, int64_t y2_fstride);

/* Self test of the bit-exact division used by bfb / cr / LinInterp (the IEEE divide with its reciprocal
 * hoisted off the dependent chain, csrc/common.cuh) against the compiler's __ddiv_rn / __fdiv_rn on n random
 * operand pairs: mode 0 = random bit patterns (all exponents, denormals, infinities), mode 1 = well-scaled
 * operands. *mismatches (device uint64) += number of bitwise differences; 0 is the only acceptable answer.
 * Modes 2 / 3: the same two populations through the CERTIFIED quotient of the bfb factorisation (reciprocal taken
 * of a predicted pivot up to 2^-22 off, quotient accepted only with its exact remainder below half an ulp): counts
 * quotients that passed the certificate and differ from __ddiv_rn / __fdiv_rn -- again 0 is the only acceptable
 * answer. Modes 4 / 5 (and 6 / 7 with a 2^-52 prediction error) count how often the certificate sends a quotient
 * to the full-range division instead (a rate, not an error). */
int ekat_b200_selftest_div_f64(uint64_t n, uint64_t seed, int mode, uint64_t* mismatches, ekat_b200_stream_t s);
int ekat_b200_selftest_div_f32(uint64_t n, uint64_t seed, int mode, uint64_t* mismatches, ekat_b200_stream_t s);

/* device memory for the host adapters' own state (LinInterp's index map), so that the C++ headers
 * need no CUDA headers: thin wrappers over cudaMalloc / cudaFree on the current device. */
int ekat_b200_device_alloc(void** ptr, size_t bytes);
int ekat_b200_device_free(void* ptr);

/* ------------------------------------------------------------------------
 * Introspection / tuning.
 * ---------------------------------------------------------------------- */
int ekat_b200_version(void);
/* text of the last CUDA error seen by this thread inside the library ("" if none) */
const char* ekat_b200_last_cuda_error(void);
/* number of kernels this library has launched in this process (all threads) */
uint64_t ekat_b200_launch_count(void);
/* name of the last kernel variant chosen by the calling thread (for bench/profiles) */
const char* ekat_b200_last_kernel(void);
/* override launch heuristics (0 = automatic): columns per CTA and load-prefetch depth (levels)
 * for the LEV_MAJOR thomas/bfb kernels. Thread-local. For tuning sweeps only. */
void ekat_b200_set_tuning(int cols_per_cta, int prefetch_levels, int variant);
/* phase trace of the last instrumented launch (variant 7) of the persistent shared-matrix kernel: copies up to
 * n int64 words (2 traced CTAs x 64 work items x 16 counters, clock cycles) to the HOST buffer `out`. Tuning only. */
int ekat_b200_debug_trace(int64_t* out, int n);

#ifdef __cplusplus
}
#endif
#endif /* EKAT_B200_H */
