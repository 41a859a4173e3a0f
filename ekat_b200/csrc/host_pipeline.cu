// host_pipeline.cu -- the same entry points for callers whose views live in HOST memory.
//
// A reference user on a CPU-only Kokkos build calls the solvers on host views
// (tests/algorithm/tridiag_tests_performance.cpp:260-303). To be a drop-in for that caller,
// these entry points take host pointers and run a column-chunked, multi-stream pipeline:
//   slot k%n :  H2D(chunk k)  ->  kernel(chunk k)  ->  D2H(chunk k)        (one stream per slot)
// so the PCIe upload of chunk k+1, the solve of chunk k and the download of chunk k-1 overlap.
// Device staging is owned by an explicit pipeline object (no hidden global state).
// Pinned host memory is recommended (pageable memory makes the copies synchronous).
#include <new>

#include "common.cuh"

namespace ekb {
template <typename R>
int lin_interp(int mode, const R* x1, const R* x2, const R* y1, R* y2, int* indx, const int* indx_in, int64_t ncol,
               int km1, int km2, int64_t pitch1, int64_t pitch2, int nfields, int64_t y1_fstride, int64_t y2_fstride,
               cudaStream_t stream);
}  // namespace ekb

// a chunk is solved through the same entry points (and kernel selection) a caller with device views uses
static int solve_chunk(int solver, double* dl, double* d, double* du, double* X, int64_t nc, int nlev, int nA, int nrhs,
                       int64_t as, int64_t xs, int layout, cudaStream_t st) {
  auto* s = reinterpret_cast<ekat_b200_stream_t>(st);
  return solver == 2 ? ekat_b200_tridiag_cr_f64(dl, d, du, X, nc, nlev, nA, nrhs, as, xs, layout, 0, s)
                     : solver == 1 ? ekat_b200_tridiag_bfb_f64(dl, d, du, X, nc, nlev, nA, nrhs, as, xs, layout, 0, s)
                                   : ekat_b200_tridiag_thomas_f64(dl, d, du, X, nc, nlev, nA, nrhs, as, xs, layout, 0, s);
}
static int solve_chunk(int solver, float* dl, float* d, float* du, float* X, int64_t nc, int nlev, int nA, int nrhs,
                       int64_t as, int64_t xs, int layout, cudaStream_t st) {
  auto* s = reinterpret_cast<ekat_b200_stream_t>(st);
  return solver == 2 ? ekat_b200_tridiag_cr_f32(dl, d, du, X, nc, nlev, nA, nrhs, as, xs, layout, 0, s)
                     : solver == 1 ? ekat_b200_tridiag_bfb_f32(dl, d, du, X, nc, nlev, nA, nrhs, as, xs, layout, 0, s)
                                   : ekat_b200_tridiag_thomas_f32(dl, d, du, X, nc, nlev, nA, nrhs, as, xs, layout, 0, s);
}

struct ekat_b200_pipeline {
  int nslots = 0;
  size_t slot_bytes = 0;
  unsigned char* dev = nullptr;
  cudaStream_t streams[8] = {};
  uint64_t h2d = 0, d2h = 0, launches = 0;
  int deferred = 0;   // 1: the host entry points return once their chunks are enqueued (ekat_b200_pipeline_sync waits)
  uint64_t next = 0;  // chunk counter across calls: consecutive calls keep rotating through the slots
};

using namespace ekb;

namespace {

int copy2d(void* dst, size_t dpitch, const void* src, size_t spitch, size_t width, size_t rows, cudaMemcpyKind kind,
           cudaStream_t st) {
  if (width == 0 || rows == 0) return EKAT_B200_OK;
  if (dpitch == width && spitch == width) return check_cuda(cudaMemcpyAsync(dst, src, width * rows, kind, st));
  return check_cuda(cudaMemcpy2DAsync(dst, dpitch, src, spitch, width, rows, kind, st));
}

// Columns per chunk: about two chunks per slot so upload / solve / download overlap, never
// below 64 Ki columns (launch + DMA set-up amortisation), never above the slot capacity `cap`.
int64_t chunk_cols(int64_t cap, int64_t ncol, int nslots) {
  const int64_t want = ((ncol + 2 * nslots - 1) / (2 * nslots) + 63) & ~(int64_t)63;
  return tmin<int64_t>(cap, tmax<int64_t>(want, tmin<int64_t>(cap, 1 << 16)));
}

// Every exit of a host entry point goes through here. On an error the copies already enqueued on the other slot
// streams still read and write the caller's host buffers, so all streams are drained before the (first) error is
// returned; on success the streams are drained unless the pipeline is in deferred mode.
int finish(ekat_b200_pipeline* p, int rc) {
  if (rc == EKAT_B200_OK && p->deferred) return EKAT_B200_OK;
  for (int s = 0; s < p->nslots; ++s) {
    const cudaError_t e = cudaStreamSynchronize(p->streams[s]);
    if (rc == EKAT_B200_OK) rc = check_cuda(e);
  }
  return rc;
}

template <typename R>
int tridiag_host(ekat_b200_pipeline* p, int solver, const R* dl, const R* d, const R* du, R* X, int64_t ncol, int nlev,
                 int nA, int nrhs, int64_t a_stride, int64_t x_stride, int layout) {
  if (!p || ncol < 0 || nlev < 1 || nrhs < 1 || (nA != 1 && nA != nrhs) || solver < 0 || solver > 2)
    return EKAT_B200_ERR_INVALID;
  if (ncol == 0) return EKAT_B200_OK;
  if (!dl || !d || !du || !X) return EKAT_B200_ERR_INVALID;
  const bool lm = layout == EKAT_B200_LAYOUT_LEV_MAJOR;
  if (!lm && layout != EKAT_B200_LAYOUT_COL_MAJOR) return EKAT_B200_ERR_INVALID;
  if (lm ? (a_stride < ncol || x_stride < ncol) : (a_stride < (int64_t)nlev * nA || x_stride < (int64_t)nlev * nrhs))
    return EKAT_B200_ERR_INVALID;
  const size_t per_col = (size_t)(3 * nA + nrhs) * nlev * sizeof(R);
  int64_t cc = (int64_t)(p->slot_bytes / per_col);
  cc &= ~(int64_t)63;  // 64-column granularity keeps every device plane 256-byte aligned
  if (cc < 64) return EKAT_B200_ERR_TOO_LARGE;
  cc = chunk_cols(cc, ncol, p->nslots);
  const size_t rowsA = (size_t)nA * nlev, rowsX = (size_t)nrhs * nlev;
  int k = 0;
  for (int64_t c0 = 0; c0 < ncol; c0 += cc, ++k) {
    const int64_t nc = tmin<int64_t>(cc, ncol - c0);
    const int s = (int)((p->next + (uint64_t)k) % (uint64_t)p->nslots);
    cudaStream_t st = p->streams[s];
    R* base = reinterpret_cast<R*>(p->dev + (size_t)s * p->slot_bytes);
    int rc;
    if (lm) {
      // device chunk: [plane][lev][cc] with pitch cc
      R *ddl = base, *dd = ddl + rowsA * cc, *ddu = dd + rowsA * cc, *dx = ddu + rowsA * cc;
      const size_t w = (size_t)nc * sizeof(R), dp = (size_t)cc * sizeof(R);
      if ((rc = copy2d(ddl, dp, dl + c0, a_stride * sizeof(R), w, rowsA, cudaMemcpyHostToDevice, st))) return finish(p, rc);
      if ((rc = copy2d(dd, dp, d + c0, a_stride * sizeof(R), w, rowsA, cudaMemcpyHostToDevice, st))) return finish(p, rc);
      if ((rc = copy2d(ddu, dp, du + c0, a_stride * sizeof(R), w, rowsA, cudaMemcpyHostToDevice, st))) return finish(p, rc);
      if ((rc = copy2d(dx, dp, X + c0, x_stride * sizeof(R), w, rowsX, cudaMemcpyHostToDevice, st))) return finish(p, rc);
      if ((rc = solve_chunk(solver, ddl, dd, ddu, dx, nc, nlev, nA, nrhs, cc, cc, layout, st))) return finish(p, rc);
      if ((rc = copy2d(X + c0, x_stride * sizeof(R), dx, dp, w, rowsX, cudaMemcpyDeviceToHost, st))) return finish(p, rc);
    } else {
      // device chunk: [col][lev][p], tightly packed
      const size_t ea = (size_t)nlev * nA, ex = (size_t)nlev * nrhs;
      R *ddl = base, *dd = ddl + ea * cc, *ddu = dd + ea * cc, *dx = ddu + ea * cc;
      const size_t wa = ea * sizeof(R), wx = ex * sizeof(R);
      if ((rc = copy2d(ddl, wa, dl + c0 * a_stride, a_stride * sizeof(R), wa, nc, cudaMemcpyHostToDevice, st))) return finish(p, rc);
      if ((rc = copy2d(dd, wa, d + c0 * a_stride, a_stride * sizeof(R), wa, nc, cudaMemcpyHostToDevice, st))) return finish(p, rc);
      if ((rc = copy2d(ddu, wa, du + c0 * a_stride, a_stride * sizeof(R), wa, nc, cudaMemcpyHostToDevice, st))) return finish(p, rc);
      if ((rc = copy2d(dx, wx, X + c0 * x_stride, x_stride * sizeof(R), wx, nc, cudaMemcpyHostToDevice, st))) return finish(p, rc);
      if ((rc = solve_chunk(solver, ddl, dd, ddu, dx, nc, nlev, nA, nrhs, ea, ex, layout, st))) return finish(p, rc);
      if ((rc = copy2d(X + c0 * x_stride, x_stride * sizeof(R), dx, wx, wx, nc, cudaMemcpyDeviceToHost, st))) return finish(p, rc);
    }
    p->h2d += (uint64_t)nc * per_col;
    p->d2h += (uint64_t)nc * rowsX * sizeof(R);
    p->launches += 1;
  }
  p->next += (uint64_t)k;
  return finish(p, EKAT_B200_OK);
}

template <typename R>
int lin_interp_host(ekat_b200_pipeline* p, const R* x1, const R* x2, const R* y1, R* y2, int64_t ncol, int km1, int km2,
                    int64_t pitch1, int64_t pitch2, int nfields, int64_t y1_fstride, int64_t y2_fstride) {
  if (!p || ncol < 0 || km1 < 0 || km2 < 0 || nfields < 1) return EKAT_B200_ERR_INVALID;
  if (ncol == 0 || km2 == 0) return EKAT_B200_OK;
  if (km1 < 2 || pitch1 < km1 || pitch2 < km2 || !x1 || !x2 || !y1 || !y2) return EKAT_B200_ERR_INVALID;
  const size_t per_col = ((size_t)pitch1 * (1 + nfields) + (size_t)pitch2 * (1 + nfields)) * sizeof(R);
  int64_t cc = (int64_t)(p->slot_bytes / per_col);
  cc &= ~(int64_t)63;
  if (cc < 64) return EKAT_B200_ERR_TOO_LARGE;
  cc = chunk_cols(cc, ncol, p->nslots);
  int k = 0;
  for (int64_t c0 = 0; c0 < ncol; c0 += cc, ++k) {
    const int64_t nc = tmin<int64_t>(cc, ncol - c0);
    const int s = (int)((p->next + (uint64_t)k) % (uint64_t)p->nslots);
    cudaStream_t st = p->streams[s];
    R* base = reinterpret_cast<R*>(p->dev + (size_t)s * p->slot_bytes);
    R *dx1 = base, *dx2 = dx1 + cc * pitch1, *dy1 = dx2 + cc * pitch2, *dy2 = dy1 + (size_t)nfields * cc * pitch1;
    int rc;
    const size_t b1 = (size_t)nc * pitch1 * sizeof(R), b2 = (size_t)nc * pitch2 * sizeof(R);
    if ((rc = check_cuda(cudaMemcpyAsync(dx1, x1 + c0 * pitch1, b1, cudaMemcpyHostToDevice, st)))) return finish(p, rc);
    if ((rc = check_cuda(cudaMemcpyAsync(dx2, x2 + c0 * pitch2, b2, cudaMemcpyHostToDevice, st)))) return finish(p, rc);
    for (int f = 0; f < nfields; ++f)
      if ((rc = check_cuda(cudaMemcpyAsync(dy1 + (size_t)f * cc * pitch1, y1 + (size_t)f * y1_fstride + c0 * pitch1, b1,
                                           cudaMemcpyHostToDevice, st))))
        return finish(p, rc);
    rc = lin_interp<R>(0, dx1, dx2, dy1, dy2, nullptr, nullptr, nc, km1, km2, pitch1, pitch2, nfields, cc * pitch1,
                       cc * pitch2, st);
    if (rc) return finish(p, rc);
    for (int f = 0; f < nfields; ++f)
      if ((rc = check_cuda(cudaMemcpyAsync(y2 + (size_t)f * y2_fstride + c0 * pitch2, dy2 + (size_t)f * cc * pitch2, b2,
                                           cudaMemcpyDeviceToHost, st))))
        return finish(p, rc);
    p->h2d += (uint64_t)(b1 * (1 + nfields) + b2);
    p->d2h += (uint64_t)b2 * nfields;
    p->launches += 1;
  }
  p->next += (uint64_t)k;
  return finish(p, EKAT_B200_OK);
}

}  // namespace

extern "C" {

int ekat_b200_pipeline_create(ekat_b200_pipeline** out, size_t device_bytes, int nslots) {
  if (!out || nslots < 1 || nslots > 8 || device_bytes < (size_t)nslots * 4096) return EKAT_B200_ERR_INVALID;
  ekat_b200_pipeline* p = new (std::nothrow) ekat_b200_pipeline();
  if (!p) return EKAT_B200_ERR_INVALID;
  p->nslots = nslots;
  p->slot_bytes = (device_bytes / nslots) & ~(size_t)255;
  int rc = check_cuda(cudaMalloc(&p->dev, p->slot_bytes * nslots));
  for (int s = 0; s < nslots && !rc; ++s) rc = check_cuda(cudaStreamCreateWithFlags(&p->streams[s], cudaStreamNonBlocking));
  if (rc) {
    ekat_b200_pipeline_destroy(p);
    return rc;
  }
  *out = p;
  return EKAT_B200_OK;
}

int ekat_b200_pipeline_destroy(ekat_b200_pipeline* p) {
  if (!p) return EKAT_B200_OK;
  for (int s = 0; s < p->nslots; ++s)
    if (p->streams[s]) cudaStreamDestroy(p->streams[s]);
  if (p->dev) cudaFree(p->dev);
  delete p;
  return EKAT_B200_OK;
}

int ekat_b200_pipeline_set_deferred(ekat_b200_pipeline* p, int on) {
  if (!p) return EKAT_B200_ERR_INVALID;
  p->deferred = on ? 1 : 0;
  return EKAT_B200_OK;
}

int ekat_b200_pipeline_sync(ekat_b200_pipeline* p) {
  if (!p) return EKAT_B200_ERR_INVALID;
  int rc = EKAT_B200_OK;
  for (int s = 0; s < p->nslots; ++s) {
    const cudaError_t e = cudaStreamSynchronize(p->streams[s]);
    if (rc == EKAT_B200_OK) rc = check_cuda(e);
  }
  return rc;
}

int ekat_b200_pipeline_stats(ekat_b200_pipeline* p, uint64_t* h2d, uint64_t* d2h, uint64_t* chunks) {
  if (!p) return EKAT_B200_ERR_INVALID;
  if (h2d) *h2d = p->h2d;
  if (d2h) *d2h = p->d2h;
  if (chunks) *chunks = p->launches;
  return EKAT_B200_OK;
}

int ekat_b200_tridiag_host_f64(ekat_b200_pipeline* p, int solver, const double* dl, const double* d, const double* du,
                               double* X, int64_t ncol, int nlev, int nA, int nrhs, int64_t a_stride, int64_t x_stride,
                               int layout) {
  return tridiag_host<double>(p, solver, dl, d, du, X, ncol, nlev, nA, nrhs, a_stride, x_stride, layout);
}
int ekat_b200_tridiag_host_f32(ekat_b200_pipeline* p, int solver, const float* dl, const float* d, const float* du,
                               float* X, int64_t ncol, int nlev, int nA, int nrhs, int64_t a_stride, int64_t x_stride,
                               int layout) {
  return tridiag_host<float>(p, solver, dl, d, du, X, ncol, nlev, nA, nrhs, a_stride, x_stride, layout);
}
int ekat_b200_lin_interp_host_f64(ekat_b200_pipeline* p, const double* x1, const double* x2, const double* y1,
                                  double* y2, int64_t ncol, int km1, int km2, int64_t pitch1, int64_t pitch2,
                                  int nfields, int64_t y1_fstride, int64_t y2_fstride) {
  return lin_interp_host<double>(p, x1, x2, y1, y2, ncol, km1, km2, pitch1, pitch2, nfields, y1_fstride, y2_fstride);
}
int ekat_b200_lin_interp_host_f32(ekat_b200_pipeline* p, const float* x1, const float* x2, const float* y1, float* y2,
                                  int64_t ncol, int km1, int km2, int64_t pitch1, int64_t pitch2, int nfields,
                                  int64_t y1_fstride, int64_t y2_fstride) {
  return lin_interp_host<float>(p, x1, x2, y1, y2, ncol, km1, km2, pitch1, pitch2, nfields, y1_fstride, y2_fstride);
}

}  // extern "C"
