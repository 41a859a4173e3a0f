// GPU check of the C++ host adapters (include/ekat_b200/ekat_tridiag.hpp, ekat_lin_interp.hpp) driven the way a
// reference user would drive them: Kokkos-style views (oracle/standin's look-alike, wrapping device memory),
// LayoutRight = the reference's batches, LayoutLeft = native level-major.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include <cuda_runtime.h>
#include <Kokkos_Core.hpp>

#include "ekat_b200/ekat_lin_interp.hpp"
#include "ekat_b200/ekat_tridiag.hpp"

#define CK(x)                                                                  \
  do {                                                                         \
    cudaError_t e_ = (x);                                                      \
    if (e_ != cudaSuccess) {                                                   \
      std::printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
      std::exit(2);                                                            \
    }                                                                          \
  } while (0)

template <class S> using VR2 = Kokkos::View<S**, Kokkos::LayoutRight, Kokkos::MemoryUnmanaged>;
template <class S> using VR3 = Kokkos::View<S***, Kokkos::LayoutRight, Kokkos::MemoryUnmanaged>;
template <class S> using VL2 = Kokkos::View<S**, Kokkos::LayoutLeft, Kokkos::MemoryUnmanaged>;
template <class S> using VL3 = Kokkos::View<S***, Kokkos::LayoutLeft, Kokkos::MemoryUnmanaged>;

// minimal duck-typed strided-rows view: what a Kokkos subview(A, ALL, k, ALL) of A[col][3][lev] looks like
struct Slice {
  using value_type = double;
  using array_layout = Kokkos::LayoutRight;
  static constexpr int rank = 2;
  double* p; int n0, n1; size_t s0;
  double* data() const { return p; }
  int extent_int(int i) const { return i == 0 ? n0 : n1; }
  size_t stride(int i) const { return i == 0 ? s0 : 1; }
};

template <class T>
static T* dalloc(size_t n) {
  T* p;
  CK(cudaMalloc(&p, n * sizeof(T)));
  return p;
}
template <class T>
static std::vector<T> tohost(const T* p, size_t n) {
  std::vector<T> h(n);
  CK(cudaMemcpy(h.data(), p, n * sizeof(T), cudaMemcpyDeviceToHost));
  return h;
}

static int fails = 0;
#define EXPECT(c, msg)                       \
  do {                                       \
    if (!(c)) {                              \
      std::printf("FAIL: %s\n", msg);        \
      ++fails;                               \
    }                                        \
  } while (0)

int main() {
  using ekat::tridiag::Batched;
  const int ncol = 1000, nlev = 72, nrhs = 4;
  // ---- tridiag: LayoutRight (reference batches) vs LayoutLeft (native), 1 A x nrhs -------------------------
  const size_t na = (size_t)ncol * nlev, nx = na * nrhs;
  double *Ar = dalloc<double>(3 * na), *Xr = dalloc<double>(nx), *Br = dalloc<double>(nx);
  double *Al = dalloc<double>(3 * na), *Xl = dalloc<double>(nx), *Bl = dalloc<double>(nx);
  // reference-layout fill: A[col][3][lev][1] -> dl = A, d = A + nlev, du = A + 2 nlev, stride 3 nlev
  EXPECT(ekat_b200_fill_tridiag_f64(Ar, Ar + nlev, Ar + 2 * nlev, Xr, ncol, 0, nlev, 1, nrhs, 3 * nlev, nlev * nrhs,
                                    EKAT_B200_LAYOUT_COL_MAJOR, 32768, nullptr) == 0, "fill col-major");
  EXPECT(ekat_b200_fill_tridiag_f64(Al, Al + na, Al + 2 * na, Xl, ncol, 0, nlev, 1, nrhs, ncol, ncol,
                                    EKAT_B200_LAYOUT_LEV_MAJOR, 32768, nullptr) == 0, "fill lev-major");
  CK(cudaMemcpy(Br, Xr, nx * sizeof(double), cudaMemcpyDeviceToDevice));
  CK(cudaMemcpy(Bl, Xl, nx * sizeof(double), cudaMemcpyDeviceToDevice));
  {
    // strided LayoutRight views into A[col][3][lev]: build with explicit strides through LayoutStride-free trick:
    // the stand-in view takes (ptr, n0, n1) with stride n1, so slice per diagonal by using a [ncol][3*nlev] parent.
    VR2<double> parent(Ar, ncol, 3 * nlev);
    Slice dl{Ar, ncol, nlev, (size_t)3 * nlev}, d{Ar + nlev, ncol, nlev, (size_t)3 * nlev}, du{Ar + 2 * nlev, ncol, nlev, (size_t)3 * nlev};
    VR3<double> X(Xr, ncol, nlev, nrhs);
    ekat::tridiag::bfb(Batched{}, dl, d, du, X);
    (void)parent;
  }
  {
    VL2<double> dl(Al, ncol, nlev), d(Al + na, ncol, nlev), du(Al + 2 * na, ncol, nlev);
    VL3<double> X(Xl, ncol, nlev, nrhs);
    ekat::tridiag::bfb(Batched{}, dl, d, du, X);
  }
  CK(cudaDeviceSynchronize());
  {
    auto xr = tohost(Xr, nx), xl = tohost(Xl, nx);
    bool same = true;
    for (int c = 0; c < ncol && same; ++c)
      for (int i = 0; i < nlev && same; ++i)
        for (int j = 0; j < nrhs; ++j)
          if (std::memcmp(&xr[((size_t)c * nlev + i) * nrhs + j], &xl[((size_t)j * nlev + i) * ncol + c], 8)) same = false;
    EXPECT(same, "bfb LayoutRight and LayoutLeft results are not bit-identical");
    double* res = dalloc<double>(1);
    CK(cudaMemset(res, 0, 8));
    EXPECT(ekat_b200_tridiag_residual_f64(Al, Al + na, Al + 2 * na, Bl, Xl, ncol, nlev, 1, nrhs, ncol, res, nullptr) == 0, "residual");
    const double r = tohost(res, 1)[0];
    EXPECT(r <= 50 * 2.220446049250313e-16, "bfb residual > 50 eps");
    std::printf("bfb residual %.3e\n", r);
    // thomas and cr through the adapters too (restore X first)
    CK(cudaMemcpy(Xl, Bl, nx * sizeof(double), cudaMemcpyDeviceToDevice));
    VL2<double> dl(Al, ncol, nlev), d(Al + na, ncol, nlev), du(Al + 2 * na, ncol, nlev);
    VL3<double> X(Xl, ncol, nlev, nrhs);
    ekat::tridiag::thomas(Batched{}, dl, d, du, X);
    CK(cudaMemset(res, 0, 8));
    ekat_b200_tridiag_residual_f64(Al, Al + na, Al + 2 * na, Bl, Xl, ncol, nlev, 1, nrhs, ncol, res, nullptr);
    EXPECT(tohost(res, 1)[0] <= 50 * 2.220446049250313e-16, "thomas residual > 50 eps");
    CK(cudaMemcpy(Xr, Br, nx * sizeof(double), cudaMemcpyDeviceToDevice));
    VR2<double> dlr(Ar, ncol, 3 * nlev);  // cr on a contiguous [ncol][nlev] copy
    double* Ac = dalloc<double>(3 * na);
    for (int k = 0; k < 3; ++k)
      CK(cudaMemcpy2D(Ac + k * na, nlev * 8, Ar + k * nlev, 3 * nlev * 8, nlev * 8, ncol, cudaMemcpyDeviceToDevice));
    VR2<double> cdl(Ac, ncol, nlev), cd(Ac + na, ncol, nlev), cdu(Ac + 2 * na, ncol, nlev);
    VR3<double> Xc(Xr, ncol, nlev, nrhs);
    ekat::tridiag::cr(Batched{}, cdl, cd, cdu, Xc);
    CK(cudaDeviceSynchronize());
    auto xc = tohost(Xr, nx);
    double num = 0, den = 0;
    for (int c = 0; c < ncol; ++c)
      for (int i = 0; i < nlev; ++i)
        for (int j = 0; j < nrhs; ++j) {
          const double a = xl[((size_t)j * nlev + i) * ncol + c], b = xc[((size_t)c * nlev + i) * nrhs + j];
          num = std::max(num, std::abs(a - b));
          den = std::max(den, std::abs(a));
        }
    EXPECT(num / den <= 1e-13, "cr differs from bfb by more than 1e-13");
    std::printf("cr vs bfb %.3e\n", num / den);
  }
  // ---- LinInterp: Pack<double,16>-shaped views, setup + lin_interp == fused --------------------------------
  {
    struct Pack16 { using scalar = double; double d[16]; };
    const int km1 = 72, km2 = 128, p1 = 80, p2 = 128;
    double *x1 = dalloc<double>((size_t)ncol * p1), *y1 = dalloc<double>((size_t)ncol * p1), *x2 = dalloc<double>((size_t)ncol * p2);
    double *ya = dalloc<double>((size_t)ncol * p2), *yb = dalloc<double>((size_t)ncol * p2);
    CK(cudaMemset(x1, 0, (size_t)ncol * p1 * 8)); CK(cudaMemset(y1, 0, (size_t)ncol * p1 * 8));
    CK(cudaMemset(ya, 0, (size_t)ncol * p2 * 8)); CK(cudaMemset(yb, 0, (size_t)ncol * p2 * 8));
    EXPECT(ekat_b200_fill_lin_interp_f64(x1, y1, x2, ncol, 0, km1, km2, p1, p2, 0, 20260411ull, nullptr) == 0, "fill li");
    ekat::b200::LinInterp<double, 16> li(ncol, km1, km2);
    EXPECT(li.km1_pack() == 5 && li.km2_pack() == 8, "km_pack");
    VR2<const Pack16> vx1(reinterpret_cast<const Pack16*>(x1), ncol, 5), vx2(reinterpret_cast<const Pack16*>(x2), ncol, 8),
        vy1(reinterpret_cast<const Pack16*>(y1), ncol, 5);
    VR2<Pack16> vya(reinterpret_cast<Pack16*>(ya), ncol, 8), vyb(reinterpret_cast<Pack16*>(yb), ncol, 8);
    li.setup(vx1, vx2);
    li.lin_interp(vx1, vx2, vy1, vya);
    li.setup_and_interp(vx1, vx2, vy1, vyb);
    CK(cudaDeviceSynchronize());
    auto ha = tohost(ya, (size_t)ncol * p2), hb = tohost(yb, (size_t)ncol * p2);
    EXPECT(!std::memcmp(ha.data(), hb.data(), ha.size() * 8), "setup+lin_interp != fused");
    auto hx1 = tohost(x1, (size_t)ncol * p1), hy1 = tohost(y1, (size_t)ncol * p1), hx2 = tohost(x2, (size_t)ncol * p2);
    bool inside = true;  // monotone property: targets inside the range => y2 within [min y1, max y1]
    for (int c = 0; c < ncol; ++c) {
      double lo = 1e300, hi = -1e300;
      for (int k = 0; k < km1; ++k) { lo = std::min(lo, hy1[(size_t)c * p1 + k]); hi = std::max(hi, hy1[(size_t)c * p1 + k]); }
      for (int k = 0; k < km2; ++k) inside = inside && ha[(size_t)c * p2 + k] >= lo && ha[(size_t)c * p2 + k] <= hi;
    }
    EXPECT(inside, "monotone property violated");
  }
  if (fails) {
    std::printf("%d adapter checks FAILED\n", fails);
    return 1;
  }
  std::printf("ALL ADAPTER CHECKS PASSED\n");
  return 0;
}
