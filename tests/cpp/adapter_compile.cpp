// Compile-and-link test of the C++ host adapters against a Kokkos::View look-alike (oracle/standin),
// plus the argument checks that need no GPU (run with --no-gpu), plus -- with a GPU and the device
// pointers handed in by the harness (tests/test_gpu_adapters.py uses the .cu twin) -- nothing else.
#include <cstdio>
#include <cstring>
#include <stdexcept>

#include <Kokkos_Core.hpp>  // the stand-in: only its View type is used

#include "ekat_b200/ekat_lin_interp.hpp"
#include "ekat_b200/ekat_pack_kokkos.hpp"
#include "ekat_b200/ekat_reduction_utils.hpp"
#include "ekat_b200/ekat_tridiag.hpp"

template <class S> using VR2 = Kokkos::View<S**, Kokkos::LayoutRight, Kokkos::MemoryUnmanaged>;
template <class S> using VR3 = Kokkos::View<S***, Kokkos::LayoutRight, Kokkos::MemoryUnmanaged>;
template <class S> using VL2 = Kokkos::View<S**, Kokkos::LayoutLeft, Kokkos::MemoryUnmanaged>;
template <class S> using VL3 = Kokkos::View<S***, Kokkos::LayoutLeft, Kokkos::MemoryUnmanaged>;

struct Pack16 {  // Pack<double,16> look-alike (src/pack/ekat_pack.hpp:219-232)
  using scalar = double;
  double d[16];
};

template <class F>
static bool throws_invalid(F f) {
  try {
    f();
  } catch (const std::invalid_argument&) {
    return true;
  } catch (...) {
    return false;
  }
  return false;
}

int main(int argc, char** argv) {
  const bool no_gpu = argc > 1 && !std::strcmp(argv[1], "--no-gpu");
  using ekat::tridiag::Batched;
  double* fake = reinterpret_cast<double*>(0x1000);  // never dereferenced in --no-gpu mode
  float* fakef = reinterpret_cast<float*>(0x1000);
  int bad = 0;

  // empty batches are no-ops in every layout / rank / precision (instantiates all adapter templates)
  {
    VR2<double> a(fake, 0, 72), x(fake, 0, 72);
    ekat::tridiag::thomas(Batched{}, a, a, a, x);
    ekat::tridiag::bfb(Batched{}, a, a, a, x);
    ekat::tridiag::cr(Batched{}, a, a, a, x);
    VR3<double> a3(fake, 0, 72, 1), x3(fake, 0, 72, 10);
    ekat::tridiag::thomas(Batched{}, a3, a3, a3, x3);
    ekat::tridiag::bfb(Batched{}, a, a, a, x3);
    VL2<float> al(fakef, 0, 72), xl(fakef, 0, 72);
    ekat::tridiag::thomas(Batched{}, al, al, al, xl);
    VL3<float> al3(fakef, 0, 72, 4), xl3(fakef, 0, 72, 4);
    ekat::tridiag::cr(Batched{nullptr, true}, al3, al3, al3, xl3);
    VR3<Pack16> ap(reinterpret_cast<Pack16*>(fake), 0, 72, 1), xp(reinterpret_cast<Pack16*>(fake), 0, 72, 1);
    ekat::tridiag::thomas(Batched{}, ap, ap, ap, xp);
  }
  // the reference's asserts become std::invalid_argument
  {
    VR2<double> a(fake, 8, 72), x(fake, 8, 71), x9(fake, 9, 72);
    bad += !throws_invalid([&] { ekat::tridiag::thomas(Batched{}, a, a, a, x); });   // nrow differs (:328-330)
    bad += !throws_invalid([&] { ekat::tridiag::bfb(Batched{}, a, a, a, x9); });     // column count differs
    VR3<double> a2(fake, 8, 72, 2), x3(fake, 8, 72, 3);
    bad += !throws_invalid([&] { ekat::tridiag::cr(Batched{}, a2, a2, a2, x3); });   // nA not in {1, nrhs}
    VL2<double> xl(fake, 8, 72);
    bad += !throws_invalid([&] { ekat::tridiag::thomas(Batched{}, a, a, a, xl); });  // mixed layouts
  }
  // LinInterp: ctor / getters / extents (lin_interp_test.cpp:47-92 runs the API on an empty problem)
  {
    ekat::b200::LinInterp<double, 16> li(0, 0, 0);
    VR2<const double> x1(fake, 0, 0), x2(fake, 0, 0), y1(fake, 0, 0);
    VR2<double> y2(fake, 0, 0);
    li.setup(x1, x2);
    li.lin_interp(x1, x2, y1, y2);
    li.setup_and_interp(x1, x2, y1, y2);
    ekat::b200::LinInterp<float, 1> lf(0, 72, 128);
    if (lf.km1_pack() != 72 || lf.km2_pack() != 128) ++bad;
    ekat::b200::LinInterp<double, 16> lp(0, 72, 128);
    if (lp.km1_pack() != 5 || lp.km2_pack() != 8) ++bad;
    VR2<const Pack16> xp(reinterpret_cast<const Pack16*>(fake), 0, 5), xq(reinterpret_cast<const Pack16*>(fake), 0, 8);
    lp.setup(xp, xq);
    VR2<const double> wrong(fake, 0, 72);  // scalar view must be npack*N = 80 wide
    bad += !throws_invalid([&] { lp.setup(wrong, xq); });
  }
  // ReductionUtils::view_reduction: empty batches in both layouts / pack sizes; range and extent checks
  {
    using V1d = Kokkos::View<double*, Kokkos::LayoutRight, Kokkos::MemoryUnmanaged>;
    using V1f = Kokkos::View<float*, Kokkos::LayoutRight, Kokkos::MemoryUnmanaged>;
    VR2<const double> d0(fake, 0, 128);
    V1d r0(fake, 0);
    ekat::b200::ReductionUtils<>::view_reduction(nullptr, 0, 128, d0, r0);
    ekat::b200::ReductionUtils<false>::view_reduction(nullptr, 3, 100, d0, r0);
    VR2<const Pack16> p0(reinterpret_cast<const Pack16*>(fake), 0, 8);
    ekat::b200::ReductionUtils<false>::view_reduction(nullptr, 0, 128, p0, r0);
    VL2<float> l0(fakef, 0, 72);
    V1f rf(fakef, 0);
    ekat::b200::ReductionUtils<true>::view_reduction(nullptr, 0, 72, l0, rf);
    VR2<const double> d8(fake, 8, 128);
    V1d r8(fake, 8), r9(fake, 9);
    bad += !throws_invalid([&] { ekat::b200::ReductionUtils<>::view_reduction(nullptr, 0, 129, d8, r8); });  // end beyond the view
    bad += !throws_invalid([&] { ekat::b200::ReductionUtils<>::view_reduction(nullptr, 5, 4, d8, r8); });    // end < begin
    bad += !throws_invalid([&] { ekat::b200::ReductionUtils<>::view_reduction(nullptr, 0, 128, d8, r9); });  // results extent
    // adj_diff: empty batches, extents, aliasing
    VR2<double> o0(fake, 0, 128), o8(fake + 4096, 8, 128), o7(fake + 4096, 8, 127), alias(fake, 8, 128);
    ekat::b200::adj_diff<true>(nullptr, d0, o0);
    ekat::b200::adj_diff<false>(nullptr, l0, VL2<float>(fakef, 0, 72));
    VR2<Pack16> q0(reinterpret_cast<Pack16*>(fake), 0, 8);
    ekat::b200::adj_diff<true>(nullptr, p0, q0);
    bad += !throws_invalid([&] { ekat::b200::adj_diff<true>(nullptr, d8, o7); });     // extents differ
    bad += !throws_invalid([&] { ekat::b200::adj_diff<true>(nullptr, d8, alias); });  // in == out
    (void)o8;
  }
  // shift_left / shift_right / index_and_shift / transpose: empty batches, extents, aliasing
  {
    VR2<const double> d0(fake, 0, 128), d8(fake, 8, 128);
    VR2<double> o0(fake, 0, 128), o7(fake + 4096, 8, 127), alias(fake, 8, 128);
    ekat::b200::shift_right(nullptr, 0.0, d0, o0);
    ekat::b200::shift_left(nullptr, 1.5, d0, o0);
    VR2<const Pack16> p0(reinterpret_cast<const Pack16*>(fake), 0, 8);
    VR2<Pack16> q0(reinterpret_cast<Pack16*>(fake), 0, 8);
    ekat::b200::shift_right(nullptr, 0.0, p0, q0);
    bad += !throws_invalid([&] { ekat::b200::shift_left(nullptr, 0.0, d8, o7); });
    bad += !throws_invalid([&] { ekat::b200::shift_right(nullptr, 0.0, d8, alias); });
    using VI2 = Kokkos::View<int**, Kokkos::LayoutRight, Kokkos::MemoryUnmanaged>;
    VI2 i0(reinterpret_cast<int*>(fake), 0, 128), i8(reinterpret_cast<int*>(fake), 8, 64);
    ekat::b200::index_and_shift<1>(nullptr, d0, i0, o0, o0);
    VR2<double> o8(fake + 4096, 8, 128);
    bad += !throws_invalid([&] { ekat::b200::index_and_shift<1>(nullptr, d8, i8, o8, o8); });  // m differs from the outputs
    ekat::b200::transpose<ekat::b200::TransposeDirection::c2f>(nullptr, fake, fake + 64, 0, 72, 10);
    ekat::b200::transpose<ekat::b200::TransposeDirection::f2c>(nullptr, fakef, fakef + 64, 0, 72);
    bad += !throws_invalid([&] { ekat::b200::transpose<ekat::b200::TransposeDirection::c2f>(nullptr, fake, fake + 64, -1, 72, 10); });
  }
  if (bad) {
    std::printf("adapter argument checks FAILED (%d)\n", bad);
    return 1;
  }
  std::printf("adapter argument checks ok\n");
  (void)no_gpu;
  return 0;
}
