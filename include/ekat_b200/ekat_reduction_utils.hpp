// ekat_reduction_utils.hpp (ekat_b200) -- host adapters that keep the ekat::ReductionUtils entry point and ekat::adj_diff
// (reference: src/algorithm/ekat_reduction_utils.hpp:156-179) on top of the C ABI in ekat_b200.h.
//
// The reference's view_reduction is a team-level device function called once per column inside the caller's
// kernel (tests/algorithm/reduction_tests.cpp:117-126):
//     results(i) = ekat::ReductionUtils<ExeSpace,Serialize>::view_reduction(team, begin, end, subview(data,i));
// The drop-in replaces that whole loop by ONE call on the batched view (one more leading dimension, the column):
//     ekat::b200::ReductionUtils<Serialize>::view_reduction(stream, begin, end, data, results);
//
// `data` is duck-typed (rank 2, data(), extent_int(i), stride(i), array_layout, value_type), value type Real or
// ekat::Pack<Real,N>; [begin, end) is the SCALAR range, as in the reference (:150-155). LayoutRight (ncol, npack)
// is the reference's layout; LayoutLeft (ncol, nlev) the native level-major one. Serialize defaults to true -- the
// reference's default is ekatBFB (src/core/ekat.hpp:14-18) and bit-for-bit reproducibility is this library's
// contract; Serialize = false reproduces the reference's packed order for the view's pack size.
// Shape violations throw std::invalid_argument; CUDA failures throw std::runtime_error.
#ifndef EKAT_B200_REDUCTION_UTILS_HPP
#define EKAT_B200_REDUCTION_UTILS_HPP

#include <stdexcept>
#include <string>
#include <type_traits>

#include "ekat_b200.h"
#include "ekat_tridiag.hpp"  // b200_impl::pack_scalar, layout_code

namespace ekat {
namespace b200 {

template <bool Serialize = true>
struct ReductionUtils {
  template <typename InView, typename OutView>
  static void view_reduction(ekat_b200_stream_t stream, int begin, int end, const InView& data, const OutView& results) {
    using value_type = typename std::remove_const<typename InView::value_type>::type;
    using scalar = typename std::remove_const<typename tridiag::b200_impl::pack_scalar<value_type>::type>::type;
    using out_scalar = typename std::remove_const<typename OutView::value_type>::type;
    static_assert(InView::rank == 2 && OutView::rank == 1, "view_reduction: data is (ncol, nlev), results is (ncol)");
    static_assert(std::is_same<scalar, out_scalar>::value, "view_reduction: result type is the scalar type of the data");
    static_assert(std::is_same<scalar, double>::value || std::is_same<scalar, float>::value, "double or float");
    constexpr int pack = sizeof(value_type) / sizeof(scalar);
    constexpr int layout = tridiag::b200_impl::layout_code<typename InView::array_layout>::value;
    const long long ncol = data.extent_int(0);
    const long long width = (long long)data.extent_int(1) * pack;  // scalars per column
    if (results.extent_int(0) != ncol) throw std::invalid_argument("view_reduction: results extent != number of columns");
    if (begin < 0 || end < begin || end > width) throw std::invalid_argument("view_reduction: [begin,end) outside the view");
    if (layout == EKAT_B200_LAYOUT_LEV_MAJOR && pack != 1)
      throw std::invalid_argument("view_reduction: packed level-major views are not supported");
    if (ncol == 0) return;
    const long long pitch = layout == EKAT_B200_LAYOUT_COL_MAJOR ? (ncol > 1 ? (long long)data.stride(0) * pack : width)
                                                                 : (long long)data.stride(1);
    const scalar* in = reinterpret_cast<const scalar*>(data.data());
    scalar* out = const_cast<scalar*>(reinterpret_cast<const scalar*>(results.data()));
    int rc;
    if constexpr (std::is_same<scalar, double>::value)
      rc = ekat_b200_view_reduction_f64(in, out, ncol, begin, end, pitch, (ekat_b200_layout)layout, Serialize ? 0 : pack, stream);
    else
      rc = ekat_b200_view_reduction_f32(in, out, ncol, begin, end, pitch, (ekat_b200_layout)layout, Serialize ? 0 : pack, stream);
    if (rc == EKAT_B200_ERR_INVALID) throw std::invalid_argument("ekat_b200_view_reduction: invalid argument");
    if (rc != EKAT_B200_OK) throw std::runtime_error(std::string("ekat_b200_view_reduction: ") + ekat_b200_last_cuda_error());
  }
};

// ekat::adj_diff<Forward>(view, index, index_end) at every index of every column (reference
// src/pack/ekat_pack_kokkos.hpp:100-164; tests/pack/pack_kokkos_utils.cpp:1055-1080 loops it over a view):
//     ekat::b200::adj_diff<Forward>(stream, in, out);       in, out: (ncol, nlev) views of Real or Pack<Real,N>
// A Pack view is differenced over its npack*N scalars, padding lanes included, as the reference's shift_left /
// shift_right with a 0 fill do. in and out must not alias.
template <bool Forward, typename InView, typename OutView>
void adj_diff(ekat_b200_stream_t stream, const InView& in, const OutView& out) {
  using value_type = typename std::remove_const<typename InView::value_type>::type;
  using out_type = typename std::remove_const<typename OutView::value_type>::type;
  using scalar = typename std::remove_const<typename tridiag::b200_impl::pack_scalar<value_type>::type>::type;
  static_assert(InView::rank == 2 && OutView::rank == 2, "adj_diff: (ncol, nlev) views");
  static_assert(std::is_same<value_type, out_type>::value, "adj_diff: in and out share a value type");
  static_assert(std::is_same<scalar, double>::value || std::is_same<scalar, float>::value, "double or float");
  static_assert(std::is_same<typename InView::array_layout, typename OutView::array_layout>::value, "adj_diff: one layout");
  constexpr int pack = sizeof(value_type) / sizeof(scalar);
  constexpr int layout = tridiag::b200_impl::layout_code<typename InView::array_layout>::value;
  const long long ncol = in.extent_int(0);
  const int n = in.extent_int(1) * pack;
  if (out.extent_int(0) != ncol || out.extent_int(1) != in.extent_int(1)) throw std::invalid_argument("adj_diff: extents differ");
  if (layout == EKAT_B200_LAYOUT_LEV_MAJOR && pack != 1) throw std::invalid_argument("adj_diff: packed level-major views are not supported");
  if (ncol == 0 || n == 0) return;
  if (static_cast<const void*>(in.data()) == static_cast<const void*>(out.data())) throw std::invalid_argument("adj_diff: in and out alias");
  long long pin, pout;
  if (layout == EKAT_B200_LAYOUT_COL_MAJOR) {
    pin = ncol > 1 ? (long long)in.stride(0) * pack : n;
    pout = ncol > 1 ? (long long)out.stride(0) * pack : n;
  } else {
    pin = in.stride(1);
    pout = out.stride(1);
  }
  const scalar* pi = reinterpret_cast<const scalar*>(in.data());
  scalar* po = const_cast<scalar*>(reinterpret_cast<const scalar*>(out.data()));
  int rc;
  if constexpr (std::is_same<scalar, double>::value)
    rc = ekat_b200_adj_diff_f64(pi, po, ncol, n, pin, pout, (ekat_b200_layout)layout, Forward ? 1 : 0, stream);
  else
    rc = ekat_b200_adj_diff_f32(pi, po, ncol, n, pin, pout, (ekat_b200_layout)layout, Forward ? 1 : 0, stream);
  if (rc == EKAT_B200_ERR_INVALID) throw std::invalid_argument("ekat_b200_adj_diff: invalid argument");
  if (rc != EKAT_B200_OK) throw std::runtime_error(std::string("ekat_b200_adj_diff: ") + ekat_b200_last_cuda_error());
}

}  // namespace b200
}  // namespace ekat

#endif  // EKAT_B200_REDUCTION_UTILS_HPP
