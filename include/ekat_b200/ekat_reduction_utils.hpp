// ekat_reduction_utils.hpp (ekat_b200) -- host adapter that keeps the ekat::ReductionUtils entry point
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

}  // namespace b200
}  // namespace ekat

#endif  // EKAT_B200_REDUCTION_UTILS_HPP
