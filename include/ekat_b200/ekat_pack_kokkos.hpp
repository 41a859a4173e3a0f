// ekat_pack_kokkos.hpp -- host adapters for the column-ops neighbours and layout marshalling of the reference's Pack
// glue on top of the C ABI (ekat_b200.h):
//   ekat::b200::shift_right / shift_left   <- ekat::shift_right / shift_left      (src/pack/ekat_pack.hpp:521-555)
//   ekat::b200::index_and_shift<Shift>     <- ekat::index_and_shift<Shift>        (src/pack/ekat_pack_kokkos.hpp:71-98)
//   ekat::b200::transpose<direction>       <- ekat::transpose<c2f|f2c> rank 2/3   (src/kokkos/ekat_math_utils.hpp:90-147)
// (ekat::b200::adj_diff lives in ekat_reduction_utils.hpp.) Where the reference is called once per pack / per index
// inside the caller's kernel, these take whole (ncol, n) views; views are duck-typed on data() / extent_int() /
// stride() (Kokkos::View with LayoutRight, Real or Pack<Real,N> values). No CPU path: every call ends in the library.
#ifndef EKAT_B200_PACK_KOKKOS_HPP
#define EKAT_B200_PACK_KOKKOS_HPP

#include <stdexcept>
#include <string>
#include <type_traits>

#include "ekat_b200.h"
#include "ekat_tridiag.hpp"  // b200_impl::pack_scalar, layout_code

namespace ekat {
namespace b200 {

namespace pk_impl {
inline void check(int rc, const char* what) {
  if (rc == EKAT_B200_ERR_INVALID) throw std::invalid_argument(std::string(what) + ": invalid argument");
  if (rc != EKAT_B200_OK) throw std::runtime_error(std::string(what) + ": " + ekat_b200_last_cuda_error());
}
template <typename View>
struct Geo {  // (ncol, scalars per column, scalar pitch) of a rank-2 LayoutRight view of Real or Pack<Real,N>
  using value_type = typename std::remove_const<typename View::value_type>::type;
  using scalar = typename std::remove_const<typename tridiag::b200_impl::pack_scalar<value_type>::type>::type;
  static constexpr int pack = sizeof(value_type) / sizeof(scalar);
  long long ncol, pitch;
  int n;
  explicit Geo(const View& v) : ncol(v.extent_int(0)), pitch(0), n(v.extent_int(1) * pack) {
    static_assert(View::rank == 2, "(ncol, n) views");
    pitch = ncol > 1 ? (long long)v.stride(0) * pack : n;
  }
};
template <typename S>
int shift_call(const S* in, S* out, long long ncol, int n, long long pin, long long pout, int by, S fill, ekat_b200_stream_t st);
template <>
inline int shift_call<double>(const double* in, double* out, long long ncol, int n, long long pin, long long pout, int by,
                              double fill, ekat_b200_stream_t st) {
  return ekat_b200_shift_f64(in, out, ncol, n, pin, pout, EKAT_B200_LAYOUT_COL_MAJOR, by, fill, st);
}
template <>
inline int shift_call<float>(const float* in, float* out, long long ncol, int n, long long pin, long long pout, int by, float fill,
                             ekat_b200_stream_t st) {
  return ekat_b200_shift_f32(in, out, ncol, n, pin, pout, EKAT_B200_LAYOUT_COL_MAJOR, by, fill, st);
}
template <typename InView, typename OutView, typename S>
void shift(ekat_b200_stream_t st, const InView& in, const OutView& out, int by, S fill) {
  Geo<InView> gi(in);
  Geo<OutView> go(out);
  using scalar = typename Geo<InView>::scalar;
  static_assert(std::is_same<scalar, typename Geo<OutView>::scalar>::value, "shift: in and out share a scalar type");
  if (gi.ncol != go.ncol || gi.n != go.n) throw std::invalid_argument("shift: extents differ");
  if (gi.ncol == 0 || gi.n == 0) return;
  if (static_cast<const void*>(in.data()) == static_cast<const void*>(out.data())) throw std::invalid_argument("shift: in and out alias");
  check(shift_call<scalar>(reinterpret_cast<const scalar*>(in.data()), const_cast<scalar*>(reinterpret_cast<const scalar*>(out.data())),
                           gi.ncol, gi.n, gi.pitch, go.pitch, by, (scalar)fill, st),
        "ekat_b200_shift");
}
}  // namespace pk_impl

// out(c, k) = shift_right(k > 0 ? in(c, k-1) : fill, in(c, k)) for every pack k of every column c, i.e. every scalar
// one place up its column and `fill` at its start (reference src/pack/ekat_pack.hpp:521-538).
template <typename InView, typename OutView, typename S>
void shift_right(ekat_b200_stream_t stream, S fill, const InView& in, const OutView& out) {
  pk_impl::shift(stream, in, out, +1, fill);
}
// ... one place down its column, `fill` at its end (:540-555).
template <typename InView, typename OutView, typename S>
void shift_left(ekat_b200_stream_t stream, S fill, const InView& in, const OutView& out) {
  pk_impl::shift(stream, in, out, -1, fill);
}

// index(c, j) = a(c, idx(c, j)), index_shift(c, j) = a(c, idx(c, j) + Shift) for every index (pack lane) of every
// column (reference src/pack/ekat_pack_kokkos.hpp:71-98). a: (ncol, n) scalars; idx: (ncol, m) ints or Pack<int,N>;
// index / index_shift: (ncol, m) scalars or Pack<Real,N>. Out-of-range lanes: index untouched, index_shift NaN.
template <int Shift, typename AView, typename IdxView, typename OutView>
void index_and_shift(ekat_b200_stream_t stream, const AView& a, const IdxView& idx, const OutView& index, const OutView& index_shift) {
  pk_impl::Geo<AView> ga(a);
  pk_impl::Geo<OutView> go(index), gs(index_shift);
  using scalar = typename pk_impl::Geo<AView>::scalar;
  static_assert(std::is_same<scalar, typename pk_impl::Geo<OutView>::scalar>::value, "index_and_shift: one scalar type");
  static_assert(IdxView::rank == 2, "index_and_shift: (ncol, m) index view");
  constexpr int ipack = sizeof(typename IdxView::value_type) / sizeof(int);
  const long long ncol = idx.extent_int(0);
  const int m = idx.extent_int(1) * ipack;
  const long long pidx = ncol > 1 ? (long long)idx.stride(0) * ipack : m;
  if (ga.ncol != ncol || go.ncol != ncol || gs.ncol != ncol || go.n != m || gs.n != m || go.pitch != gs.pitch)
    throw std::invalid_argument("index_and_shift: extents differ");
  if (ncol == 0 || m == 0) return;
  const scalar* pa = reinterpret_cast<const scalar*>(a.data());
  const int* pi = reinterpret_cast<const int*>(idx.data());
  scalar* p0 = const_cast<scalar*>(reinterpret_cast<const scalar*>(index.data()));
  scalar* p1 = const_cast<scalar*>(reinterpret_cast<const scalar*>(index_shift.data()));
  int rc;
  if constexpr (std::is_same<scalar, double>::value)
    rc = ekat_b200_index_and_shift_f64(pa, pi, p0, p1, ncol, ga.n, m, ga.pitch, pidx, go.pitch, EKAT_B200_LAYOUT_COL_MAJOR, Shift, stream);
  else
    rc = ekat_b200_index_and_shift_f32(pa, pi, p0, p1, ncol, ga.n, m, ga.pitch, pidx, go.pitch, EKAT_B200_LAYOUT_COL_MAJOR, Shift, stream);
  pk_impl::check(rc, "ekat_b200_index_and_shift");
}

// ekat::transpose<direction>(sv, dv, ni, nk[, nj]) on DEVICE pointers (reference src/kokkos/ekat_math_utils.hpp:97-147).
struct TransposeDirection {
  enum Enum { c2f, f2c };
};
template <TransposeDirection::Enum direction, typename Scalar>
void transpose(ekat_b200_stream_t stream, const Scalar* sv, Scalar* dv, int ni, int nk, int nj = 1) {
  static_assert(std::is_same<Scalar, double>::value || std::is_same<Scalar, float>::value, "double or float");
  if (ni < 0 || nk < 0 || nj < 1) throw std::invalid_argument("transpose: extents");
  const int c2f = direction == TransposeDirection::c2f ? 1 : 0;
  int rc;
  if constexpr (std::is_same<Scalar, double>::value) rc = ekat_b200_transpose3_f64(sv, dv, ni, nk, nj, (long long)nk * nj, ni, c2f, stream);
  else rc = ekat_b200_transpose3_f32(sv, dv, ni, nk, nj, (long long)nk * nj, ni, c2f, stream);
  pk_impl::check(rc, "ekat_b200_transpose3");
}

}  // namespace b200
}  // namespace ekat

#endif  // EKAT_B200_PACK_KOKKOS_HPP
