// ekat_tridiag.hpp (ekat_b200) -- host adapters that keep the ekat::tridiag entry points
// (reference: src/algorithm/ekat_tridiag.hpp:51-88) on top of the C ABI in ekat_b200.h.
//
// The reference's solvers are team-level device functions called once per column inside the
// caller's kernel:
//     Kokkos::parallel_for(policy, KOKKOS_LAMBDA(const MT& team) {
//       const int ip = team.league_rank();
//       ekat::tridiag::thomas(team, get_diag(A,ip,0), get_diag(A,ip,1), get_diag(A,ip,2), get_xs(X,ip));
//     });                                   // tests/algorithm/tridiag_tests_performance.cpp:218-226
// The drop-in replaces that whole loop by ONE call on the batched views (one more leading
// dimension, the column), tagged with ekat::tridiag::Batched so the overloads cannot collide
// with the reference's:
//     ekat::tridiag::thomas(ekat::tridiag::Batched{stream}, dl, d, du, X);
//
// Views are duck-typed (anything with rank, data(), extent_int(i), stride(i), array_layout and
// value_type -- e.g. Kokkos::View<Real**, LayoutRight|LayoutLeft, CudaSpace>):
//     rank 2  (ncol, nlev)          one matrix / one rhs per column
//     rank 3  (ncol, nlev, n)       n matrices / n right-hand sides per column
//   LayoutRight  -> the reference's column-major batches        (EKAT_B200_LAYOUT_COL_MAJOR)
//   LayoutLeft   -> native level-major, columns fastest          (EKAT_B200_LAYOUT_LEV_MAJOR)
//   LayoutStride is rejected at compile time, as in the reference (:33-35, :97-103).
// Value types may be Real or ekat::Pack<Real,N> (the scalarized extent is used, as ekat::scalarize does,
// src/pack/ekat_pack_kokkos.hpp:166-240); dl,d,du must share one type, X may differ only in pack size (:46-47).
//
// Semantics kept: X = B on input, X = A\B on output, in place. Unlike the reference, (dl,d,du) are
// left untouched unless Batched::clobber is set. Shape violations (the reference's debug asserts,
// :328-330) throw std::invalid_argument; CUDA failures throw std::runtime_error.
#ifndef EKAT_B200_TRIDIAG_HPP
#define EKAT_B200_TRIDIAG_HPP

#include <stdexcept>
#include <string>
#include <type_traits>

#include "ekat_b200.h"

namespace ekat {
namespace tridiag {

struct Batched {
  ekat_b200_stream_t stream = nullptr;
  bool clobber = false;
  Batched() = default;
  explicit Batched(ekat_b200_stream_t s, bool clobber_diags = false) : stream(s), clobber(clobber_diags) {}
};

namespace b200_impl {

template <typename T> struct scalar_of { using type = T; };
template <typename T> struct scalar_of<const T> { using type = typename scalar_of<T>::type; };
// Pack<T,N>-like types expose `scalar` (src/pack/ekat_pack.hpp:219-232)
template <typename T, typename = void> struct pack_scalar { using type = T; };
template <typename T> struct pack_scalar<T, std::void_t<typename T::scalar>> { using type = typename T::scalar; };

template <typename View> struct Geometry {
  using value_type = typename std::remove_const<typename View::value_type>::type;
  using scalar = typename std::remove_const<typename pack_scalar<value_type>::type>::type;
  static constexpr int pack = sizeof(value_type) / sizeof(scalar);
  scalar* ptr;
  long long ncol;
  int nlev, n;
  long long stride;  // column stride (COL_MAJOR) or column pitch (LEV_MAJOR), in scalars
  int layout;
};

template <typename Layout> struct layout_code;

template <typename View>
Geometry<View> geometry(const View& v, const char* what) {
  using G = Geometry<View>;
  using Layout = typename View::array_layout;
  constexpr int rank = View::rank;
  static_assert(rank == 2 || rank == 3, "batched tridiag views are (ncol,nlev) or (ncol,nlev,n)");
  // Same rule as impl::EnableIfCanUsePointer (ekat_tridiag.hpp:97-103): LayoutStride has no layout_code.
  G g;
  g.ptr = const_cast<typename G::scalar*>(reinterpret_cast<const typename G::scalar*>(v.data()));
  g.ncol = v.extent_int(0);
  g.nlev = v.extent_int(1);
  g.n = (rank == 3 ? v.extent_int(2) : 1) * G::pack;
  g.layout = layout_code<Layout>::value;
  if (g.layout == EKAT_B200_LAYOUT_COL_MAJOR) {
    // (c,i,p) at c*stride(0) + i*n + p : rows of one column must be contiguous
    if (g.ncol > 0 && g.nlev > 0 && static_cast<long long>(v.stride(1)) * G::pack != g.n)
      throw std::invalid_argument(std::string(what) + ": LayoutRight column is not contiguous");
    g.stride = g.ncol > 1 ? static_cast<long long>(v.stride(0)) * G::pack : static_cast<long long>(g.nlev) * g.n;
  } else {
    if (G::pack != 1 && rank == 3)
      throw std::invalid_argument(std::string(what) + ": packed values need LayoutRight (pack lanes are the fastest index)");
    if (G::pack != 1) throw std::invalid_argument(std::string(what) + ": LayoutLeft views must be scalar-valued");
    if (g.ncol > 0 && v.stride(0) != 1) throw std::invalid_argument(std::string(what) + ": LayoutLeft column stride must be 1");
    g.stride = g.nlev > 1 ? static_cast<long long>(v.stride(1)) : g.ncol;
    if (rank == 3 && g.n > 1 && static_cast<long long>(v.stride(2)) != g.stride * g.nlev)
      throw std::invalid_argument(std::string(what) + ": LayoutLeft planes are not evenly pitched");
  }
  return g;
}

inline void check(int rc, const char* fn) {
  if (rc == EKAT_B200_OK) return;
  if (rc == EKAT_B200_ERR_INVALID) throw std::invalid_argument(std::string(fn) + ": invalid arguments");
  if (rc == EKAT_B200_ERR_TOO_LARGE) throw std::runtime_error(std::string(fn) + ": column too large for on-chip staging");
  if (rc == EKAT_B200_ERR_CUDA) throw std::runtime_error(std::string(fn) + ": CUDA error: " + ekat_b200_last_cuda_error());
  throw std::runtime_error(std::string(fn) + ": unsupported");
}

using fn_f64 = int (*)(double*, double*, double*, double*, int64_t, int, int, int, int64_t, int64_t, int, unsigned,
                       ekat_b200_stream_t);
using fn_f32 = int (*)(float*, float*, float*, float*, int64_t, int, int, int, int64_t, int64_t, int, unsigned,
                       ekat_b200_stream_t);

template <typename TridiagDiag, typename DataArray>
void solve(fn_f64 f64, fn_f32 f32, const char* name, const Batched& b, const TridiagDiag& dl, const TridiagDiag& d,
           const TridiagDiag& du, const DataArray& X) {
  const auto gl = geometry(dl, "dl"), gd = geometry(d, "d"), gu = geometry(du, "du");
  const auto gx = geometry(X, "X");
  using S = typename decltype(gx)::scalar;
  static_assert(std::is_same<S, typename decltype(gd)::scalar>::value, "A and X must share the scalar type");
  static_assert(std::is_same<S, double>::value || std::is_same<S, float>::value, "float or double only");
  // the reference's asserts (ekat_tridiag.hpp:328-330, :374-379)
  if (gl.ncol != gd.ncol || gu.ncol != gd.ncol || gx.ncol != gd.ncol) throw std::invalid_argument("tridiag: column counts differ");
  if (gl.nlev != gd.nlev || gu.nlev != gd.nlev || gx.nlev != gd.nlev) throw std::invalid_argument("tridiag: nrow differs");
  if (gl.n != gd.n || gu.n != gd.n || gl.stride != gd.stride || gu.stride != gd.stride || gl.layout != gd.layout ||
      gu.layout != gd.layout)
    throw std::invalid_argument("tridiag: dl, d, du must have identical shape and layout");
  if (gx.layout != gd.layout) throw std::invalid_argument("tridiag: A and X must use the same layout");
  // Pack padding makes the trailing extent npack*N: padded matrices/rhs are solved too, exactly as the reference's
  // pack arithmetic does; nA must still be 1 or nrhs.
  if (gd.n != 1 && gd.n != gx.n) throw std::invalid_argument("tridiag: nA must be 1 or nrhs");
  const unsigned flags = b.clobber ? EKAT_B200_CLOBBER_DIAGS : 0u;
  int rc;
  if constexpr (std::is_same<S, double>::value)
    rc = f64(gl.ptr, gd.ptr, gu.ptr, gx.ptr, gd.ncol, gd.nlev, gd.n, gx.n, gd.stride, gx.stride, gd.layout, flags, b.stream);
  else
    rc = f32(gl.ptr, gd.ptr, gu.ptr, gx.ptr, gd.ncol, gd.nlev, gd.n, gx.n, gd.stride, gx.stride, gd.layout, flags, b.stream);
  check(rc, name);
}

}  // namespace b200_impl

// a. Thomas (reference :51-62). FMA/reciprocal arithmetic: agrees to 1e-13 / 1e-5 relative.
template <typename TridiagDiag, typename DataArray>
void thomas(const Batched& b, TridiagDiag dl, TridiagDiag d, TridiagDiag du, DataArray X) {
  b200_impl::solve(&ekat_b200_tridiag_thomas_f64, &ekat_b200_tridiag_thomas_f32, "ekat::tridiag::thomas", b, dl, d, du, X);
}
// c. Cyclic reduction (reference :64-69). Bit-identical to the reference's CR without FMA contraction.
template <typename TridiagDiag, typename DataArray>
void cr(const Batched& b, TridiagDiag dl, TridiagDiag d, TridiagDiag du, DataArray X) {
  b200_impl::solve(&ekat_b200_tridiag_cr_f64, &ekat_b200_tridiag_cr_f32, "ekat::tridiag::cr", b, dl, d, du, X);
}
// bfb (reference :81-88). Bit-identical to the reference without FMA contraction.
template <typename TridiagDiag, typename DataArray>
void bfb(const Batched& b, TridiagDiag dl, TridiagDiag d, TridiagDiag du, DataArray X) {
  b200_impl::solve(&ekat_b200_tridiag_bfb_f64, &ekat_b200_tridiag_bfb_f32, "ekat::tridiag::bfb", b, dl, d, du, X);
}

}  // namespace tridiag
}  // namespace ekat

// Layout codes for Kokkos' (or any look-alike's) layout tags. LayoutStride has no specialisation on purpose.
namespace Kokkos {
struct LayoutRight;
struct LayoutLeft;
}  // namespace Kokkos
namespace ekat { namespace tridiag { namespace b200_impl {
template <> struct layout_code<Kokkos::LayoutRight> { static constexpr int value = EKAT_B200_LAYOUT_COL_MAJOR; };
template <> struct layout_code<Kokkos::LayoutLeft> { static constexpr int value = EKAT_B200_LAYOUT_LEV_MAJOR; };
}}}  // namespace ekat::tridiag::b200_impl

#endif  // EKAT_B200_TRIDIAG_HPP
