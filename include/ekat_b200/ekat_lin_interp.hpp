// ekat_lin_interp.hpp (ekat_b200) -- host adapter that keeps the ekat::LinInterp interface
// (reference: src/algorithm/ekat_lin_interp.hpp:38-157) on top of the C ABI in ekat_b200.h.
//
// The reference is used from inside a team kernel, one column per team:
//     LinInterp<Real,N> li(ncol, km1, km2);
//     Kokkos::parallel_for(li.policy(), KOKKOS_LAMBDA(const MemberType& team) {
//       const int i = team.league_rank();
//       li.setup(team, subview(x1,i), subview(x2,i));  team.team_barrier();
//       li.lin_interp(team, subview(x1,i), subview(x2,i), subview(y1,i), subview(y2,i));
//     });                                               // tests/algorithm/lin_interp_test.cpp:141-154
// The drop-in keeps the class, its constructor, km1_pack()/km2_pack(), setup() and lin_interp(), but each
// call takes the WHOLE 2D views (ncol, km_pack) and replaces the loop:
//     ekat::b200::LinInterp<Real,N> li(ncol, km1, km2);
//     li.setup(x1, x2);  li.lin_interp(x1, x2, y1, y2);         // or li.setup_and_interp(x1,x2,y1,y2)
// Stream order provides the setup -> lin_interp barrier the reference leaves to the caller (:148).
// Define EKAT_B200_REPLACE_LIN_INTERP before including to also get the name ekat::LinInterp.
//
// Views: rank-2 LayoutRight, value type Real or ekat::Pack<Real,N> (const or not, :71-90 of the reference
// test), extent(1) = npack(km) packs; as in the reference only the first km1/km2 scalars of a row are data.
#ifndef EKAT_B200_LIN_INTERP_HPP
#define EKAT_B200_LIN_INTERP_HPP

#include <stdexcept>
#include <string>
#include <type_traits>

#include "ekat_b200.h"

namespace ekat {
namespace b200 {

template <typename ScalarT, int PackSize>
class LinInterp {
  static_assert(std::is_same<ScalarT, double>::value || std::is_same<ScalarT, float>::value, "float or double only");
  static_assert(PackSize > 0 && (PackSize & (PackSize - 1)) == 0, "pack size must be a power of two (ekat_pack.hpp:635)");

 public:
  using Scalar = ScalarT;
  static constexpr int LI_PACKN = PackSize;

  // reference ctor: src/algorithm/ekat_lin_interp_impl.hpp:12-19 (allocates m_indx_map on the device)
  LinInterp(int ncol, int km1, int km2)
      : m_ncol(ncol), m_km1(km1), m_km2(km2), m_km1_pack((km1 + PackSize - 1) / PackSize),
        m_km2_pack((km2 + PackSize - 1) / PackSize) {
    if (ncol < 0 || km1 < 0 || km2 < 0) throw std::invalid_argument("LinInterp: negative extent");
    check(ekat_b200_device_alloc(reinterpret_cast<void**>(&m_indx_map), sizeof(int) * static_cast<size_t>(ncol) * km2),
          "LinInterp::LinInterp");
  }
  ~LinInterp() { ekat_b200_device_free(m_indx_map); }
  LinInterp(const LinInterp&) = delete;
  LinInterp& operator=(const LinInterp&) = delete;

  int km1_pack() const { return m_km1_pack; }
  int km2_pack() const { return m_km2_pack; }
  int ncol() const { return m_ncol; }
  const int* indx_map() const { return m_indx_map; }  // [ncol][km2] device ints (m_indx_map, :156)

  // setup for every column (reference :82-100)
  template <typename V1, typename V2>
  void setup(const V1& x1, const V2& x2, ekat_b200_stream_t stream = nullptr) const {
    const Scalar* px1 = coord(x1, m_km1_pack, "x1");
    const Scalar* px2 = coord(x2, m_km2_pack, "x2");
    if constexpr (std::is_same<Scalar, double>::value)
      check(ekat_b200_lin_interp_setup_f64(px1, px2, m_indx_map, m_ncol, m_km1, m_km2, pitch1(), pitch2(), stream), "LinInterp::setup");
    else
      check(ekat_b200_lin_interp_setup_f32(px1, px2, m_indx_map, m_ncol, m_km1, m_km2, pitch1(), pitch2(), stream), "LinInterp::setup");
  }

  // lin_interp for every column with the map from setup (reference :102-126)
  template <typename V1, typename V2, typename V3, typename V4>
  void lin_interp(const V1& x1, const V2& x2, const V3& y1, const V4& y2, ekat_b200_stream_t stream = nullptr) const {
    const Scalar* px1 = coord(x1, m_km1_pack, "x1");
    const Scalar* px2 = coord(x2, m_km2_pack, "x2");
    const Scalar* py1 = coord(y1, m_km1_pack, "y1");
    Scalar* py2 = const_cast<Scalar*>(coord(y2, m_km2_pack, "y2"));
    static_assert(!std::is_const<typename V4::value_type>::value, "y2 must be writable");
    if constexpr (std::is_same<Scalar, double>::value)
      check(ekat_b200_lin_interp_apply_f64(px1, px2, py1, py2, m_indx_map, m_ncol, m_km1, m_km2, pitch1(), pitch2(), 1, 0, 0, stream), "LinInterp::lin_interp");
    else
      check(ekat_b200_lin_interp_apply_f32(px1, px2, py1, py2, m_indx_map, m_ncol, m_km1, m_km2, pitch1(), pitch2(), 1, 0, 0, stream), "LinInterp::lin_interp");
  }

  // setup + lin_interp in one kernel; the index map stays on chip (it is still written if keep_map)
  template <typename V1, typename V2, typename V3, typename V4>
  void setup_and_interp(const V1& x1, const V2& x2, const V3& y1, const V4& y2, ekat_b200_stream_t stream = nullptr,
                        bool keep_map = false) const {
    const Scalar* px1 = coord(x1, m_km1_pack, "x1");
    const Scalar* px2 = coord(x2, m_km2_pack, "x2");
    const Scalar* py1 = coord(y1, m_km1_pack, "y1");
    Scalar* py2 = const_cast<Scalar*>(coord(y2, m_km2_pack, "y2"));
    int* map = keep_map ? m_indx_map : nullptr;
    if constexpr (std::is_same<Scalar, double>::value)
      check(ekat_b200_lin_interp_f64(px1, px2, py1, py2, map, m_ncol, m_km1, m_km2, pitch1(), pitch2(), 1, 0, 0, stream), "LinInterp::setup_and_interp");
    else
      check(ekat_b200_lin_interp_f32(px1, px2, py1, py2, map, m_ncol, m_km1, m_km2, pitch1(), pitch2(), 1, 0, 0, stream), "LinInterp::setup_and_interp");
  }

 private:
  int64_t pitch1() const { return static_cast<int64_t>(m_km1_pack) * PackSize; }
  int64_t pitch2() const { return static_cast<int64_t>(m_km2_pack) * PackSize; }

  // (ncol, npack) LayoutRight view of Scalar (then extent(1) is in scalars and must equal npack*N, the repack
  // divisibility rule of src/pack/ekat_pack_kokkos.hpp:266-267) or of Pack<Scalar,PackSize>.
  template <typename V>
  const Scalar* coord(const V& v, int npack, const char* what) const {
    static_assert(V::rank == 2, "LinInterp batched views are (ncol, km_pack)");
    using value_type = typename std::remove_const<typename V::value_type>::type;
    constexpr int lanes = sizeof(value_type) / sizeof(Scalar);
    static_assert(lanes == 1 || lanes == PackSize, "view pack size must be 1 (scalarized) or the LinInterp pack size");
    const long long want = lanes == 1 ? static_cast<long long>(npack) * PackSize : npack;
    if (v.extent_int(0) != m_ncol || v.extent_int(1) != want)
      throw std::invalid_argument(std::string("LinInterp: ") + what + " has the wrong extents (need ncol x npack*N)");
    if (m_ncol > 1 && want > 0 && (static_cast<long long>(v.stride(0)) != want || v.stride(1) != 1))
      throw std::invalid_argument(std::string("LinInterp: ") + what + " must be a contiguous LayoutRight view");
    return reinterpret_cast<const Scalar*>(v.data());
  }

  static void check(int rc, const char* fn) {
    if (rc == EKAT_B200_OK) return;
    if (rc == EKAT_B200_ERR_INVALID) throw std::invalid_argument(std::string(fn) + ": invalid arguments");
    if (rc == EKAT_B200_ERR_CUDA) throw std::runtime_error(std::string(fn) + ": CUDA error: " + ekat_b200_last_cuda_error());
    throw std::runtime_error(std::string(fn) + ": failed");
  }

  int m_ncol, m_km1, m_km2, m_km1_pack, m_km2_pack;
  int* m_indx_map = nullptr;  // [x2_idx] -> x1_idx, per column
};

}  // namespace b200

#ifdef EKAT_B200_REPLACE_LIN_INTERP
template <typename ScalarT, int PackSize, typename DeviceT = void>
using LinInterp = b200::LinInterp<ScalarT, PackSize>;
#endif

}  // namespace ekat

#endif  // EKAT_B200_LIN_INTERP_HPP
