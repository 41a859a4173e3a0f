// Serial/OpenMP stand-in for the subset of <Kokkos_Core.hpp> that EKAT's
// ekat_tridiag.hpp and ekat_lin_interp.hpp touch.
//
// TEST INFRASTRUCTURE ONLY. This file exists so that the *unchanged* reference
// headers under /root/reference/src can be compiled with plain g++ into
// oracle/_ref/libekat_ref.so (Kokkos itself is an un-vendored, absent
// submodule: /root/reference/.gitmodules:1-3). Kokkos contributes no
// arithmetic to the hot path -- only indexing, loop distribution and barriers --
// so a one-thread team is exact. Nothing in the product
// (ekat_b200/, include/) includes this header.
//
// This is an original, minimal re-statement of public Kokkos API *signatures*;
// it contains no Kokkos source code.
#ifndef EKAT_B200_ORACLE_KOKKOS_STANDIN_HPP
#define EKAT_B200_ORACLE_KOKKOS_STANDIN_HPP

#include <cassert>
#include <cmath>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <memory>
#include <string>
#include <type_traits>
#include <utility>

#define KOKKOS_INLINE_FUNCTION inline
#define KOKKOS_FORCEINLINE_FUNCTION inline __attribute__((always_inline))
#define KOKKOS_FUNCTION
#define KOKKOS_LAMBDA [=]
#define KOKKOS_COMPILER_GNU 1330

namespace Kokkos {

// ---------------------------------------------------------------- layouts
namespace standin_impl {
struct LayoutBase {
  size_t dimension[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  LayoutBase() = default;
  LayoutBase(size_t a, size_t b = 0, size_t c = 0, size_t d = 0) {
    dimension[0] = a; dimension[1] = b; dimension[2] = c; dimension[3] = d;
  }
};
}  // namespace standin_impl
struct LayoutRight : standin_impl::LayoutBase { using LayoutBase::LayoutBase; using array_layout = LayoutRight; };
struct LayoutLeft  : standin_impl::LayoutBase { using LayoutBase::LayoutBase; using array_layout = LayoutLeft; };
struct LayoutStride : standin_impl::LayoutBase {
  using array_layout = LayoutStride;
  size_t stride[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  LayoutStride() = default;
  LayoutStride(size_t n0, size_t s0) { dimension[0] = n0; stride[0] = s0; }
};

// ----------------------------------------------------- spaces and devices
struct HostSpace { using memory_space = HostSpace; };
struct Serial {
  using execution_space = Serial;
  using memory_space = HostSpace;
  using array_layout = LayoutRight;
  using device_type = Serial;
  int concurrency() const { return 1; }
};
using DefaultExecutionSpace = Serial;
using DefaultHostExecutionSpace = Serial;
template <class E, class M>
struct Device {
  using execution_space = E;
  using memory_space = M;
  using device_type = Device;
};

// ----------------------------------------------------------- memory traits
enum MemoryTraitsFlags : unsigned { Unmanaged = 1, RandomAccess = 2, Atomic = 4, Restrict = 8, Aligned = 16 };
template <unsigned F>
struct MemoryTraits {
  enum : bool {
    is_unmanaged = (F & Unmanaged) != 0,
    is_random_access = (F & RandomAccess) != 0,
    is_atomic = (F & Atomic) != 0,
    is_restrict = (F & Restrict) != 0,
    is_aligned = (F & Aligned) != 0
  };
  static constexpr unsigned value = F;
};
using MemoryUnmanaged = MemoryTraits<Unmanaged>;
using MemoryManaged = MemoryTraits<0>;

// --------------------------------------------------------------- View type
namespace standin_impl {
template <class T> struct DataInfo { using value_type = T; static constexpr int rank = 0; };
template <class T> struct DataInfo<T*> {
  using value_type = typename DataInfo<T>::value_type;
  static constexpr int rank = DataInfo<T>::rank + 1;
};
template <class T> struct is_layout : std::false_type {};
template <> struct is_layout<LayoutRight> : std::true_type {};
template <> struct is_layout<LayoutLeft> : std::true_type {};
template <> struct is_layout<LayoutStride> : std::true_type {};
template <class T> struct is_memtraits : std::false_type {};
template <unsigned F> struct is_memtraits<MemoryTraits<F>> : std::true_type {};

// Pick the first property satisfying Pred, else Default.
template <template <class> class Pred, class Default, class... Ps> struct pick { using type = Default; };
template <template <class> class Pred, class Default, class P, class... Ps>
struct pick<Pred, Default, P, Ps...> {
  using type = typename std::conditional<Pred<P>::value, P, typename pick<Pred, Default, Ps...>::type>::type;
};
}  // namespace standin_impl

struct ALL_t {};
inline ALL_t ALL() { return ALL_t(); }

template <class DataType, class... Props>
class View {
  using info = standin_impl::DataInfo<DataType>;

 public:
  struct traits {
    using data_type = DataType;
    using scalar_array_type = DataType;
    using value_type = typename info::value_type;
    using non_const_value_type = typename std::remove_const<value_type>::type;
    using array_layout = typename standin_impl::pick<standin_impl::is_layout, LayoutRight, Props...>::type;
    using memory_traits = typename standin_impl::pick<standin_impl::is_memtraits, MemoryTraits<0>, Props...>::type;
    using device_type = Device<Serial, HostSpace>;
    using execution_space = Serial;
    using memory_space = HostSpace;
  };
  using data_type = DataType;
  using value_type = typename traits::value_type;
  using non_const_value_type = typename traits::non_const_value_type;
  using const_value_type = typename std::add_const<value_type>::type;
  using array_layout = typename traits::array_layout;
  using device_type = typename traits::device_type;
  using execution_space = Serial;
  using memory_space = HostSpace;
  using pointer_type = value_type*;
  using reference_type = value_type&;
  using HostMirror = View;

  // Kokkos exposes rank both as a value and as a callable.
  static constexpr std::integral_constant<int, info::rank> rank = {};

  View() = default;

  // Allocating constructor (managed, zero-initialised like Kokkos).
  explicit View(const std::string&, size_t n0 = 0, size_t n1 = 0, size_t n2 = 0, size_t n3 = 0) {
    set_dims(n0, n1, n2, n3);
    size_t n = span();
    using NC = non_const_value_type;
    NC* p = n ? new NC[n]() : nullptr;
    m_owner = std::shared_ptr<NC>(p, std::default_delete<NC[]>());
    m_data = p;
  }
  // Wrapping constructors (unmanaged).
  View(pointer_type p, size_t n0 = 0, size_t n1 = 0, size_t n2 = 0, size_t n3 = 0) : m_data(p) {
    set_dims(n0, n1, n2, n3);
  }
  View(pointer_type p, const array_layout& l) : m_data(p) { set_layout(l); }

  // Converting copy (non-const -> const value, managed -> unmanaged, ...).
  template <class DT2, class... P2,
            class = typename std::enable_if<
                standin_impl::DataInfo<DT2>::rank == info::rank &&
                std::is_convertible<typename View<DT2, P2...>::pointer_type, pointer_type>::value>::type>
  View(const View<DT2, P2...>& o) : m_data(o.data()) {
    for (int i = 0; i < 4; ++i) { m_dim[i] = o.extent(i); m_stride[i] = o.stride(i); }
    m_owner = o.owner();
  }

  KOKKOS_FORCEINLINE_FUNCTION size_t extent(int i) const { return i < info::rank ? m_dim[i] : 1; }
  KOKKOS_FORCEINLINE_FUNCTION int extent_int(int i) const { return static_cast<int>(extent(i)); }
  KOKKOS_FORCEINLINE_FUNCTION size_t stride(int i) const { return m_stride[i]; }
  KOKKOS_FORCEINLINE_FUNCTION size_t size() const {
    size_t n = 1;
    for (int i = 0; i < info::rank; ++i) n *= m_dim[i];
    return n;
  }
  KOKKOS_FORCEINLINE_FUNCTION size_t span() const { return size(); }
  KOKKOS_FORCEINLINE_FUNCTION pointer_type data() const { return m_data; }
  std::shared_ptr<void> owner() const { return m_owner; }

  KOKKOS_FORCEINLINE_FUNCTION reference_type operator()(size_t i0) const { return m_data[i0 * m_stride[0]]; }
  KOKKOS_FORCEINLINE_FUNCTION reference_type operator()(size_t i0, size_t i1) const {
    return m_data[i0 * m_stride[0] + i1 * m_stride[1]];
  }
  KOKKOS_FORCEINLINE_FUNCTION reference_type operator()(size_t i0, size_t i1, size_t i2) const {
    return m_data[i0 * m_stride[0] + i1 * m_stride[1] + i2 * m_stride[2]];
  }
  KOKKOS_FORCEINLINE_FUNCTION reference_type operator()(size_t i0, size_t i1, size_t i2, size_t i3) const {
    return m_data[i0 * m_stride[0] + i1 * m_stride[1] + i2 * m_stride[2] + i3 * m_stride[3]];
  }
  KOKKOS_FORCEINLINE_FUNCTION reference_type operator[](size_t i0) const { return m_data[i0 * m_stride[0]]; }

  // Only what the reference perf-test idiom needs: impl_map().reference(...).
  struct Map {
    const View* v;
    template <class... I> reference_type reference(I... i) const { return (*v)(static_cast<size_t>(i)...); }
  };
  Map impl_map() const { return Map{this}; }

 private:
  void set_dims(size_t n0, size_t n1, size_t n2, size_t n3) {
    m_dim[0] = n0; m_dim[1] = n1; m_dim[2] = n2; m_dim[3] = n3;
    for (int i = info::rank; i < 4; ++i) m_dim[i] = 1;
    compute_strides();
  }
  void set_layout(const LayoutStride& l) {
    for (int i = 0; i < 4; ++i) { m_dim[i] = i < info::rank ? l.dimension[i] : 1; m_stride[i] = l.stride[i]; }
  }
  void set_layout(const standin_impl::LayoutBase& l) {
    set_dims(l.dimension[0], l.dimension[1], l.dimension[2], l.dimension[3]);
  }
  void compute_strides() {
    constexpr int r = info::rank;
    if (std::is_same<array_layout, LayoutLeft>::value) {
      size_t s = 1;
      for (int i = 0; i < r; ++i) { m_stride[i] = s; s *= m_dim[i]; }
    } else {
      size_t s = 1;
      for (int i = r - 1; i >= 0; --i) { m_stride[i] = s; s *= m_dim[i]; }
    }
  }

  pointer_type m_data = nullptr;
  size_t m_dim[4] = {1, 1, 1, 1};
  size_t m_stride[4] = {1, 1, 1, 1};
  std::shared_ptr<void> m_owner;
};

template <class V> struct is_view : std::false_type {};
template <class D, class... P> struct is_view<View<D, P...>> : std::true_type {};

// ---- subview: only the forms on the hot path ---------------------------
// subview(X, ALL(), j) : column j of a rank-2 view -> strided rank-1 view.
template <class DT, class... P>
KOKKOS_INLINE_FUNCTION View<typename View<DT, P...>::value_type*, LayoutStride, MemoryTraits<Unmanaged>>
subview(const View<DT, P...>& v, ALL_t, int j) {
  static_assert(View<DT, P...>::rank == 2, "stand-in subview(ALL,j) is rank-2 only");
  LayoutStride l(v.extent(0), v.stride(0));
  return View<typename View<DT, P...>::value_type*, LayoutStride, MemoryTraits<Unmanaged>>(&v(0, j), l);
}
// subview(X, i, ALL()) : row i of a rank-2 view.
template <class DT, class... P>
KOKKOS_INLINE_FUNCTION View<typename View<DT, P...>::value_type*, LayoutRight, MemoryTraits<Unmanaged>>
subview(const View<DT, P...>& v, int i, ALL_t) {
  static_assert(View<DT, P...>::rank == 2, "stand-in subview(i,ALL) is rank-2 only");
  return View<typename View<DT, P...>::value_type*, LayoutRight, MemoryTraits<Unmanaged>>(&v(i, 0), v.extent(1));
}

template <class V> V create_mirror_view(const V& v) { return v; }
template <class A, class B> void deep_copy(const A&, const B&) {}
inline void fence() {}
[[noreturn]] inline void abort(const char* msg) { std::fprintf(stderr, "Kokkos::abort: %s\n", msg); std::abort(); }

// ------------------------------------------------------------- team policy
namespace standin_impl {
struct TeamMember {
  int m_league_rank = 0, m_league_size = 1;
  int league_rank() const { return m_league_rank; }
  int league_size() const { return m_league_size; }
  int team_rank() const { return 0; }
  int team_size() const { return 1; }
  void team_barrier() const {}
};
}  // namespace standin_impl

template <class... Props>
class TeamPolicy {
 public:
  using member_type = standin_impl::TeamMember;
  using execution_space = Serial;
  TeamPolicy() = default;
  TeamPolicy(int league, int team, int vec = 1) : m_league(league), m_team(team), m_vec(vec) {}
  int league_size() const { return m_league; }
  int team_size() const { return 1; }
 private:
  int m_league = 0, m_team = 1, m_vec = 1;
};
template <class... Props>
class RangePolicy {
 public:
  RangePolicy(long b, long e) : m_b(b), m_e(e) {}
  long begin() const { return m_b; }
  long end() const { return m_e; }
 private:
  long m_b, m_e;
};

template <class T> struct reduction_identity;
namespace standin_impl {
struct PerTeamTag {};
struct RangeBoundary { int begin, end; };
}  // namespace standin_impl
inline standin_impl::PerTeamTag PerTeam(const standin_impl::TeamMember&) { return {}; }
inline standin_impl::PerTeamTag PerThread(const standin_impl::TeamMember&) { return {}; }
template <class F> KOKKOS_FORCEINLINE_FUNCTION void single(standin_impl::PerTeamTag, const F& f) { f(); }

inline standin_impl::RangeBoundary TeamVectorRange(const standin_impl::TeamMember&, int n) { return {0, n}; }
inline standin_impl::RangeBoundary TeamThreadRange(const standin_impl::TeamMember&, int n) { return {0, n}; }
inline standin_impl::RangeBoundary ThreadVectorRange(const standin_impl::TeamMember&, int n) { return {0, n}; }
inline standin_impl::RangeBoundary TeamVectorRange(const standin_impl::TeamMember&, int b, int e) { return {b, e}; }
inline standin_impl::RangeBoundary TeamThreadRange(const standin_impl::TeamMember&, int b, int e) { return {b, e}; }
inline standin_impl::RangeBoundary ThreadVectorRange(const standin_impl::TeamMember&, int b, int e) { return {b, e}; }

template <class F>
KOKKOS_FORCEINLINE_FUNCTION void parallel_for(const standin_impl::RangeBoundary& r, const F& f) {
  for (int i = r.begin; i < r.end; ++i) f(i);
}
// Team-level reduction of a 1-thread team: the thread's own accumulation, in index order, starting from the
// identity (what Kokkos::parallel_reduce(TeamThreadRange) gives a team of one).
template <class F, class T>
KOKKOS_FORCEINLINE_FUNCTION void parallel_reduce(const standin_impl::RangeBoundary& r, const F& f, T& result) {
  T acc = reduction_identity<T>::sum();
  for (int i = r.begin; i < r.end; ++i) f(i, acc);
  result = acc;
}
// League loop: one 1-thread team per league entry, distributed over OpenMP
// threads exactly as Kokkos::OpenMP distributes TeamPolicy(n, 1).
template <class... P, class F>
void parallel_for(const TeamPolicy<P...>& p, const F& f) {
  const int n = p.league_size();
#if defined(_OPENMP)
#pragma omp parallel for schedule(static)
#endif
  for (int i = 0; i < n; ++i) {
    standin_impl::TeamMember m;
    m.m_league_rank = i;
    m.m_league_size = n;
    f(m);
  }
}
template <class... P, class F>
void parallel_for(const std::string&, const TeamPolicy<P...>& p, const F& f) { parallel_for(p, f); }
template <class... P, class F>
void parallel_for(const RangePolicy<P...>& p, const F& f) {
  for (long i = p.begin(); i < p.end(); ++i) f(static_cast<int>(i));
}

// ------------------------------------------------------------------- misc
template <class A, class B> struct pair { A first; B second; };
template <class A, class B> pair<A, B> make_pair(A a, B b) { return pair<A, B>{a, b}; }
template <class T, size_t N> struct Array {
  T m_data[N ? N : 1];
  T& operator[](size_t i) { return m_data[i]; }
  const T& operator[](size_t i) const { return m_data[i]; }
  static constexpr size_t size() { return N; }
  T* data() { return m_data; }
  const T* data() const { return m_data; }
};
template <class T> struct reduction_identity {
  static T sum() { return T(0); }
  static T prod() { return T(1); }
  static T max() { return std::numeric_limits<T>::lowest(); }
  static T min() { return std::numeric_limits<T>::max(); }
};

using std::abs; using std::ceil; using std::cbrt; using std::exp; using std::expm1; using std::floor;
using std::isfinite; using std::isnan; using std::log; using std::log10; using std::log1p; using std::pow;
using std::round; using std::sqrt; using std::tanh; using std::tgamma; using std::trunc; using std::erf;
using std::fmax; using std::fmin; using std::copysign; using std::isinf; using std::sin; using std::cos;
using std::max; using std::min;
template <class T> inline T sign(T x) { return (T(0) < x) - (x < T(0)); }

namespace Experimental {
template <class T> inline constexpr T quiet_NaN_v = std::numeric_limits<T>::quiet_NaN();
template <class T> inline constexpr T finite_max_v = std::numeric_limits<T>::max();
template <class T> inline constexpr T finite_min_v = std::numeric_limits<T>::lowest();
template <class T> inline constexpr T epsilon_v = std::numeric_limits<T>::epsilon();
}  // namespace Experimental

}  // namespace Kokkos

#endif  // EKAT_B200_ORACLE_KOKKOS_STANDIN_HPP
