// oracle/ref_driver.cpp -- C-ABI driver around the UNCHANGED reference headers.
//
// TEST INFRASTRUCTURE ONLY (checker + timed CPU baseline). Built by
// oracle/Makefile into oracle/_ref/libekat_ref.so from the reference sources
// where they lie (-I/root/reference/src/{algorithm,pack,kokkos,core} and
// -I/root/reference/tests/algorithm) against oracle/standin/Kokkos_Core.hpp.
// No reference source is copied here: this file only *calls*
//   ekat::tridiag::{thomas,cr,bfb}       (src/algorithm/ekat_tridiag.hpp:323-678)
//   ekat::LinInterp::{setup,lin_interp}  (src/algorithm/ekat_lin_interp.hpp:69-126)
//   ekat::adj_diff<Forward>              (src/pack/ekat_pack_kokkos.hpp:100-164)
//   ekat::shift_left / shift_right       (src/pack/ekat_pack.hpp:521-555)
//   ekat::index_and_shift<Shift>         (src/pack/ekat_pack_kokkos.hpp:71-98)
//   ekat::transpose<c2f|f2c> (rank 2, 3) (src/kokkos/ekat_math_utils.hpp:90-147)
//   ekat::impl::view_reduction<Serialize> (src/algorithm/ekat_reduction_utils.hpp:67-144)
//   ekat::test::fill_tridiag_matrix / fill_data_matrix / matvec
//                                        (tests/algorithm/tridiag_tests.hpp:11-88)
// through the same batch idiom as the reference's own drivers
// (tests/algorithm/tridiag_tests_performance.cpp:204-303,
//  tests/algorithm/lin_interp_test.cpp:141-154): one 1-thread team per column,
// league loop distributed with OpenMP (what Kokkos::OpenMP does for
// TeamPolicy(ncol,1), src/kokkos/ekat_team_policy_utils.hpp:22-30).
//
// Data layout is the reference's: A[ncol][3][nrow][nA], X[ncol][nrow][nrhs]
// (tests/algorithm/tridiag_tests_performance.cpp:181-185); LinInterp views are
// [ncol][npack(km)*N] scalars (src/algorithm/ekat_lin_interp_impl.hpp:12-19).

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>
#if defined(_OPENMP)
#include <omp.h>
#endif

#include <Kokkos_Core.hpp>

// The LinInterp index map is a private member; the checker wants to compare it.
#define private public
#include "ekat_lin_interp.hpp"
#undef private
#include "ekat_tridiag.hpp"
#include "ekat_reduction_utils.hpp"
#include "ekat_math_utils.hpp"
#include "ekat_pack_kokkos.hpp"
#include "tridiag_tests.hpp"

namespace {

using Kokkos::LayoutRight;
using Kokkos::MemoryUnmanaged;
using Kokkos::View;

template <class S> using V1 = View<S*, LayoutRight, MemoryUnmanaged>;
template <class S> using V2 = View<S**, LayoutRight, MemoryUnmanaged>;
using Policy = Kokkos::TeamPolicy<Kokkos::DefaultExecutionSpace>;
using Member = Policy::member_type;

enum Solver { THOMAS_SERIAL = 0, THOMAS_TEAM = 1, CR = 2, BFB = 3 };

template <class S>
int tridiag(int solver, S* A, S* X, int64_t ncol, int nrow, int nA, int nrhs) {
  if (ncol < 0 || nrow < 1 || nrhs < 1 || (nA != 1 && nA != nrhs)) return -1;
  if (ncol > 0x7fffffff) return -2;
  const size_t astride = size_t(3) * nrow * nA, xstride = size_t(nrow) * nrhs;
  Policy policy(static_cast<int>(ncol), 1, 1);
  int err = 0;
  const auto f = [=, &err](const Member& team) {
    const size_t ip = team.league_rank();
    S* a = A + ip * astride;
    S* x = X + ip * xstride;
    const size_t dsz = size_t(nrow) * nA;
    if (nA == 1) {
      V1<S> dl(a, nrow), d(a + dsz, nrow), du(a + 2 * dsz, nrow);
      V2<S> x2(x, nrow, nrhs);
      switch (solver) {
        case THOMAS_SERIAL:
          if (nrhs == 1) {
            V1<S> x1(x, nrow);
            Kokkos::single(Kokkos::PerTeam(team), [&]() { ekat::tridiag::thomas(dl, d, du, x1); });
          } else {
            Kokkos::single(Kokkos::PerTeam(team), [&]() { ekat::tridiag::thomas(dl, d, du, x2); });
          }
          break;
        case THOMAS_TEAM: ekat::tridiag::thomas(team, dl, d, du, x2); break;
        case CR:
          if (nrhs == 1) {
            V1<S> x1(x, nrow);
            ekat::tridiag::cr(team, dl, d, du, x1);
          } else {
            ekat::tridiag::cr(team, dl, d, du, x2);
          }
          break;
        case BFB: ekat::tridiag::bfb(team, dl, d, du, x2); break;
        default: err = -3;
      }
    } else {
      V2<S> dl(a, nrow, nA), d(a + dsz, nrow, nA), du(a + 2 * dsz, nrow, nA);
      V2<S> x2(x, nrow, nrhs);
      switch (solver) {
        case THOMAS_SERIAL:
          Kokkos::single(Kokkos::PerTeam(team), [&]() { ekat::tridiag::thomas(dl, d, du, x2); });
          break;
        case THOMAS_TEAM: err = -4; break;  // rank-2 A unsupported by team thomas (ekat_tridiag.hpp:323-327)
        case CR: ekat::tridiag::cr(team, dl, d, du, x2); break;
        case BFB: ekat::tridiag::bfb(team, dl, d, du, x2); break;
        default: err = -3;
      }
    }
  };
  Kokkos::parallel_for(policy, f);
  return err;
}

// Reference fills, per column c: fill_tridiag_matrix(dl,d,du,nA,seed=(col0+c)%seed_mod),
// fill_data_matrix(X, seed=nrhs) -- tests/algorithm/tridiag_tests_performance.cpp:190-196.
template <class S>
void fill(S* A, S* X, int64_t ncol, int64_t col0, int nrow, int nA, int nrhs, int seed_mod) {
  const size_t astride = size_t(3) * nrow * nA, xstride = size_t(nrow) * nrhs, dsz = size_t(nrow) * nA;
#if defined(_OPENMP)
#pragma omp parallel for schedule(static)
#endif
  for (int64_t c = 0; c < ncol; ++c) {
    S* a = A + c * astride;
    V2<S> dl(a, nrow, nA), d(a + dsz, nrow, nA), du(a + 2 * dsz, nrow, nA);
    const int seed = static_cast<int>(seed_mod > 0 ? (col0 + c) % seed_mod : (col0 + c));
    ekat::test::fill_tridiag_matrix(dl, d, du, nA, seed);
    if (X) {
      V2<S> x(X + c * xstride, nrow, nrhs);
      ekat::test::fill_data_matrix(x, nrhs);
    }
  }
}

// max_c [ max|B - A X| / max|B| ] using the reference's matvec (tridiag_tests.hpp:42-88)
// and its norm (tridiag_tests.hpp:89-106).
template <class S>
double residual(const S* A, const S* B, const S* X, int64_t ncol, int nrow, int nA, int nrhs) {
  const size_t astride = size_t(3) * nrow * nA, xstride = size_t(nrow) * nrhs, dsz = size_t(nrow) * nA;
  double worst = 0;
#if defined(_OPENMP)
#pragma omp parallel for schedule(static) reduction(max : worst)
#endif
  for (int64_t c = 0; c < ncol; ++c) {
    S* a = const_cast<S*>(A) + c * astride;
    V2<S> dl(a, nrow, nA), d(a + dsz, nrow, nA), du(a + 2 * dsz, nrow, nA);
    V2<S> x(const_cast<S*>(X) + c * xstride, nrow, nrhs);
    std::vector<S> ybuf(xstride);
    V2<S> y(ybuf.data(), nrow, nrhs);
    ekat::test::matvec(dl, d, du, x, y, nA, nrhs);
    const S* b = B + c * xstride;
    double num = 0, den = 0;
    for (size_t k = 0; k < xstride; ++k) {
      if (!std::isfinite(double(ybuf[k]))) { num = INFINITY; den = 1; break; }
      num = std::max(num, std::abs(double(b[k]) - double(ybuf[k])));
      den = std::max(den, std::abs(double(b[k])));
    }
    const double r = den > 0 ? num / den : num;
    worst = std::max(worst, r);
  }
  return worst;
}

template <class S, int N>
int lin_interp_n(const S* x1, const S* x2, const S* y1, S* y2, int* indx_map, int64_t ncol, int km1, int km2,
                 int nfields) {
  using LI = ekat::LinInterp<S, N>;
  using Pack = typename LI::Pack;
  LI li(static_cast<int>(ncol), km1, km2);
  const int p1 = li.km1_pack(), p2 = li.km2_pack();
  const size_t pitch1 = size_t(p1) * N, pitch2 = size_t(p2) * N;
  using CV = View<const Pack**, LayoutRight, MemoryUnmanaged>;
  using MV = View<Pack**, LayoutRight, MemoryUnmanaged>;
  CV x1v(reinterpret_cast<const Pack*>(x1), ncol, p1), x2v(reinterpret_cast<const Pack*>(x2), ncol, p2);
  const auto f = [=](const typename LI::MemberType& team) {
    const int i = team.league_rank();
    const auto x1c = Kokkos::subview(x1v, i, Kokkos::ALL());
    const auto x2c = Kokkos::subview(x2v, i, Kokkos::ALL());
    li.setup(team, x1c, x2c);
    team.team_barrier();
    for (int fld = 0; fld < nfields; ++fld) {
      CV y1v(reinterpret_cast<const Pack*>(y1 + size_t(fld) * ncol * pitch1), ncol, p1);
      MV y2v(reinterpret_cast<Pack*>(y2 + size_t(fld) * ncol * pitch2), ncol, p2);
      li.lin_interp(team, x1c, x2c, Kokkos::subview(y1v, i, Kokkos::ALL()), Kokkos::subview(y2v, i, Kokkos::ALL()));
    }
  };
  Kokkos::parallel_for(li.policy(), f);
  if (indx_map) {
    const auto ims = ekat::scalarize(li.m_indx_map);
    for (int64_t c = 0; c < ncol; ++c)
      for (int k = 0; k < km2; ++k) indx_map[c * km2 + k] = ims(c, k);
  }
  return 0;
}

template <class S>
int lin_interp(int packn, const S* x1, const S* x2, const S* y1, S* y2, int* indx_map, int64_t ncol, int km1,
               int km2, int nfields) {
  if (ncol < 0 || km1 < 0 || km2 < 0 || nfields < 0 || ncol > 0x7fffffff) return -1;
  switch (packn) {
    case 1: return lin_interp_n<S, 1>(x1, x2, y1, y2, indx_map, ncol, km1, km2, nfields);
    case 2: return lin_interp_n<S, 2>(x1, x2, y1, y2, indx_map, ncol, km1, km2, nfields);
    case 4: return lin_interp_n<S, 4>(x1, x2, y1, y2, indx_map, ncol, km1, km2, nfields);
    case 8: return lin_interp_n<S, 8>(x1, x2, y1, y2, indx_map, ncol, km1, km2, nfields);
    case 16: return lin_interp_n<S, 16>(x1, x2, y1, y2, indx_map, ncol, km1, km2, nfields);
    default: return -2;
  }
}

}  // namespace

namespace {
// view_reduction for every column: in [ncol][pitch] scalars = [ncol][pitch/N] packs (pitch % N == 0), out [ncol];
// called as the reference's test calls it (tests/algorithm/reduction_tests.cpp:117-126), the view as provider.
template <class S, int N, bool Serialize>
int view_reduction_n(const S* in, S* out, int64_t ncol, int begin, int end, int64_t pitch) {
  using P = ekat::Pack<S, N>;
  if (pitch % N) return -3;
  Kokkos::parallel_for(Policy((int)ncol, 1), [&](const Member& team) {
    const int c = team.league_rank();
    V1<const P> row(reinterpret_cast<const P*>(in + (size_t)c * pitch), pitch / N);
    out[c] = ekat::impl::view_reduction<Serialize>(team, begin, end, row);
  });
  return 0;
}
// ekat::adj_diff<Forward> at every index of every column (src/pack/ekat_pack_kokkos.hpp:100-164), called as the
// reference's test calls it on a view (tests/pack/pack_kokkos_utils.cpp:1061-1077): in, out [ncol][pitch] scalars.
template <class S, int N, bool Forward>
int adj_diff_n(const S* in, S* out, int64_t ncol, int64_t pitch) {
  using P = ekat::Pack<S, N>;
  if (pitch % N) return -3;
  const int np = (int)(pitch / N);
  Kokkos::parallel_for(Policy((int)ncol, 1), [&](const Member& team) {
    const int c = team.league_rank();
    V1<const P> row(reinterpret_cast<const P*>(in + (size_t)c * pitch), np);
    P* o = reinterpret_cast<P*>(out + (size_t)c * pitch);
    for (int i = 0; i < np; ++i) o[i] = ekat::adj_diff<Forward>(row, i, np);
  });
  return 0;
}
template <class S, bool Forward>
int adj_diff_s(const S* in, S* out, int64_t ncol, int64_t pitch) {  // scalar views take the non-pack overload (:139-157)
  Kokkos::parallel_for(Policy((int)ncol, 1), [&](const Member& team) {
    const int c = team.league_rank();
    V1<const S> row(in + (size_t)c * pitch, pitch);
    for (int i = 0; i < (int)pitch; ++i) out[(size_t)c * pitch + i] = ekat::adj_diff<Forward>(row, i, (int)pitch);
  });
  return 0;
}
template <class S>
int adj_diff(int packn, int forward, const S* in, S* out, int64_t ncol, int64_t pitch) {
#define EKAT_REF_AD(N_) \
  case N_: return forward ? adj_diff_n<S, N_, true>(in, out, ncol, pitch) : adj_diff_n<S, N_, false>(in, out, ncol, pitch);
  switch (packn) {
    case 0: return forward ? adj_diff_s<S, true>(in, out, ncol, pitch) : adj_diff_s<S, false>(in, out, ncol, pitch);
    EKAT_REF_AD(1) EKAT_REF_AD(2) EKAT_REF_AD(4) EKAT_REF_AD(8) EKAT_REF_AD(16)
    default: return -4;
  }
#undef EKAT_REF_AD
}

// ekat::shift_right / shift_left pack after pack along a scalarized column (src/pack/ekat_pack.hpp:521-555), the way
// adj_diff chains them (ekat_pack_kokkos.hpp:118-126): the neighbouring pack supplies the incoming lane, the scalar
// `fill` at the end of the column. dir = +1: shift_right, -1: shift_left. in, out: [ncol][pitch], pitch % N == 0.
template <class S, int N>
int shift_n(int dir, S fill, const S* in, S* out, int64_t ncol, int64_t pitch) {
  using P = ekat::Pack<S, N>;
  if (pitch % N) return -3;
  const int np = (int)(pitch / N);
  Kokkos::parallel_for(Policy((int)ncol, 1), [&](const Member& team) {
    const int c = team.league_rank();
    const P* v = reinterpret_cast<const P*>(in + (size_t)c * pitch);
    P* o = reinterpret_cast<P*>(out + (size_t)c * pitch);
    for (int k = 0; k < np; ++k) {
      if (dir > 0) o[k] = k > 0 ? ekat::shift_right(v[k - 1], v[k]) : ekat::shift_right(fill, v[k]);
      else o[k] = k < np - 1 ? ekat::shift_left(v[k + 1], v[k]) : ekat::shift_left(fill, v[k]);
    }
  });
  return 0;
}
template <class S>
int shift(int packn, int dir, S fill, const S* in, S* out, int64_t ncol, int64_t pitch) {
  switch (packn) {
    case 1: return shift_n<S, 1>(dir, fill, in, out, ncol, pitch);
    case 2: return shift_n<S, 2>(dir, fill, in, out, ncol, pitch);
    case 4: return shift_n<S, 4>(dir, fill, in, out, ncol, pitch);
    case 8: return shift_n<S, 8>(dir, fill, in, out, ncol, pitch);
    case 16: return shift_n<S, 16>(dir, fill, in, out, ncol, pitch);
    default: return -4;
  }
}

// ekat::index_and_shift<Shift>(a, idx_pack, index, index_shift) (src/pack/ekat_pack_kokkos.hpp:71-98) for every index
// pack of every column: a [ncol][n] scalars, idx [ncol][m] ints (m % N == 0), out / out_shift [ncol][m]. This library
// is built without NDEBUG, so out-of-range lanes behave as the header's debug branch says: `index` keeps what the
// caller put there, `index_shift` is ekat::invalid.
template <class S, int N, int SH>
int index_and_shift_n(const S* a, const int* idx, S* out, S* out_shift, int64_t ncol, int n, int m) {
  using P = ekat::Pack<S, N>;
  using IP = ekat::Pack<int, N>;
  if (m % N) return -3;
  Kokkos::parallel_for(Policy((int)ncol, 1), [&](const Member& team) {
    const int c = team.league_rank();
    V1<const S> row(a + (size_t)c * n, n);
    for (int j = 0; j < m / N; ++j) {
      IP ip;
      P p0, p1;
      for (int l = 0; l < N; ++l) {
        ip[l] = idx[(size_t)c * m + j * N + l];
        p0[l] = out[(size_t)c * m + j * N + l];
        p1[l] = out_shift[(size_t)c * m + j * N + l];
      }
      ekat::index_and_shift<SH>(row, ip, p0, p1);
      for (int l = 0; l < N; ++l) {
        out[(size_t)c * m + j * N + l] = p0[l];
        out_shift[(size_t)c * m + j * N + l] = p1[l];
      }
    }
  });
  return 0;
}
template <class S>
int index_and_shift(int packn, int sh, const S* a, const int* idx, S* out, S* out_shift, int64_t ncol, int n, int m) {
#define EKAT_REF_IS(N_)                                                                                      \
  case N_:                                                                                                   \
    return sh == 1 ? index_and_shift_n<S, N_, 1>(a, idx, out, out_shift, ncol, n, m)                          \
                   : sh == -1 ? index_and_shift_n<S, N_, -1>(a, idx, out, out_shift, ncol, n, m)              \
                              : sh == 2 ? index_and_shift_n<S, N_, 2>(a, idx, out, out_shift, ncol, n, m) : -5;
  switch (packn) {
    EKAT_REF_IS(1) EKAT_REF_IS(4) EKAT_REF_IS(16)
    default: return -4;
  }
#undef EKAT_REF_IS
}

template <class S>
int view_reduction(int packn, int serialize, const S* in, S* out, int64_t ncol, int begin, int end, int64_t pitch) {
#define EKAT_REF_VR(N_) \
  case N_: return serialize ? view_reduction_n<S, N_, true>(in, out, ncol, begin, end, pitch) \
                            : view_reduction_n<S, N_, false>(in, out, ncol, begin, end, pitch);
  switch (packn) {
    EKAT_REF_VR(1) EKAT_REF_VR(2) EKAT_REF_VR(4) EKAT_REF_VR(8) EKAT_REF_VR(16)
    default: return -4;
  }
#undef EKAT_REF_VR
}

}  // namespace

extern "C" {
int ekat_ref_shift_f64(int packn, int dir, double fill, const double* in, double* out, int64_t ncol, int64_t pitch) {
  return shift<double>(packn, dir, fill, in, out, ncol, pitch);
}
int ekat_ref_shift_f32(int packn, int dir, float fill, const float* in, float* out, int64_t ncol, int64_t pitch) {
  return shift<float>(packn, dir, fill, in, out, ncol, pitch);
}
int ekat_ref_index_and_shift_f64(int packn, int sh, const double* a, const int* idx, double* out, double* out_shift,
                                 int64_t ncol, int n, int m) {
  return index_and_shift<double>(packn, sh, a, idx, out, out_shift, ncol, n, m);
}
int ekat_ref_index_and_shift_f32(int packn, int sh, const float* a, const int* idx, float* out, float* out_shift,
                                 int64_t ncol, int n, int m) {
  return index_and_shift<float>(packn, sh, a, idx, out, out_shift, ncol, n, m);
}
// ekat::transpose (src/kokkos/ekat_math_utils.hpp:90-147): c2f != 0 -> TransposeDirection::c2f; nj == 0 -> the rank-2 form
int ekat_ref_transpose_f64(const double* sv, double* dv, int ni, int nk, int nj, int c2f) {
  const auto dir = c2f ? ekat::TransposeDirection::c2f : ekat::TransposeDirection::f2c;
  if (nj == 0) ekat::transpose(sv, dv, ni, nk, dir);
  else ekat::transpose(sv, dv, ni, nk, nj, dir);
  return 0;
}
int ekat_ref_transpose_f32(const float* sv, float* dv, int ni, int nk, int nj, int c2f) {
  const auto dir = c2f ? ekat::TransposeDirection::c2f : ekat::TransposeDirection::f2c;
  if (nj == 0) ekat::transpose(sv, dv, ni, nk, dir);
  else ekat::transpose(sv, dv, ni, nk, nj, dir);
  return 0;
}
int ekat_ref_adj_diff_f64(int packn, int forward, const double* in, double* out, int64_t ncol, int64_t pitch) {
  return adj_diff<double>(packn, forward, in, out, ncol, pitch);
}
int ekat_ref_adj_diff_f32(int packn, int forward, const float* in, float* out, int64_t ncol, int64_t pitch) {
  return adj_diff<float>(packn, forward, in, out, ncol, pitch);
}
int ekat_ref_view_reduction_f64(int packn, int serialize, const double* in, double* out, int64_t ncol, int begin, int end,
                                int64_t pitch) {
  return view_reduction<double>(packn, serialize, in, out, ncol, begin, end, pitch);
}
int ekat_ref_view_reduction_f32(int packn, int serialize, const float* in, float* out, int64_t ncol, int begin, int end,
                                int64_t pitch) {
  return view_reduction<float>(packn, serialize, in, out, ncol, begin, end, pitch);
}

int ekat_ref_tridiag_f64(int solver, double* A, double* X, int64_t ncol, int nrow, int nA, int nrhs) {
  return tridiag<double>(solver, A, X, ncol, nrow, nA, nrhs);
}
int ekat_ref_tridiag_f32(int solver, float* A, float* X, int64_t ncol, int nrow, int nA, int nrhs) {
  return tridiag<float>(solver, A, X, ncol, nrow, nA, nrhs);
}
void ekat_ref_fill_tridiag_f64(double* A, double* X, int64_t ncol, int64_t col0, int nrow, int nA, int nrhs,
                               int seed_mod) {
  fill<double>(A, X, ncol, col0, nrow, nA, nrhs, seed_mod);
}
void ekat_ref_fill_tridiag_f32(float* A, float* X, int64_t ncol, int64_t col0, int nrow, int nA, int nrhs,
                               int seed_mod) {
  fill<float>(A, X, ncol, col0, nrow, nA, nrhs, seed_mod);
}
double ekat_ref_residual_f64(const double* A, const double* B, const double* X, int64_t ncol, int nrow, int nA,
                             int nrhs) {
  return residual<double>(A, B, X, ncol, nrow, nA, nrhs);
}
double ekat_ref_residual_f32(const float* A, const float* B, const float* X, int64_t ncol, int nrow, int nA,
                             int nrhs) {
  return residual<float>(A, B, X, ncol, nrow, nA, nrhs);
}
// x1,y1: [ncol][npack(km1)*packn]; x2,y2: [ncol][npack(km2)*packn]; nfields y1/y2 slabs back to back.
// indx_map (nullable): [ncol][km2] receives LinInterp::m_indx_map.
int ekat_ref_lin_interp_f64(int packn, const double* x1, const double* x2, const double* y1, double* y2,
                            int* indx_map, int64_t ncol, int km1, int km2, int nfields) {
  return lin_interp<double>(packn, x1, x2, y1, y2, indx_map, ncol, km1, km2, nfields);
}
int ekat_ref_lin_interp_f32(int packn, const float* x1, const float* x2, const float* y1, float* y2,
                            int* indx_map, int64_t ncol, int km1, int km2, int nfields) {
  return lin_interp<float>(packn, x1, x2, y1, y2, indx_map, ncol, km1, km2, nfields);
}
int ekat_ref_max_threads(void) {
#if defined(_OPENMP)
  return omp_get_max_threads();
#else
  return 1;
#endif
}
void ekat_ref_set_threads(int n) {
#if defined(_OPENMP)
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}
const char* ekat_ref_build_info(void) {
  return "reference headers (unchanged) + oracle/standin Kokkos; g++ " __VERSION__
#if defined(EKAT_REF_CONTRACT_OFF)
         "; -ffp-contract=off"
#else
         "; fp-contract default"
#endif
      ;
}

}  // extern "C"
