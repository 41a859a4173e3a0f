// Stand-in: the math functions live in Kokkos_Core.hpp here. TEST INFRASTRUCTURE ONLY.
#ifndef EKAT_B200_ORACLE_KOKKOS_MATHFN_STANDIN_HPP
#define EKAT_B200_ORACLE_KOKKOS_MATHFN_STANDIN_HPP
#include <Kokkos_Core.hpp>
#endif
