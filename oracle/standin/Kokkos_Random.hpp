// Empty stand-in: ekat_team_policy_utils.hpp includes <Kokkos_Random.hpp> but the
// LinInterp/tridiag hot path uses none of it. TEST INFRASTRUCTURE ONLY.
#ifndef EKAT_B200_ORACLE_KOKKOS_RANDOM_STANDIN_HPP
#define EKAT_B200_ORACLE_KOKKOS_RANDOM_STANDIN_HPP
#include <Kokkos_Core.hpp>
#endif
