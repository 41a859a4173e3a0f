// Stand-in for the CMake-generated tests/ekat_test_config.h(.in) of the reference.
// TEST INFRASTRUCTURE ONLY.
#ifndef EKAT_TEST_CONFIG_H
#define EKAT_TEST_CONFIG_H
using Real = double;
#define EKAT_TEST_PACK_SIZE 16
#endif
