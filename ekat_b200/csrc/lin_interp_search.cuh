// lin_interp_search.cuh -- the bracket search of ekat::LinInterp::setup (reference src/algorithm/ekat_lin_interp_impl.hpp:160-186
// with src/kokkos/ekat_upper_bound.hpp:28-50) on a column staged in shared memory; shared by the reference-layout and
// the level-major interpolation kernels.
#pragma once
#include "common.cuh"

namespace ekb {

// upper_bound over sorted x1[0..km1): the number of entries <= v, found with the predicate `v >= *it` of the
// reference's EKAT_ENABLE_GPU variant (ekat_upper_bound.hpp:28-50) -- the variant this library replaces. Ties
// resolve identically in both reference variants; a NaN target fails every `>=` and maps to index 0 here and in
// the GPU variant, whereas the host variant (std::upper_bound, `!(v < *it)`, :51-53) would map it to km1-1
// (tests/test_gpu_lin_interp.py::test_nan_targets_follow_the_gpu_upper_bound_variant). Then the reference's
// "if (idx > 0) --idx" (:179-181). Branch-free: one probe picks the power-of-two window [0,P) or [km1-P,km1)
// (P = bit_floor(km1)) that must contain the boundary, then log2(P) probes with halving steps and a last one --
// each a shared-memory load, a compare and a predicated add. LOG2P > 0: unrolled for that window; 0: loop.
__device__ __forceinline__ double lds_at(unsigned addr, double) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ float lds_at(unsigned addr, float) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
  return v;
}
// The running position is a 32-bit shared-memory byte address and a probe is spelled in PTX as exactly: load
// [p + const], compare, predicated add (left to itself the compiler emits an add, a compare and a predicated move,
// and five more instructions per probe for generic 64-bit pointers). Returns the BYTE offset of the bracket k1
// within the column (k1 * STRIDE): every later gather is [column base + offset] with no index arithmetic.
template <int OFF, int STEP>
__device__ __forceinline__ void probe(unsigned& p, double v) {
  asm("{\n.reg .pred q;\n.reg .f64 t;\nld.shared.f64 t, [%0+%2];\nsetp.ge.f64 q, %1, t;\n@q add.u32 %0, %0, %3;\n}\n"
      : "+r"(p)
      : "d"(v), "n"(OFF), "n"(STEP));
}
template <int OFF, int STEP>
__device__ __forceinline__ void probe(unsigned& p, float v) {
  asm("{\n.reg .pred q;\n.reg .f32 t;\nld.shared.f32 t, [%0+%2];\nsetp.ge.f32 q, %1, t;\n@q add.u32 %0, %0, %3;\n}\n"
      : "+r"(p)
      : "f"(v), "n"(OFF), "n"(STEP));
}
template <typename R, int BIT, int B>
struct Probes {  // probes with steps 2^BIT, 2^(BIT-1), ..., 1 elements: the window [p, p + 2^(BIT+1)) shrinks to one slot
  static __device__ __forceinline__ void run(unsigned& p, R v) {
    probe<((1 << BIT) - 1) * B, B << BIT>(p, v);
    Probes<R, BIT - 1, B>::run(p, v);
  }
};
template <typename R, int B>
struct Probes<R, -1, B> {
  static __device__ __forceinline__ void run(unsigned&, R) {}
};
// STRIDE = bytes between consecutive source levels of the column: sizeof(R) for the reference's [column][level]
// rows, 32 * sizeof(R) for a level-major tile [level][32 columns] (each lane then stays in its own bank).
template <typename R, int LOG2P, int STRIDE>
__device__ __forceinline__ unsigned bracket_off(unsigned x1, int km1, int P, R v) {
  constexpr unsigned B = STRIDE;
  unsigned p = x1;
  if (km1 > P) {
    if (v >= lds_at(x1 + (unsigned)(km1 - P - 1) * B, R())) p += (unsigned)(km1 - P) * B;
  }
  if (LOG2P > 0) {
    Probes<R, LOG2P - 1, STRIDE>::run(p, v);  // LOG2P probes: steps P/2 ... 1
  } else {
    for (int step = P >> 1; step >= 1; step >>= 1)
      if (v >= lds_at(p + (unsigned)(step - 1) * B, R())) p += (unsigned)step * B;
  }
  probe<0, B>(p, v);  // the last slot itself
  // pos = (p - x1) / B entries are <= v; the reference's "if (idx > 0) --idx" (ekat_lin_interp_impl.hpp:179-181)
  return max(p - x1, B) - B;
}

}  // namespace ekb
