"""Generate tests/golden/column_ops.npz from oracle/_ref (the UNCHANGED reference headers src/pack/ekat_pack.hpp,
src/pack/ekat_pack_kokkos.hpp, src/kokkos/ekat_math_utils.hpp compiled in place). Run in the build container only:

    python -m oracle.gen_golden_column_ops

TEST INFRASTRUCTURE. Cases: ekat::shift_right / shift_left chained over columns of Pack<Real,1|4|16> (the reference's
own test data i^2 in row 0, tests/pack/pack_kokkos_utils.cpp:1040-1052, seeded columns below), ekat::index_and_shift
<+1|-1|+2> with in-range and out-of-range indices, ekat::transpose<c2f|f2c> of rank-2 and rank-3 arrays."""
import os

import numpy as np

import oracle

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "column_ops.npz")


def main():
    r = oracle.ref()
    assert r is not None, "oracle/_ref/libekat_ref.so missing: run `make -C oracle` where /root/reference exists"
    out = {}
    for dt in ("f64", "f32"):
        dtype = np.float64 if dt == "f64" else np.float32
        for pitch, N in [(16, 1), (16, 4), (128, 16), (80, 16), (72, 1), (8, 4)]:
            rng = np.random.default_rng(500 + pitch + N)
            x = (rng.standard_normal((20, pitch)) * 50).astype(dtype)
            x[0] = (np.arange(pitch, dtype=np.float64) ** 2).astype(dtype)
            key = f"shift_{dt}_{pitch}_{N}"
            out[key + "_x"] = x
            out[key + "_right"] = r.shift(x, +1, 0.0, N)
            out[key + "_left"] = r.shift(x, -1, 0.0, N)
            out[key + "_right_fill"] = r.shift(x, +1, -3.5, N)
            out[key + "_left_fill"] = r.shift(x, -1, 7.25, N)
        for n, m, N, sh in [(72, 128, 16, 1), (128, 128, 4, 1), (33, 16, 1, -1), (64, 32, 16, 2), (10, 16, 4, 1)]:
            rng = np.random.default_rng(900 + n + m + N)
            a = (rng.standard_normal((12, n)) * 10).astype(dtype)
            idx = rng.integers(0, n, (12, m)).astype(np.int32)
            idx[1, :4] = [n - 1, 0, n - 2, 1]       # the ends: shifted index leaves the column for some of these
            # an index that is itself out of range is undefined in the reference (only the SHIFTED index is checked,
            # ekat_pack_kokkos.hpp:91), so the fixture holds none; these leave the column through the shift
            idx[2, 0] = n - 1 if sh > 0 else 0
            idx[2, 1] = n - sh if sh > 1 else idx[2, 1]
            key = f"ias_{dt}_{n}_{m}_{N}_{sh}"
            o0, o1 = r.index_and_shift(a, idx, sh, packn=N)
            out[key + "_a"], out[key + "_idx"], out[key + "_out"], out[key + "_shift"] = a, idx, o0, o1
        for shape in [(7, 5), (33, 72), (5, 4, 3), (12, 72, 10), (3, 128, 1)]:
            rng = np.random.default_rng(1300 + sum(shape))
            sv = rng.standard_normal(int(np.prod(shape))).astype(dtype)
            key = f"tr_{dt}_" + "x".join(str(v) for v in shape)
            out[key + "_src"] = sv
            out[key + "_c2f"] = r.transpose(sv, shape, True)
            out[key + "_f2c"] = r.transpose(sv, shape, False)
    np.savez_compressed(OUT, **out)
    print(OUT, len(out), "arrays")


if __name__ == "__main__":
    main()
