"""Generate tests/golden/reduction.npz from oracle/_ref (the UNCHANGED reference header
src/algorithm/ekat_reduction_utils.hpp compiled in place). Run in the build container only:

    python -m oracle.gen_golden_reduction

TEST INFRASTRUCTURE. Cases = the reference's own (tests/algorithm/reduction_tests.cpp:139-175: total size, pack size,
[begin,end), data(k) = 1/(k+1), NaN in the pack padding) plus column-shaped ones (72 and 128 levels, Pack 16, ragged
ranges, begin and end inside one pack) on seeded random data. Stored: inputs and the reference's results for
Serialize = true and false."""
import os

import numpy as np

import oracle

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "reduction.npz")

# (total scalars, pack size, begin, end)
REF_CASES = [(8, 1, 0, 8), (8, 1, 2, 5), (8, 4, 0, 8), (7, 4, 0, 7), (3, 4, 0, 3), (16, 4, 4, 11)]
COL_CASES = [(72, 16, 0, 72), (128, 16, 0, 128), (128, 16, 5, 117), (128, 8, 3, 4), (64, 4, 1, 3), (129, 2, 1, 129),
             (100, 1, 0, 100)]


def inputs(dtype, tot, N, ncol, seed):
    pitch = -(-tot // N) * N
    x = np.full((ncol, pitch), np.nan, dtype)
    if seed is None:  # the reference's series
        x[:, :tot] = (1.0 / (np.arange(tot) + 1.0)).astype(dtype)
    else:
        rng = np.random.default_rng(seed)
        x[:, :tot] = ((rng.random((ncol, tot)) * 2 - 1) * 10.0 ** rng.integers(-3, 4, (ncol, tot))).astype(dtype)
    return x


def main():
    r = oracle.ref()
    assert r is not None, "oracle/_ref/libekat_ref.so missing: run `make -C oracle` where /root/reference exists"
    out = {}
    n = 0
    for dt in ("f64", "f32"):
        dtype = np.float64 if dt == "f64" else np.float32
        for cases, ncol, seeded in ((REF_CASES, 1, False), (COL_CASES, 40, True)):
            for (tot, N, b, e) in cases:
                x = inputs(dtype, tot, N, ncol, 1000 + n if seeded else None)
                key = f"{dt}_{tot}_{N}_{b}_{e}"
                out[key + "_x"] = x
                out[key + "_ser"] = r.view_reduction(x, b, e, 0)
                out[key + "_pack"] = r.view_reduction(x, b, e, N)
                n += 1
    # ekat::adj_diff (src/pack/ekat_pack_kokkos.hpp:100-164): the reference test's data i^2
    # (tests/pack/pack_kokkos_utils.cpp:1040-1052) and seeded columns, scalar views and Pack 4 / 16 views
    for dt in ("f64", "f32"):
        dtype = np.float64 if dt == "f64" else np.float32
        for pitch, N in [(16, 0), (16, 4), (128, 16), (80, 16), (72, 0), (7, 1), (1, 0)]:
            rng = np.random.default_rng(77 + pitch + N)
            x = (rng.standard_normal((24, pitch)) * 50).astype(dtype)
            x[0] = (np.arange(pitch, dtype=np.float64) ** 2).astype(dtype)
            key = f"adj_{dt}_{pitch}_{N}"
            out[key + "_x"] = x
            out[key + "_fwd"] = r.adj_diff(x, True, N)
            out[key + "_bwd"] = r.adj_diff(x, False, N)
    np.savez_compressed(OUT, **out)
    print(OUT, n, "cases")


if __name__ == "__main__":
    main()
