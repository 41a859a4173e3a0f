"""Generate tests/golden/*.npz and known_answers.json from oracle/_ref (the UNCHANGED reference headers
compiled in place). Run in the build container only (needs /root/reference at build time of _ref):

    python -m oracle.gen_golden

TEST INFRASTRUCTURE. The fixtures pin (a) the plain-C port on any box and (b) the CUDA path on the GPU box,
where /root/reference does not exist. Inputs are the reference's deterministic fills
(tests/algorithm/tridiag_tests.hpp:11-40) or the counter-based generator of SURVEY.md section 8d, so only
shapes + outputs need to be stored; a few raw inputs are stored too so the fills themselves are pinned.
"""
import json
import os

import numpy as np

import oracle
from ekat_b200 import synth

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")

# nrow list of tests/algorithm/tridiag_tests_correctness.cpp:410-414 (2048 kept out of the fixture for size)
NROWS = [1, 2, 3, 4, 5, 8, 10, 16, 32, 43, 63, 64, 65, 111, 128, 129]


def u64sum(a):
    a = np.ascontiguousarray(a)
    if a.dtype == np.float64:
        return int(a.view(np.uint64).sum(dtype=np.uint64))
    return int(a.view(np.uint32).astype(np.uint64).sum(dtype=np.uint64))


def main():
    r = oracle.ref()
    assert r is not None, "oracle/_ref/libekat_ref.so missing: run `make -C oracle` where /root/reference exists"
    os.makedirs(OUT, exist_ok=True)
    known = {"source": r.info(), "tridiag": [], "lin_interp": []}

    # --- SURVEY.md section 8(c) known answers, regenerated (must equal the table there) -------------------
    for dt, nrow, nrhs, nA, seed in [("f64", 72, 1, 1, 0), ("f64", 128, 10, 1, 5), ("f64", 128, 10, 10, 5),
                                     ("f64", 129, 3, 1, 7), ("f32", 72, 1, 1, 0), ("f32", 128, 10, 1, 5)]:
        dtype = np.float64 if dt == "f64" else np.float32
        A, X = r.fill_tridiag(dtype, 1, nrow, nA, nrhs, col0=seed, seed_mod=0)
        Xb = r.tridiag("bfb", A.copy(), X.copy())
        Xt = r.tridiag("thomas", A.copy(), X.copy())
        Xc = r.tridiag("cr", A.copy(), X.copy())
        known["tridiag"].append({
            "dtype": dt, "nrow": nrow, "nrhs": nrhs, "nA": nA, "seed": seed,
            "thomas_eq_bfb": bool(np.array_equal(Xb, Xt)),
            "x_first": float(Xb.flat[0]).hex(), "x_last": float(Xb.flat[-1]).hex(),
            "u64sum_bfb": hex(u64sum(Xb)), "u64sum_cr": hex(u64sum(Xc)),
            "cr_vs_bfb": float(np.abs(Xc.astype(np.float64) - Xb).max() / np.abs(Xb).max()),
        })
    km1, km2 = 72, 128
    for N in (1, 16):
        p1 = (km1 + N - 1) // N * N
        x1 = np.zeros((1, p1)); y1 = np.zeros((1, p1)); x2 = np.zeros((1, km2))
        k = np.arange(km1)
        x1[0, :km1] = k; y1[0, :km1] = 1.5 * ((k * k) % 17); x2[0] = np.arange(km2, dtype=np.float64) * 71.0 / 127.0
        y2, im = r.lin_interp(x1, x2, y1, km1, km2, N)
        known["lin_interp"].append({"pack": N, "y2_1": float(y2[0, 1]).hex(), "y2_64": float(y2[0, 64]).hex(),
                                    "y2_127": float(y2[0, 127]).hex(), "u64sum": hex(u64sum(y2[0, :km2]))})
    with open(os.path.join(OUT, "known_answers.json"), "w") as f:
        json.dump(known, f, indent=1)

    # --- tridiag fixture: every solver x shape of the reference's property test, 3 columns each ----------
    tri = {}
    meta = []
    for dt in ("f64", "f32"):
        dtype = np.float64 if dt == "f64" else np.float32
        for nrow in NROWS:
            for nrhs, nA in [(1, 1), (3, 1), (3, 3)]:
                ncol, col0 = 3, 11
                A, X = r.fill_tridiag(dtype, ncol, nrow, nA, nrhs, col0=col0)
                for s in ("thomas", "cr", "bfb"):
                    a, x = A.copy(), X.copy()
                    r.tridiag(s, a, x)
                    key = f"{dt}_{s}_n{nrow}_r{nrhs}_a{nA}"
                    tri[key] = x
                    if s == "bfb":
                        tri[key + "_dl"] = a[:, 0].copy()
                        tri[key + "_d"] = a[:, 1].copy()
                meta.append([dt, nrow, nrhs, nA, ncol, col0])
    # raw inputs for one case per dtype pin the fills themselves
    for dt in ("f64", "f32"):
        dtype = np.float64 if dt == "f64" else np.float32
        A, X = r.fill_tridiag(dtype, 3, 43, 3, 3, col0=11)
        tri[f"{dt}_fillA"] = A
        tri[f"{dt}_fillX"] = X
    np.savez_compressed(os.path.join(OUT, "tridiag.npz"), meta=np.array(meta, dtype=object), **tri)

    # --- LinInterp fixture ---------------------------------------------------------------------------------
    li = {}
    meta = []
    for dt in ("f64", "f32"):
        dtype = np.float64 if dt == "f64" else np.float32
        for km1, km2, N in [(72, 128, 1), (72, 128, 16), (128, 128, 1), (10, 10, 1), (100, 37, 8), (2, 5, 1), (33, 70, 4)]:
            ncol, col0 = 5, 3
            x1, y1, x2 = synth.lin_interp_inputs(dtype, ncol, km1, km2, N, col0=col0)
            # widen the coverage: column 1 extrapolates on both sides, column 2 hits source points exactly
            x2[1, 0] = x1[1, 0] - 0.75
            x2[1, km2 - 1] = x1[1, km1 - 1] + 1.25
            m = min(km1, km2)
            x2[2, :m] = x1[2, :m]
            if km2 > m:
                x2[2, m:km2] = x1[2, km1 - 1]
            y2, im = r.lin_interp(x1, x2, np.stack([y1, 2.0 * y1 + 1.0]).astype(dtype), km1, km2, N)
            key = f"{dt}_k{km1}_{km2}_p{N}"
            li[key + "_x2"] = x2
            li[key + "_y2"] = y2
            li[key + "_im"] = im
            meta.append([dt, km1, km2, N, ncol, col0])
    for dt in ("f64", "f32"):
        dtype = np.float64 if dt == "f64" else np.float32
        x1, y1, x2 = synth.lin_interp_inputs(dtype, 5, 33, 70, 4, col0=3)
        li[f"{dt}_gen_x1"], li[f"{dt}_gen_y1"], li[f"{dt}_gen_x2"] = x1, y1, x2
    np.savez_compressed(os.path.join(OUT, "lin_interp.npz"), meta=np.array(meta, dtype=object), **li)
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
