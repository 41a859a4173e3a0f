"""CPU suite, part 1: pin the oracle.

The plain-C port (oracle/ekat_oracle.c) is checked against
  (1) the known answers of SURVEY.md section 8(c) (tests/golden/known_answers.json),
  (2) the golden fixtures generated from the unchanged reference headers (oracle/gen_golden.py),
  (3) oracle/_ref itself, bit for bit, whenever it is present,
and then re-states the reference's own property tests against it.
"""
import json
import os

import numpy as np
import pytest

from helpers import NROWS, NROWS_BIG, bits_equal, eps, rel_diff, u64sum
from ekat_b200 import synth

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
DT = {"f64": np.float64, "f32": np.float32}


def test_known_answers_survey_8c(port):
    known = json.load(open(os.path.join(GOLD, "known_answers.json")))
    # the table printed in SURVEY.md section 8(c): spot-pin two literal rows here as well
    assert known["tridiag"][0]["u64sum_bfb"] == "0xfe02368e7d0c3a5c"
    assert known["tridiag"][4]["u64sum_bfb"] == "0x2bf011b5c5"
    assert known["lin_interp"][0]["u64sum"] == "0xcf15bf7efdfbf8fe"
    for row in known["tridiag"]:
        dtype = DT[row["dtype"]]
        A, X = port.fill_tridiag(dtype, 1, row["nrow"], row["nA"], row["nrhs"], col0=row["seed"], seed_mod=0)
        Xb = port.tridiag("bfb", A.copy(), X.copy())
        Xt = port.tridiag("thomas", A.copy(), X.copy())
        Xc = port.tridiag("cr", A.copy(), X.copy())
        assert bits_equal(Xb, Xt) == row["thomas_eq_bfb"]
        assert float(Xb.flat[0]).hex() == row["x_first"] and float(Xb.flat[-1]).hex() == row["x_last"]
        assert hex(u64sum(Xb)) == row["u64sum_bfb"]
        assert hex(u64sum(Xc)) == row["u64sum_cr"]
    km1, km2 = 72, 128
    for row in known["lin_interp"]:
        N = row["pack"]
        p1 = (km1 + N - 1) // N * N
        x1 = np.zeros((1, p1)); y1 = np.zeros((1, p1)); x2 = np.zeros((1, km2))
        k = np.arange(km1)
        x1[0, :km1] = k; y1[0, :km1] = 1.5 * ((k * k) % 17); x2[0] = np.arange(km2, dtype=np.float64) * 71.0 / 127.0
        y2, _ = port.lin_interp(x1, x2, y1, km1, km2, N)
        assert float(y2[0, 1]).hex() == row["y2_1"] and float(y2[0, 64]).hex() == row["y2_64"]
        assert float(y2[0, 127]).hex() == row["y2_127"] == (13.5).hex()  # exact end point
        assert hex(u64sum(y2[0, :km2])) == row["u64sum"]


def test_port_matches_golden_tridiag(port):
    g = np.load(os.path.join(GOLD, "tridiag.npz"), allow_pickle=True)
    for dt, nrow, nrhs, nA, ncol, col0 in g["meta"]:
        A, X = port.fill_tridiag(DT[dt], ncol, nrow, nA, nrhs, col0=col0)
        for s in ("thomas", "cr", "bfb"):
            a, x = A.copy(), X.copy()
            port.tridiag(s, a, x)
            key = f"{dt}_{s}_n{nrow}_r{nrhs}_a{nA}"
            assert bits_equal(x, g[key]), key
            if s == "bfb":  # clobbered diagonals: multipliers in dl, pivots in d (ekat_tridiag.hpp:299-302)
                assert bits_equal(a[:, 0], g[key + "_dl"]) and bits_equal(a[:, 1], g[key + "_d"]), key
    for dt in DT:
        A, X = port.fill_tridiag(DT[dt], 3, 43, 3, 3, col0=11)
        assert bits_equal(A, g[f"{dt}_fillA"]) and bits_equal(X, g[f"{dt}_fillX"])


def _golden_li_inputs(dt, km1, km2, N, ncol, col0, g):
    x1, y1, _ = synth.lin_interp_inputs(DT[dt], ncol, km1, km2, N, col0=col0)
    x2 = g[f"{dt}_k{km1}_{km2}_p{N}_x2"]
    return x1, np.stack([y1, 2.0 * y1 + 1.0]).astype(DT[dt]), x2


def test_port_matches_golden_lin_interp(port):
    g = np.load(os.path.join(GOLD, "lin_interp.npz"), allow_pickle=True)
    for dt, km1, km2, N, ncol, col0 in g["meta"]:
        x1, y1, x2 = _golden_li_inputs(dt, km1, km2, N, ncol, col0, g)
        y2, im = port.lin_interp(x1, x2, y1, km1, km2, N)
        key = f"{dt}_k{km1}_{km2}_p{N}"
        assert bits_equal(y2, g[key + "_y2"]), key
        assert np.array_equal(im, g[key + "_im"]), key
    for dt in DT:  # the synthetic generator itself is pinned
        x1, y1, x2 = synth.lin_interp_inputs(DT[dt], 5, 33, 70, 4, col0=3)
        assert bits_equal(x1, g[f"{dt}_gen_x1"]) and bits_equal(y1, g[f"{dt}_gen_y1"]) and bits_equal(x2, g[f"{dt}_gen_x2"])


@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_port_equals_compiled_reference(port, ref, dt):
    if ref is None:
        pytest.skip("oracle/_ref not built on this box")
    dtype = DT[dt]
    for nrow in NROWS_BIG:
        for nrhs, nA in [(1, 1), (5, 1), (5, 5), (13, 1)]:
            if nrow == 2048 and nrhs > 5:
                continue
            A, X = ref.fill_tridiag(dtype, 4, nrow, nA, nrhs, col0=32765)  # seeds wrap through 32768
            A2, X2 = port.fill_tridiag(dtype, 4, nrow, nA, nrhs, col0=32765)
            assert bits_equal(A, A2) and bits_equal(X, X2)
            for s in ("thomas", "thomas_team", "cr", "bfb"):
                if s == "thomas_team" and nA > 1:
                    continue
                a1, x1 = A.copy(), X.copy()
                a2, x2 = A.copy(), X.copy()
                ref.tridiag(s, a1, x1)
                port.tridiag(s, a2, x2)
                assert bits_equal(x1, x2) and bits_equal(a1, a2), (s, nrow, nrhs, nA)
    rng = np.random.default_rng(7)
    for km1, km2, N in [(72, 128, 16), (72, 128, 1), (10, 91, 2), (100, 100, 8)]:
        x1, y1, x2 = synth.lin_interp_inputs(dtype, 9, km1, km2, N, col0=100)
        x2[0, :km2] = np.sort(rng.uniform(x1[0, 0] - 2, x1[0, km1 - 1] + 2, km2)).astype(dtype)
        ya, ia = ref.lin_interp(x1, x2, y1, km1, km2, N)
        yb, ib = port.lin_interp(x1, x2, y1, km1, km2, N)
        assert bits_equal(ya, yb) and np.array_equal(ia, ib)


# ---------------------------------------------------------------------------------------------------------
# the reference's property tests, against the oracle (tests/algorithm/tridiag_tests_correctness.cpp:403-484)
# ---------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("dt", ["f64", "f32"])
@pytest.mark.parametrize("solver", ["thomas", "thomas_team", "cr", "bfb"])
def test_tridiag_property_residual(port, dt, solver):
    dtype = DT[dt]
    for nrow in NROWS_BIG:
        for nrhs in (1, 5, 9, 13):  # nrhs = 1..13 step 4 (:415-421)
            for nA in {1, nrhs}:
                if solver == "thomas_team" and nA > 1:
                    continue  # rank-2 A unsupported by the team version (:430-433)
                A, B = port.fill_tridiag(dtype, 1, nrow, nA, nrhs, col0=0, seed_mod=0)
                X = port.tridiag(solver, A.copy(), B.copy())
                assert port.residual(A, B, X) <= 50 * eps(dtype), (nrow, nrhs, nA)


def test_thomas_equals_bfb_and_cr_within_tolerance(port):
    for dtype, tolr in ((np.float64, 1e-13), (np.float32, 1e-5)):
        for nrow in NROWS:
            A, B = port.fill_tridiag(dtype, 2, nrow, 1, 3, col0=9)
            xb = port.tridiag("bfb", A.copy(), B.copy())
            xt = port.tridiag("thomas", A.copy(), B.copy())
            xc = port.tridiag("cr", A.copy(), B.copy())
            assert bits_equal(xb, xt)
            assert rel_diff(xb, xc) <= tolr


# tests/algorithm/lin_interp_test.cpp:94-166 / :168-258 / :260-335 / :337-422
@pytest.mark.parametrize("dt", ["f64", "f32"])
@pytest.mark.parametrize("N", [1, 16])
def test_lin_interp_properties(port, dt, N):
    dtype = DT[dt]
    rng = np.random.default_rng(3)
    for it in range(20):
        km1 = int(rng.integers(10, 101))
        ncol = 10
        x1, y1, _ = synth.lin_interp_inputs(dtype, ncol, km1, km1, N, col0=it * 10)
        # identity: x2 == x1  =>  y2 == y1 exactly
        y2, _ = port.lin_interp(x1, x1, y1, km1, km1, N)
        assert bits_equal(y2[:, :km1], y1[:, :km1])
        # averaging: x2 = midpoints, km2 = km1-1  =>  y2 ~ (y1[k]+y1[k+1])/2
        km2 = km1 - 1
        p2 = (km2 + N - 1) // N * N
        x2 = np.zeros((ncol, p2), dtype)
        x2[:, :km2] = (x1[:, :km2].astype(np.float64) + x1[:, 1:km1]) / 2
        y2, _ = port.lin_interp(x1, x2, y1, km1, km2, N)
        avg = (y1[:, :km2].astype(np.float64) + y1[:, 1:km1]) / 2
        assert np.allclose(y2[:, :km2], avg, rtol=200 * eps(dtype), atol=200 * eps(dtype) * 100)
        # monotone: targets strictly inside the source range => min(y1) <= y2 <= max(y1)
        _, _, x2m = synth.lin_interp_inputs(dtype, ncol, km1, 2 * km1, N, col0=it * 10)
        y2, _ = port.lin_interp(x1, x2m, y1, km1, 2 * km1, N)
        lo, hi = y1[:, :km1].min(1, keepdims=True), y1[:, :km1].max(1, keepdims=True)
        assert (y2[:, :2 * km1] >= lo).all() and (y2[:, :2 * km1] <= hi).all()
    # down-sampling: km1 = 2*km2, x2 = x1[::2]  =>  y2 == y1[::2] exactly
    km2 = 37
    km1 = 2 * km2
    x1, y1, _ = synth.lin_interp_inputs(dtype, 10, km1, km2, N)
    p2 = (km2 + N - 1) // N * N
    x2 = np.zeros((10, p2), dtype)
    x2[:, :km2] = x1[:, 0:km1:2]
    y2, _ = port.lin_interp(x1, x2, y1, km1, km2, N)
    assert bits_equal(y2[:, :km2], np.ascontiguousarray(y1[:, 0:km1:2]))


def test_lin_interp_pack_size_invariance_and_edges(port):
    x1a, y1a, x2a = synth.lin_interp_inputs(np.float64, 4, 72, 128, 1)
    x1b, y1b, x2b = synth.lin_interp_inputs(np.float64, 4, 72, 128, 16)
    ya, ia = port.lin_interp(x1a, x2a, y1a, 72, 128, 1)
    yb, ib = port.lin_interp(x1b, x2b, y1b, 72, 128, 16)
    assert bits_equal(ya[:, :128], yb[:, :128]) and np.array_equal(ia, ib)
    # extrapolation below / above the range uses the first / last interval (impl :119-131,178-183)
    x1 = np.array([[0.0, 1.0, 3.0]]); y1 = np.array([[10.0, 20.0, 0.0]]); x2 = np.array([[-1.0, 4.0, 3.0, 0.0]])
    y2, im = port.lin_interp(x1, x2, y1, 3, 4, 1)
    assert im.tolist() == [[0, 2, 2, 0]]
    assert y2.tolist() == [[0.0, -10.0, 0.0, 10.0]]
    # empty problems are a silent no-op (lin_interp_test.cpp:47-92)
    y2, im = port.lin_interp(np.zeros((0, 0)), np.zeros((0, 0)), np.zeros((0, 0)), 0, 0, 1)
    assert y2.shape == (0, 0)
