"""GPU parity tests for LinInterp: fused / setup / apply entry points against the CPU oracle, bit for bit
(the reference's 4-statement formula in order, IEEE division), for scalar and Pack<.,N> layouts, plus the
reference's own property tests (tests/algorithm/lin_interp_test.cpp) run on the GPU."""
import os

import numpy as np
import pytest
import torch

from helpers import bits_equal, eps
from ekat_b200 import synth
from ekat_b200.lin_interp import LinInterp, fill

pytestmark = pytest.mark.gpu
DT = {"f64": (np.float64, torch.float64), "f32": (np.float32, torch.float32)}
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def gpu_interp(dev, tdt, x1, x2, y1, km1, km2, N, mode="fused"):
    ncol = x1.shape[0]
    li = LinInterp(ncol, km1, km2, tdt, N, device=dev)
    tx1, tx2, ty1 = (torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (x1, x2, y1))
    ty2 = torch.full(ty1.shape[:-1] + (li.pitch2,), -777.0, dtype=tdt, device=dev)
    if mode == "fused":
        li.run(tx1, tx2, ty1, ty2, write_map=True)
    else:
        li.setup(tx1, tx2)
        li.lin_interp(tx1, tx2, ty1, ty2)
    torch.cuda.synchronize()
    return ty2.cpu().numpy(), li.indx_map.cpu().numpy()


@pytest.mark.parametrize("dt", ["f64", "f32"])
@pytest.mark.parametrize("mode", ["fused", "setup_apply"])
def test_parity_bitwise(cuda, port, dt, mode):
    npdt, tdt = DT[dt]
    rng = np.random.default_rng(11)
    for km1, km2, N, ncol in [(72, 128, 1, 301), (72, 128, 16, 301), (128, 128, 1, 77), (10, 10, 1, 5), (100, 37, 8, 64),
                              (2, 5, 1, 3), (33, 70, 4, 130), (73, 129, 1, 41), (257, 300, 2, 19)]:
        x1, y1, x2 = synth.lin_interp_inputs(npdt, ncol, km1, km2, N, col0=17)
        x2[1, 0] = x1[1, 0] - 0.75                      # extrapolate below
        x2[1, km2 - 1] = x1[1, km1 - 1] + 1.25          # and above
        m = min(km1, km2)
        x2[2, :m] = x1[2, :m]                           # exact source points
        x2[0, :km2] = rng.permutation(x2[0, :km2])      # targets need not be sorted
        y1f = np.stack([y1, 2 * y1 + 1, -y1]).astype(npdt)
        want, wim = port.lin_interp(x1, x2, y1f, km1, km2, N)
        got, gim = gpu_interp(cuda, tdt, x1, x2, y1f, km1, km2, N, mode)
        assert np.array_equal(gim, wim), (km1, km2, N)
        assert bits_equal(got[..., :km2], want[..., :km2]), (km1, km2, N)
        assert (got[..., km2:] == -777.0).all(), "pack padding lanes must not be written"


@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_golden_fixture(cuda, dt):
    npdt, tdt = DT[dt]
    g = np.load(os.path.join(GOLD, "lin_interp.npz"), allow_pickle=True)
    for gdt, km1, km2, N, ncol, col0 in g["meta"]:
        if gdt != dt:
            continue
        x1, y1, _ = fill(tdt, ncol, km1, km2, N, col0=col0, device=cuda)
        key = f"{dt}_k{km1}_{km2}_p{N}"
        x1, y1 = x1.cpu().numpy(), y1.cpu().numpy()
        y1f = np.stack([y1, 2.0 * y1 + 1.0]).astype(npdt)
        got, gim = gpu_interp(cuda, tdt, x1, g[key + "_x2"], y1f, km1, km2, N)
        assert np.array_equal(gim, g[key + "_im"]), key
        assert bits_equal(got[..., :km2], g[key + "_y2"][..., :km2]), key


@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_device_generator_equals_numpy_generator(cuda, dt):
    npdt, tdt = DT[dt]
    for km1, km2, N in [(72, 128, 16), (33, 70, 4), (128, 128, 1)]:
        a = fill(tdt, 200, km1, km2, N, col0=5, device=cuda)
        b = synth.lin_interp_inputs(npdt, 200, km1, km2, N, col0=5)
        for x, y in zip(a, b):
            assert bits_equal(x.cpu().numpy(), y)


@pytest.mark.parametrize("dt", ["f64", "f32"])
@pytest.mark.parametrize("N", [1, 16])
def test_reference_properties_on_gpu(cuda, dt, N):
    """lin_interp_identity / _avg / _down_sampling / _monotone (lin_interp_test.cpp:94-422)."""
    npdt, tdt = DT[dt]
    rng = np.random.default_rng(5)
    for it in range(6):
        km1 = int(rng.integers(10, 101))
        ncol = 10
        x1, y1, _ = synth.lin_interp_inputs(npdt, ncol, km1, km1, N, col0=it * 10)
        y2, _ = gpu_interp(cuda, tdt, x1, x1, y1, km1, km1, N)
        assert bits_equal(y2[:, :km1], y1[:, :km1])                                         # identity, exact
        km2 = km1 - 1
        p2 = (km2 + N - 1) // N * N
        x2 = np.zeros((ncol, p2), npdt)
        x2[:, :km2] = (x1[:, :km2].astype(np.float64) + x1[:, 1:km1]) / 2
        y2, _ = gpu_interp(cuda, tdt, x1, x2, y1, km1, km2, N)
        avg = (y1[:, :km2].astype(np.float64) + y1[:, 1:km1]) / 2
        assert np.allclose(y2[:, :km2], avg, rtol=200 * eps(npdt), atol=2e4 * eps(npdt))     # averaging
        _, _, x2m = synth.lin_interp_inputs(npdt, ncol, km1, 2 * km1, N, col0=it * 10)
        y2, _ = gpu_interp(cuda, tdt, x1, x2m, y1, km1, 2 * km1, N)
        lo, hi = y1[:, :km1].min(1, keepdims=True), y1[:, :km1].max(1, keepdims=True)
        assert (y2[:, :2 * km1] >= lo).all() and (y2[:, :2 * km1] <= hi).all()              # monotone
    km2, km1 = 37, 74
    x1, y1, _ = synth.lin_interp_inputs(npdt, 10, km1, km2, N)
    p2 = (km2 + N - 1) // N * N
    x2 = np.zeros((10, p2), npdt)
    x2[:, :km2] = x1[:, 0:km1:2]
    y2, _ = gpu_interp(cuda, tdt, x1, x2, y1, km1, km2, N)
    assert bits_equal(y2[:, :km2], np.ascontiguousarray(y1[:, 0:km1:2]))                    # down-sampling, exact


def test_edge_cases_and_tma_fallback(cuda, port):
    from ekat_b200._lib import lib
    # empty problem (lin_interp_test.cpp:47-92)
    li = LinInterp(0, 0, 0, torch.float64, 16, device=cuda)
    e = torch.empty((0, 0), dtype=torch.float64, device=cuda)
    li.setup(e, e)
    li.lin_interp(e, e, e, e.clone())
    li.run(e, e, e, e.clone())
    # wrong pitch is rejected (the reference's repack divisibility requirement)
    li = LinInterp(4, 72, 128, torch.float64, 16, device=cuda)
    with pytest.raises(ValueError):
        li.setup(torch.zeros((4, 72), dtype=torch.float64, device=cuda), torch.zeros((4, 128), dtype=torch.float64, device=cuda))
    # the TMA path and the plain-load path produce the same bits
    x1, y1, x2 = synth.lin_interp_inputs(np.float64, 1000, 72, 128, 1)
    want, _ = port.lin_interp(x1, x2, y1, 72, 128, 1)
    try:
        for variant in (0, 1):
            for cb in (0, 2, 7, 16):
                lib().ekat_b200_set_tuning(cb, 0, variant)
                got, _ = gpu_interp(cuda, torch.float64, x1, x2, y1, 72, 128, 1)
                assert bits_equal(got, want), (variant, cb)
    finally:
        lib().ekat_b200_set_tuning(0, 0, 0)


def test_config2_full_size(cuda, port):
    """BASELINE config 2: 21,600 columns, 72 -> 128 levels, Pack<double,16> vs scalar layouts give the same bits,
    and a slab agrees with the oracle bit for bit."""
    outs = []
    for N in (1, 16):
        x1, y1, x2 = fill(torch.float64, 21600, 72, 128, N, device=cuda)
        li = LinInterp(21600, 72, 128, torch.float64, N, device=cuda)
        y2 = torch.zeros((21600, li.pitch2), dtype=torch.float64, device=cuda)
        li.run(x1, x2, y1, y2)
        outs.append(y2[:, :128].cpu().numpy())
        c0 = 20000
        want, _ = port.lin_interp(x1[c0:c0 + 50].cpu().numpy(), x2[c0:c0 + 50].cpu().numpy(), y1[c0:c0 + 50].cpu().numpy(), 72, 128, N)
        assert bits_equal(outs[-1][c0:c0 + 50], want[:, :128])
    assert bits_equal(np.ascontiguousarray(outs[0]), np.ascontiguousarray(outs[1]))


@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_level_major_parity_bitwise(cuda, port, dt):
    """ekat_b200_lin_interp_lm_*: the same interpolation on level-major operands ([level][column], the tridiagonal
    kernels' native layout), several fields sharing the coordinates, ragged column counts, in place when km2 == km1."""
    from ekat_b200.lin_interp import lin_interp_lm
    npdt, tdt = DT[dt]
    rng = np.random.default_rng(5)
    for km1, km2, ncol, pad, nf in [(128, 128, 1000, 0, 10), (72, 128, 301, 3, 3), (10, 10, 5, 3, 1), (100, 37, 64, 0, 2),
                                    (2, 5, 3, 1, 1), (73, 129, 41, 3, 4), (256, 200, 70, 2, 2)]:
        x1, y1, x2 = synth.lin_interp_inputs(npdt, ncol, km1, km2, 1, col0=3)
        x2[1, 0] = x1[1, 0] - 0.75
        x2[1, km2 - 1] = x1[1, km1 - 1] + 1.25
        m = min(km1, km2)
        x2[2, :m] = x1[2, :m]
        x2[0, :km2] = rng.permutation(x2[0, :km2])
        y1f = np.stack([(f + 1) * y1 - f for f in range(nf)]).astype(npdt)          # [nf, ncol, km1]
        want, _ = port.lin_interp(x1, x2, y1f, km1, km2, 1)                          # [nf, ncol, km2]
        lev = lambda a: torch.from_numpy(np.pad(np.ascontiguousarray(np.swapaxes(a, -1, -2)), [(0, 0)] * (a.ndim - 1) + [(0, pad)])).to(cuda)
        tx1, tx2, ty1 = lev(x1), lev(x2), lev(y1f)
        ty2 = torch.full((nf, km2, ncol + pad), -777.0, dtype=tdt, device=cuda)
        lin_interp_lm(tx1, tx2, ty1, ty2, ncol=ncol)
        got = ty2.cpu().numpy()
        assert bits_equal(np.ascontiguousarray(got[:, :, :ncol].transpose(0, 2, 1)), want), (km1, km2, ncol, nf)
        assert (got[:, :, ncol:] == -777.0).all(), "columns beyond ncol must not be written"
        if km1 == km2:  # in place
            lin_interp_lm(tx1, tx2, ty1, ty1, ncol=ncol)
            assert bits_equal(np.ascontiguousarray(ty1.cpu().numpy()[:, :, :ncol].transpose(0, 2, 1)), want)


@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_nan_targets_follow_the_gpu_upper_bound_variant(cuda, port, dt):
    """ekat::upper_bound has two variants (src/kokkos/ekat_upper_bound.hpp:28-53): the EKAT_ENABLE_GPU build's own
    loop with the predicate `value >= *it`, and std::upper_bound (`!(value < *it)`) on host builds. They agree on
    every ordered input and differ only for a NaN target: every `>=` is false, so the GPU variant returns index 0
    where std::upper_bound returns km1 - 1. This library replaces the GPU build, so it matches the GPU variant
    (and so does the oracle port, which restates that loop): the index map of a NaN target is 0 and its value NaN.
    The other targets of the column are unaffected."""
    npdt, tdt = DT[dt]
    for km1, km2, N in [(72, 128, 1), (128, 128, 16), (33, 70, 4)]:
        x1, y1, x2 = synth.lin_interp_inputs(npdt, 40, km1, km2, N, col0=3)
        x2[5, 7] = np.nan
        x2[9, 0] = np.nan
        x2[9, km2 - 1] = np.nan
        want, wim = port.lin_interp(x1, x2, y1, km1, km2, N)
        for mode in ("fused", "setup_apply"):
            got, gim = gpu_interp(cuda, tdt, x1, x2, y1, km1, km2, N, mode)
            assert np.array_equal(gim, wim)
            assert gim[5, 7] == 0 and gim[9, 0] == 0 and gim[9, km2 - 1] == 0
            assert np.isnan(got[5, 7]) and np.isnan(got[9, 0]) and np.isnan(got[9, km2 - 1])
            ok = ~np.isnan(want[:, :km2])
            assert bits_equal(got[:, :km2][ok], want[:, :km2][ok])
