"""ReductionUtils::view_reduction: the oracle against the golden fixtures generated from the
unchanged reference header (oracle/gen_golden_reduction.py) and against oracle/_ref itself; the reference's own test
(tests/algorithm/reduction_tests.cpp:62-175) restated; and -- on a GPU -- the CUDA path against all of it, bit for bit,
through the C ABI."""
import os

import numpy as np
import pytest

from helpers import bits_equal

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reduction.npz")
DT = {"f64": np.float64, "f32": np.float32}


def gold_cases():
    g = np.load(GOLD)
    for key in sorted(k[:-2] for k in g.files if k.endswith("_x") and not k.startswith("adj_")):
        dt, tot, N, b, e = key.split("_")
        yield key, dt, int(tot), int(N), int(b), int(e), g[key + "_x"], g[key + "_ser"], g[key + "_pack"]


def test_port_matches_golden_reduction(port):
    n = 0
    for key, dt, tot, N, b, e, x, ser, pack in gold_cases():
        assert bits_equal(port.view_reduction(x, b, e, 0), ser), key
        assert bits_equal(port.view_reduction(x, b, e, N), pack), key
        n += 1
    assert n == 26


@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_port_equals_compiled_reference_reduction(port, ref, dt):
    if ref is None:
        pytest.skip("oracle/_ref not built on this box")
    rng = np.random.default_rng(7)
    for tot, N in [(128, 16), (72, 16), (129, 4), (33, 8), (5, 2), (1, 1)]:
        pitch = -(-tot // N) * N
        x = np.full((64, pitch), np.nan, DT[dt])
        x[:, :tot] = (rng.standard_normal((64, tot)) * 100).astype(DT[dt])
        for b, e in [(0, tot), (tot // 3, tot - tot // 4), (min(1, tot), tot), (0, 0)]:
            for packn in (0, N):
                assert bits_equal(port.view_reduction(x, b, e, packn), ref.view_reduction(x, b, e, packn)), (tot, N, b, e, packn)


@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_reference_view_reduction_test_restated(port, dt):
    """tests/algorithm/reduction_tests.cpp:62-175: data(k) = 1/(k+1), NaN pack padding; Serialize == the serial sum
    exactly, else within 10 eps."""
    for tot, N, b, e in [(8, 1, 0, 8), (8, 1, 2, 5), (8, 4, 0, 8), (7, 4, 0, 7), (3, 4, 0, 3), (16, 4, 4, 11)]:
        pitch = -(-tot // N) * N
        x = np.full((1, pitch), np.nan, DT[dt])
        x[0, :tot] = (1.0 / (np.arange(tot) + 1.0)).astype(DT[dt])
        serial = DT[dt](0)
        for k in range(b, e):
            serial = DT[dt](serial + x[0, k])
        assert port.view_reduction(x, b, e, 0)[0] == serial
        assert abs(port.view_reduction(x, b, e, N)[0] - serial) <= 10 * np.finfo(DT[dt]).eps


# ------------------------------------------------------------------------------------------------ GPU
def _gpu(x_np, b, e, serialize, N, layout, cuda, strided=False):
    import torch
    import ekat_b200
    from ekat_b200.reduction import view_reduction
    if layout == ekat_b200.COL_MAJOR:
        if strided:
            wide = torch.full((x_np.shape[0], x_np.shape[1] + 2 * max(N, 1)), float("nan"), dtype=torch.from_numpy(x_np).dtype, device=cuda)
            wide[:, :x_np.shape[1]] = torch.from_numpy(x_np).to(cuda)
            x = wide[:, :x_np.shape[1]]
        else:
            x = torch.from_numpy(x_np).to(cuda)
        return view_reduction(x, b, e, serialize=serialize, pack=N, layout=layout).cpu().numpy()
    x = torch.from_numpy(np.ascontiguousarray(x_np.T)).to(cuda)  # [nlev, ncol]
    return view_reduction(x, b, e, serialize=serialize, pack=N, layout=layout).cpu().numpy()


@pytest.mark.gpu
def test_gpu_view_reduction_matches_golden(cuda):
    import ekat_b200
    for key, dt, tot, N, b, e, x, ser, pack in gold_cases():
        for layout in (ekat_b200.COL_MAJOR, ekat_b200.LEV_MAJOR):
            assert bits_equal(_gpu(x, b, e, True, N, layout, cuda), ser), (key, layout)
            assert bits_equal(_gpu(x, b, e, False, N, layout, cuda), pack), (key, layout)
        assert bits_equal(_gpu(x, b, e, False, N, ekat_b200.COL_MAJOR, cuda, strided=True), pack), key


@pytest.mark.gpu
@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_gpu_view_reduction_matches_oracle(cuda, port, dt):
    """Bit for bit against the oracle on seeded columns: ragged column counts (tiles of 32), every pack size the
    reference instantiates, ranges with partial first/last packs, long columns (direct fallback), empty batches."""
    import ekat_b200
    rng = np.random.default_rng(11)
    for ncol, tot, N in [(1, 128, 16), (31, 72, 16), (33, 72, 16), (1000, 128, 16), (4097, 128, 1), (257, 129, 4),
                         (65, 43, 8), (70, 2048, 16), (40, 20000, 16), (0, 128, 16)]:
        pitch = -(-tot // N) * N
        x = np.full((ncol, pitch), np.nan, DT[dt])
        x[:, :tot] = ((rng.random((ncol, tot)) * 2 - 1) * 10.0 ** rng.integers(-2, 3, (ncol, tot))).astype(DT[dt])
        for b, e in [(0, tot), (tot // 5, tot - tot // 7), (1, min(3, tot))]:
            for serialize in (True, False):
                want = port.view_reduction(x, b, e, 0 if serialize else N)
                for layout in (ekat_b200.COL_MAJOR, ekat_b200.LEV_MAJOR):
                    got = _gpu(x, b, e, serialize, N, layout, cuda)
                    assert bits_equal(got, want), (ncol, tot, N, b, e, serialize, layout)


@pytest.mark.gpu
def test_gpu_view_reduction_full_size_property(cuda):
    """BASELINE config 3 size (345,600 x 128): columns of ones sum to the level count exactly; a permutation of the
    columns permutes the results; both orders; both layouts."""
    import torch
    import ekat_b200
    from ekat_b200.reduction import view_reduction
    ncol, nlev = 345_600, 128
    g = torch.Generator(device=cuda).manual_seed(5)
    x = torch.rand((ncol, nlev), dtype=torch.float64, device=cuda, generator=g)
    ones = torch.ones_like(x)
    for ser in (True, False):
        assert torch.equal(view_reduction(ones, 0, nlev, serialize=ser, pack=16), torch.full((ncol,), float(nlev), dtype=torch.float64, device=cuda))
        r = view_reduction(x, 0, nlev, serialize=ser, pack=16)
        perm = torch.randperm(ncol, device=cuda, generator=g)
        assert torch.equal(view_reduction(x[perm].contiguous(), 0, nlev, serialize=ser, pack=16), r[perm])
        rl = view_reduction(x.t().contiguous(), 0, nlev, serialize=ser, pack=16, layout=ekat_b200.LEV_MAJOR)
        assert torch.equal(rl, r)
        assert torch.allclose(r, x.sum(1), rtol=1e-13, atol=0)


# ------------------------------------------------------------------------------------------------ adj_diff (8f-3)
def adj_cases():
    g = np.load(GOLD)
    for key in sorted(k[:-2] for k in g.files if k.startswith("adj_") and k.endswith("_x")):
        yield key, g[key + "_x"], g[key + "_fwd"], g[key + "_bwd"]


def test_port_matches_golden_adj_diff(port):
    n = 0
    for key, x, fwd, bwd in adj_cases():
        assert bits_equal(port.adj_diff(x, True), fwd) and bits_equal(port.adj_diff(x, False), bwd), key
        n += 1
    assert n == 14
    # the reference's own check (tests/pack/pack_kokkos_utils.cpp:1051-1095): (i+1)^2 - i^2 = 2i+1, i^2 - (i-1)^2 = 2i-1
    x = (np.arange(16, dtype=np.float64) ** 2)[None]
    f, b = port.adj_diff(x, True)[0], port.adj_diff(x, False)[0]
    assert np.array_equal(f[:-1], 2 * np.arange(15) + 1.0) and f[-1] == -225.0
    assert np.array_equal(b[1:], 2 * np.arange(1, 16) - 1.0) and b[0] == 0.0


@pytest.mark.gpu
def test_gpu_adj_diff_matches_golden_and_oracle(cuda, port):
    import torch
    import ekat_b200
    from ekat_b200.reduction import adj_diff
    for key, x, fwd, bwd in adj_cases():
        xd = torch.from_numpy(x).to(cuda)
        assert bits_equal(adj_diff(xd, True).cpu().numpy(), fwd), key
        assert bits_equal(adj_diff(xd, False).cpu().numpy(), bwd), key
        xl = xd.t().contiguous()  # [n, ncol]
        assert bits_equal(adj_diff(xl, True, layout=ekat_b200.LEV_MAJOR).t().contiguous().cpu().numpy(), fwd), key
        assert bits_equal(adj_diff(xl, False, layout=ekat_b200.LEV_MAJOR).t().contiguous().cpu().numpy(), bwd), key
    rng = np.random.default_rng(5)
    for dt in (np.float64, np.float32):
        for ncol, n in [(1, 1), (33, 129), (1000, 128), (4097, 72), (5, 3000), (0, 128)]:
            x = (rng.standard_normal((ncol, n)) * 10).astype(dt)
            wide = torch.zeros((ncol, n + 3), dtype=torch.from_numpy(x).dtype, device=cuda)  # odd pitch: the scalar path
            wide[:, :n] = torch.from_numpy(x).to(cuda)
            for f in (True, False):
                want = port.adj_diff(x, f)
                assert bits_equal(adj_diff(torch.from_numpy(x).to(cuda), f).cpu().numpy(), want), (ncol, n, f)
                assert bits_equal(adj_diff(wide[:, :n], f).cpu().numpy(), want), (ncol, n, f, "strided")
                if ncol:
                    got = adj_diff(torch.from_numpy(np.ascontiguousarray(x.T)).to(cuda), f, layout=ekat_b200.LEV_MAJOR)
                    assert bits_equal(got.cpu().numpy(), np.ascontiguousarray(want.T)), (ncol, n, f, "lev")
