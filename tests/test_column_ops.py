"""SURVEY.md section 8f-3 / 8f-1 rows built in round 2: ekat::shift_left / shift_right, ekat::index_and_shift,
ekat::transpose (rank 2 and 3) and the host_to_device row repacking -- the oracle port against golden fixtures
generated from the unchanged reference headers (oracle/gen_golden_column_ops.py) and against oracle/_ref itself, and --
on a GPU -- the CUDA path against all of it, bit for bit, through the C ABI, in both layouts."""
import os

import numpy as np
import pytest

from helpers import bits_equal

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "column_ops.npz")
DT = {"f64": np.float64, "f32": np.float32}


def same(a, b):
    """bitwise equal, NaNs (ekat::invalid) compared as NaNs"""
    a, b = np.ascontiguousarray(a), np.ascontiguousarray(b)
    na, nb = np.isnan(a), np.isnan(b)
    return a.shape == b.shape and np.array_equal(na, nb) and bits_equal(np.where(na, 0, a).astype(a.dtype), np.where(nb, 0, b).astype(b.dtype))


def shift_cases():
    g = np.load(GOLD)
    for key in sorted(k[:-2] for k in g.files if k.startswith("shift_") and k.endswith("_x")):
        _, dt, pitch, N = key.split("_")
        yield key, dt, int(pitch), int(N), g


def ias_cases():
    g = np.load(GOLD)
    for key in sorted(k[:-2] for k in g.files if k.startswith("ias_") and k.endswith("_a")):
        _, dt, n, m, N, sh = key.split("_")
        yield key, dt, int(n), int(m), int(N), int(sh), g


def tr_cases():
    g = np.load(GOLD)
    for key in sorted(k[:-4] for k in g.files if k.startswith("tr_") and k.endswith("_src")):
        _, dt, shp = key.split("_")
        yield key, dt, tuple(int(v) for v in shp.split("x")), g


def test_port_matches_golden_column_ops(port):
    n = 0
    for key, dt, pitch, N, g in shift_cases():
        x = g[key + "_x"]
        assert bits_equal(port.shift(x, +1, 0.0), g[key + "_right"]), key
        assert bits_equal(port.shift(x, -1, 0.0), g[key + "_left"]), key
        assert bits_equal(port.shift(x, +1, -3.5), g[key + "_right_fill"]), key
        assert bits_equal(port.shift(x, -1, 7.25), g[key + "_left_fill"]), key
        n += 1
    for key, dt, nn, m, N, sh, g in ias_cases():
        o0, o1 = port.index_and_shift(g[key + "_a"], g[key + "_idx"], sh)
        assert same(o0, g[key + "_out"]) and same(o1, g[key + "_shift"]), key
        n += 1
    for key, dt, shape, g in tr_cases():
        assert bits_equal(port.transpose(g[key + "_src"], shape, True), g[key + "_c2f"]), key
        assert bits_equal(port.transpose(g[key + "_src"], shape, False), g[key + "_f2c"]), key
        n += 1
    assert n == 2 * (6 + 5 + 5)


@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_port_equals_compiled_reference_column_ops(port, ref, dt):
    if ref is None:
        pytest.skip("oracle/_ref not built on this box")
    rng = np.random.default_rng(3)
    for pitch, N in [(128, 16), (80, 16), (12, 4), (5, 1)]:
        x = (rng.standard_normal((30, pitch)) * 9).astype(DT[dt])
        for s, fill in [(1, 0.0), (-1, 0.0), (1, 2.5), (-1, -1.0)]:
            assert bits_equal(port.shift(x, s, fill), ref.shift(x, s, fill, N)), (pitch, N, s)
    for n, m, N, sh in [(72, 128, 16, 1), (50, 64, 4, -1), (9, 16, 1, 2)]:
        a = rng.standard_normal((17, n)).astype(DT[dt])
        idx = rng.integers(0, n, (17, m)).astype(np.int32)  # in range themselves; the shift may leave the column
        p0, p1 = port.index_and_shift(a, idx, sh)
        r0, r1 = ref.index_and_shift(a, idx, sh, packn=N)
        assert same(p0, r0) and same(p1, r1), (n, m, N, sh)
    for shape in [(6, 7), (9, 8, 5), (64, 72, 10)]:
        sv = rng.standard_normal(int(np.prod(shape))).astype(DT[dt])
        for c2f in (True, False):
            assert bits_equal(port.transpose(sv, shape, c2f), ref.transpose(sv, shape, c2f)), (shape, c2f)


def test_transpose_is_numpy_transpose(port):
    """The reference's index formulas (ekat_math_utils.hpp:97-147) are the axis reversal of a C array."""
    rng = np.random.default_rng(5)
    a = rng.standard_normal((6, 5, 4))
    assert np.array_equal(port.transpose(a, a.shape, True).reshape(4, 5, 6), a.transpose(2, 1, 0))
    assert np.array_equal(port.transpose(a.transpose(2, 1, 0).copy(), a.shape, False).reshape(6, 5, 4), a)


# ------------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
def test_gpu_shift_matches_reference_headers(cuda):
    import torch
    import ekat_b200
    from ekat_b200.reduction import shift
    for key, dt, pitch, N, g in shift_cases():
        x = g[key + "_x"]
        xd = torch.from_numpy(x).to(cuda)
        xl = torch.from_numpy(np.ascontiguousarray(x.T)).to(cuda)
        for s, fill, name in [(1, 0.0, "_right"), (-1, 0.0, "_left"), (1, -3.5, "_right_fill"), (-1, 7.25, "_left_fill")]:
            assert bits_equal(shift(xd, s, fill).cpu().numpy(), g[key + name]), (key, name)
            got = shift(xl, s, fill, layout=ekat_b200.LEV_MAJOR).cpu().numpy().T
            assert bits_equal(got, g[key + name]), (key, name, "level-major")


@pytest.mark.gpu
@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_gpu_shift_general(cuda, port, dt):
    """Any shift distance, strided rows, big batch: against the oracle port."""
    import torch
    from ekat_b200.reduction import shift
    rng = np.random.default_rng(21)
    x = rng.standard_normal((5000, 80)).astype(DT[dt])
    big = torch.from_numpy(np.pad(x, ((0, 0), (0, 16)))).to(cuda)
    for s, fill in [(1, 0.0), (-1, 0.0), (3, 1.5), (-7, -2.0), (0, 9.0), (80, 4.0), (-200, 5.0)]:
        got = shift(big[:, :80], s, fill).cpu().numpy()
        assert bits_equal(got, port.shift(x, s, fill)), (s, fill)


@pytest.mark.gpu
def test_gpu_index_and_shift_matches_reference_headers(cuda):
    import torch
    import ekat_b200
    from ekat_b200.reduction import index_and_shift
    for key, dt, n, m, N, sh, g in ias_cases():
        a, idx = g[key + "_a"], g[key + "_idx"]
        ad, idd = torch.from_numpy(a).to(cuda), torch.from_numpy(idx).to(cuda)
        o0 = torch.full(idx.shape, -777.0, dtype=ad.dtype, device=cuda)
        o0, o1 = index_and_shift(ad, idd, sh, out=o0)
        assert same(o0.cpu().numpy(), g[key + "_out"]) and same(o1.cpu().numpy(), g[key + "_shift"]), key
        al, il = torch.from_numpy(np.ascontiguousarray(a.T)).to(cuda), torch.from_numpy(np.ascontiguousarray(idx.T)).to(cuda)
        l0 = torch.full(il.shape, -777.0, dtype=ad.dtype, device=cuda)
        l0, l1 = index_and_shift(al, il, sh, layout=ekat_b200.LEV_MAJOR, out=l0)
        assert same(l0.cpu().numpy().T, g[key + "_out"]) and same(l1.cpu().numpy().T, g[key + "_shift"]), (key, "level-major")


@pytest.mark.gpu
@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_gpu_index_and_shift_is_the_lin_interp_gather(cuda, port, dt):
    """index_and_shift<1> with LinInterp's index map gathers (y1(k1), y1(k1+1)) (ekat_pack_kokkos.hpp:71-79): with those
    two and the coordinates the interpolated value follows from the reference's 4 statements -- compare with LinInterp."""
    import torch
    from ekat_b200 import synth
    from ekat_b200.reduction import index_and_shift
    x1, y1, x2 = synth.lin_interp_inputs(DT[dt], 300, 72, 128, 1, col0=2)
    want, im = port.lin_interp(x1, x2, y1, 72, 128, 1)
    k1 = np.minimum(im, 70).astype(np.int32)  # interior brackets only: k1 + 1 stays in range
    yk, yk1 = index_and_shift(torch.from_numpy(y1).to(cuda), torch.from_numpy(k1).to(cuda), 1)
    xk, xk1 = index_and_shift(torch.from_numpy(x1).to(cuda), torch.from_numpy(k1).to(cuda), 1)
    yk, yk1, xk, xk1 = (t.cpu().numpy() for t in (yk, yk1, xk, xk1))
    v = yk1 - yk
    v = v * (x2[:, :128] - xk)
    v = v / (xk1 - xk)
    v = v + yk
    sel = im <= 70
    assert bits_equal(v[sel], want[:, :128][sel])


@pytest.mark.gpu
def test_gpu_transposes_match_reference_headers(cuda):
    import torch
    from ekat_b200 import tridiag as T
    for key, dt, shape, g in tr_cases():
        sv = g[key + "_src"]
        ni, nk = shape[0], shape[1]
        nj = shape[2] if len(shape) == 3 else 1
        c = torch.from_numpy(sv.reshape(ni, nk, nj)).to(cuda)
        got = T.transpose3(c, c2f=True).cpu().numpy().ravel()
        assert bits_equal(got, g[key + "_c2f"]), (key, "c2f")
        f = torch.from_numpy(sv.reshape(nj, nk, ni)).to(cuda)
        got = T.transpose3(f, c2f=False).cpu().numpy().ravel()
        assert bits_equal(got, g[key + "_f2c"]), (key, "f2c")


@pytest.mark.gpu
@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_gpu_reference_batch_through_transpose3_and_native_kernels(cuda, port, dt):
    """The marshalling row's purpose: a caller holding the reference batch A[col][3][lev][nA], X[col][lev][nrhs]
    (tests/algorithm/tridiag_tests_performance.cpp:181-185) reaches the native level-major kernels through
    ekat_b200_transpose3 and gets the oracle's bits back (bfb), padded pitches and strided sources included."""
    import torch
    import ekat_b200
    from ekat_b200 import tridiag as T
    tdt = torch.float64 if dt == "f64" else torch.float32
    for ncol, nlev, nrhs in [(1000, 72, 1), (3001, 128, 10), (77, 43, 3)]:
        A, B = port.fill_tridiag(DT[dt], ncol, nlev, 1, nrhs, col0=5)
        want = port.tridiag("bfb", A.copy(), B.copy())
        At, Xt = torch.from_numpy(A).to(cuda), torch.from_numpy(B).to(cuda)
        pitch = (ncol + 3) // 4 * 4
        dl, d, du = (T.transpose3(At[:, k], c2f=True, f_pitch=pitch) for k in range(3))  # strided [col][lev][1] slices
        X = T.transpose3(Xt, c2f=True, f_pitch=pitch)                                     # [nrhs][lev][pitch]
        T.bfb(dl, d, du, X, ncol=ncol)
        back = T.transpose3(X, c2f=False, ni=ncol)
        assert bits_equal(back.cpu().numpy(), want), (ncol, nlev, nrhs)


@pytest.mark.gpu
def test_gpu_repack_rows(cuda):
    import torch
    from ekat_b200 import tridiag as T
    for tdt in (torch.float64, torch.float32):
        x = torch.rand((1234, 72), dtype=tdt, device=cuda)
        p = T.repack(x, 80)                                  # Pack<.,16> storage of 72 levels: 5 packs
        assert torch.equal(p[:, :72], x) and bool((p[:, 72:] == 0).all())
        q = T.repack(p[:, :72], 72, zero_pad=False)          # and back (device_to_host drops the padding)
        assert torch.equal(q, x)
