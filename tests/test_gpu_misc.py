"""GPU tests for everything around the two kernels: device fills vs the reference fills, layout marshalling,
verification reductions, the fused solve+remap kernel, the host-buffer pipeline and the C++ adapters."""
import os
import subprocess

import numpy as np
import pytest
import torch

from helpers import a_to_lev, bits_equal, rel_diff, tol, u64sum, x_from_lev, x_to_lev
import ekat_b200
from ekat_b200 import synth, tridiag as T
from ekat_b200.host import Pipeline
from ekat_b200.lin_interp import tridiag_lin_interp

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DT = {"f64": (np.float64, torch.float64), "f32": (np.float32, torch.float32)}


@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_device_fills_equal_reference_fills(cuda, port, dt):
    npdt, tdt = DT[dt]
    for nrow, nA, nrhs in [(72, 1, 1), (43, 3, 3), (128, 1, 10)]:
        A, X = port.fill_tridiag(npdt, 300, nrow, nA, nrhs, col0=32700)  # crosses the seed wrap at 32768
        dl, d, du, Xd = T.fill(tdt, 300, nrow, nA, nrhs, layout=ekat_b200.LEV_MAJOR, col0=32700, device=cuda)
        for got, want in zip((dl, d, du), a_to_lev(A)):
            assert bits_equal(got.cpu().numpy(), want)
        assert bits_equal(Xd.cpu().numpy(), x_to_lev(X))
        dl, d, du, Xd = T.fill(tdt, 300, nrow, nA, nrhs, layout=ekat_b200.COL_MAJOR, col0=32700, device=cuda)
        assert bits_equal(torch.stack([dl, d, du], 1).cpu().numpy(), A) and bits_equal(Xd.cpu().numpy(), X)


@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_transposes_roundtrip(cuda, dt):
    _, tdt = DT[dt]
    for ncol, n in [(1, 1), (37, 72), (1000, 129), (33, 31)]:
        src = torch.randn((ncol, n + 5), dtype=tdt, device=cuda)[:, :n]  # strided rows
        lev = T.col_to_lev(src, pitch=ncol + 3)
        assert torch.equal(lev[:, :ncol], src.t())
        back = T.lev_to_col(lev, ncol=ncol)
        assert torch.equal(back, src)


@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_checksum_and_residual_equal_oracle(cuda, port, dt):
    npdt, tdt = DT[dt]
    for nA, nrhs in [(1, 1), (1, 4), (4, 4)]:
        A, B = port.fill_tridiag(npdt, 500, 72, nA, nrhs, col0=2)
        X = port.tridiag("bfb", A.copy(), B.copy())
        dl, d, du = (torch.from_numpy(a).to(cuda) for a in a_to_lev(A))
        Bd, Xd = torch.from_numpy(x_to_lev(B)).to(cuda), torch.from_numpy(x_to_lev(X)).to(cuda)
        assert (int(T.checksum(Xd).item()) & (2**64 - 1)) == u64sum(X)
        assert float(T.residual(dl, d, du, Bd, Xd).item()) == port.residual(A, B, X)
        Xbad = Xd.clone()
        Xbad[0, 3, 77] = float("nan")
        assert float(T.residual(dl, d, du, Bd, Xbad).item()) == float("inf")


@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_fused_solve_and_remap(cuda, port, dt):
    """BASELINE config 4 shape (reduced column count): 10-RHS shared-A solve + remap of the 10 solutions."""
    npdt, tdt = DT[dt]
    for ncol, nlev, km2, nrhs in [(200, 128, 128, 10), (37, 72, 128, 3), (50, 33, 17, 1), (4744, 128, 128, 10), (1000, 72, 72, 4),
                                  (5000, 33, 17, 5), (2500, 100, 60, 7), (148 * 32 * 2 + 8, 64, 128, 4),
                                  (64, 5, 9, 6), (96, 2, 3, 5), (1024, 256, 128, 5), (320, 16, 128, 12)]:
        A, B = port.fill_tridiag(npdt, ncol, nlev, 1, nrhs, col0=4)
        x1, _, x2 = synth.lin_interp_inputs(npdt, ncol, nlev, km2, 1, col0=4)
        Xref = port.tridiag("thomas", A.copy(), B.copy())                       # [ncol,nlev,nrhs]
        y1 = np.ascontiguousarray(Xref.transpose(2, 0, 1))                      # [nrhs,ncol,nlev]
        Yref, _ = port.lin_interp(x1, x2, y1, nlev, km2, 1)                     # [nrhs,ncol,km2]
        dl, d, du = (torch.from_numpy(a[0]).to(cuda) for a in a_to_lev(A))
        Bd = torch.from_numpy(x_to_lev(B)).to(cuda)
        x1d = torch.from_numpy(np.ascontiguousarray(x1.T)).to(cuda)
        x2d = torch.from_numpy(np.ascontiguousarray(x2.T)).to(cuda)
        Y2 = torch.zeros((nrhs, km2, ncol), dtype=tdt, device=cuda)
        Xd = torch.zeros_like(Bd)
        tridiag_lin_interp(dl, d, du, Bd, x1d, x2d, Y2, Xd)
        assert rel_diff(Xref, x_from_lev(Xd.cpu().numpy())) <= tol(npdt)
        assert rel_diff(Yref, Y2.cpu().numpy().transpose(0, 2, 1)) <= 10 * tol(npdt)
        # the remap half is bit-identical to LinInterp applied to the solutions the kernel itself produced
        Xg = np.ascontiguousarray(x_from_lev(Xd.cpu().numpy()).transpose(2, 0, 1))
        Yg, _ = port.lin_interp(x1, x2, Xg, nlev, km2, 1)
        assert bits_equal(Yg, np.ascontiguousarray(Y2.cpu().numpy().transpose(0, 2, 1)))
        # and it equals the unfused GPU path (thomas, then LinInterp on the transposed solution) to tolerance
        Y2b = torch.zeros_like(Y2)
        tridiag_lin_interp(dl, d, du, Bd, x1d, x2d, Y2b, None)
        assert rel_diff(Y2.cpu().numpy(), Y2b.cpu().numpy()) <= 10 * tol(npdt)  # (may take the one-kernel path: other thomas rounding)


@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_fused_repeatable_and_equal_to_two_pass(cuda, dt):
    """The one-kernel path hands shared-memory planes between warps through mbarriers, several tiles per CTA: any
    missed hand-over shows as run-to-run differences. Ten runs must be bit-identical, and agree with the two-pass path
    (other thomas rounding: tolerance) -- at a size where every CTA walks 6+ tiles."""
    from ekat_b200._lib import lib
    from ekat_b200.lin_interp import fill as li_fill
    tdt = DT[dt][1]
    ncol, nlev, km2, nrhs = 148 * 32 * 6 + 8, 128, 128, 10  # (pitch = ncol must keep rows 16-byte aligned for TMA)
    dl, d, du, B = T.fill(tdt, ncol, nlev, 1, nrhs, device=cuda)
    x1, _, x2 = li_fill(tdt, ncol, nlev, km2, 1, device=cuda, lev_major=True)
    Y, X = torch.zeros((nrhs, km2, ncol), dtype=tdt, device=cuda), torch.zeros_like(B)
    tridiag_lin_interp(dl[0], d[0], du[0], B, x1, x2, Y, X)
    assert lib().ekat_b200_last_kernel().decode().startswith("solve_remap_tma")
    for _ in range(9):
        Y2, X2 = torch.full_like(Y, float("nan")), torch.full_like(X, float("nan"))
        tridiag_lin_interp(dl[0], d[0], du[0], B, x1, x2, Y2, X2)
        assert torch.equal(Y2.view(torch.uint8), Y.view(torch.uint8)) and torch.equal(X2.view(torch.uint8), X.view(torch.uint8))
    lib().ekat_b200_set_tuning(0, 0, 3)  # two passes: persistent solve out of place, then level-major LinInterp
    try:
        Y3 = torch.zeros_like(Y)
        tridiag_lin_interp(dl[0], d[0], du[0], B, x1, x2, Y3, None)
        assert not lib().ekat_b200_last_kernel().decode().startswith("solve_remap_tma")
    finally:
        lib().ekat_b200_set_tuning(0, 0, 0)
    scale = float(Y3.abs().max())
    assert float((Y - Y3).abs().max()) <= 10 * tol(DT[dt][0]) * scale


@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_host_pipeline(cuda, port, dt):
    """Host buffers in, host buffers out, several chunks through 3 slots; same results as the device entry points."""
    npdt, tdt = DT[dt]
    pipe = Pipeline(device_bytes=48 << 20, nslots=3)
    try:
        ncol, nlev = 70001, 72
        A, B = port.fill_tridiag(npdt, ncol, nlev, 1, 1, col0=0)
        want = port.tridiag("bfb", A.copy(), B.copy())
        dl, d, du = (torch.from_numpy(a).pin_memory() for a in a_to_lev(A))
        X = torch.from_numpy(x_to_lev(B)).pin_memory()
        pipe.tridiag("bfb", dl, d, du, X, layout=ekat_b200.LEV_MAJOR)
        assert bits_equal(x_from_lev(X.numpy()), want)
        At, Xt = torch.from_numpy(A.copy()).pin_memory(), torch.from_numpy(B.copy()).pin_memory()
        pipe.tridiag("cr", At[:, 0], At[:, 1], At[:, 2], Xt, layout=ekat_b200.COL_MAJOR)
        assert bits_equal(Xt.numpy(), port.tridiag("cr", A.copy(), B.copy()))
        st = pipe.stats()
        assert st["chunks"] >= 4 and st["h2d_bytes"] == 2 * ncol * nlev * 4 * np.dtype(npdt).itemsize
        x1, y1, x2 = synth.lin_interp_inputs(npdt, 50001, 72, 128, 1)
        wy, _ = port.lin_interp(x1, x2, y1, 72, 128, 1)
        y2 = torch.zeros((50001, 128), dtype=tdt).pin_memory()
        pipe.lin_interp(torch.from_numpy(x1).pin_memory(), torch.from_numpy(x2).pin_memory(), torch.from_numpy(y1).pin_memory(), y2, 72, 128)
        assert bits_equal(y2.numpy(), wy)
    finally:
        pipe.close()


def test_cpp_host_adapters_on_device(cuda):
    """The ekat::tridiag / ekat::b200::LinInterp adapters driven from C++ on real device views."""
    src = os.path.join(ROOT, "tests", "cpp", "adapter_gpu.cu")
    exe = os.path.join(ROOT, "tests", "cpp", "adapter_gpu")
    cmd = ["nvcc", "-std=c++17", "-O1", "-gencode", "arch=compute_100a,code=sm_100a", "-ccbin", "g++",
           "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(ROOT, "oracle", "standin"), src, "-o", exe,
           "-L" + os.path.join(ROOT, "ekat_b200", "lib"), "-lekat_b200", "-Xlinker", "-rpath=" + os.path.join(ROOT, "ekat_b200", "lib")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "ALL ADAPTER CHECKS PASSED" in r.stdout
