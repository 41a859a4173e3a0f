"""BASELINE configs 4 and 5 at their stated sizes (configs 1-3 at full size: test_gpu_tridiag.py
test_full_size_configs_properties, test_gpu_lin_interp.py test_config2_full_size). The oracle cannot run 10^6 columns in
seconds, so a full-size run is checked through size-independent properties on EVERY column (the reference's residual
norm, tests/algorithm/tridiag_tests.hpp:42-106, as a device reduction; partition-invariance of the checksum) and against
the oracle on slabs of columns cut from the device operands themselves: first columns, the 32768 seed wrap of the
reference fill, the middle (a tile boundary of the persistent kernels) and the last columns."""
import numpy as np
import pytest
import torch

from helpers import bits_equal, eps, rel_diff, tol, x_from_lev
from ekat_b200 import tridiag as T
from ekat_b200._lib import lib
from ekat_b200.lin_interp import LinInterp, fill as li_fill, tridiag_lin_interp

pytestmark = pytest.mark.gpu
NLEV = 128
W = 48  # columns per slab


def slabs(ncol):
    return [0, 32768 - W // 2, ncol // 2 - W // 2 + 5, ncol - W]


def host_matrix(dl, d, du, c0):
    """Level-major device diagonals [nlev][ncol] -> the oracle's A[W][3][nlev][1]."""
    return np.ascontiguousarray(np.stack([t[:, c0:c0 + W].cpu().numpy().T for t in (dl, d, du)], axis=1)[..., None])


def test_config5_shard_full_size(cuda, port):
    """3,145,728 columns x 128 levels, double: thomas and bfb (1 rhs) and LinInterp 128 -> 128 as bench.py runs them."""
    ncol, dt, npdt = 3145728, torch.float64, np.float64
    dl, d, du, B = T.fill(dt, ncol, NLEV, 1, 1, device=cuda)
    for solver in ("thomas", "bfb"):
        X = B.clone()
        getattr(T, solver)(dl, d, du, X)
        assert float(T.residual(dl, d, du, B, X).item()) <= 50 * eps(npdt), solver
        for c0 in slabs(ncol):
            A_h = host_matrix(dl[0], d[0], du[0], c0)
            B_h = x_from_lev(B[:, :, c0:c0 + W].cpu().numpy())
            want = port.tridiag(solver, A_h, B_h.copy())
            got = x_from_lev(X[:, :, c0:c0 + W].cpu().numpy())
            if solver == "bfb":
                assert bits_equal(got, want), (solver, c0)
            else:
                assert rel_diff(want, got) <= tol(npdt), (solver, c0)
        whole = int(T.checksum(X).item())
        parts = torch.zeros(1, dtype=torch.int64, device=cuda)
        for lo, hi in [(0, ncol // 3), (ncol // 3, ncol // 2 + 7), (ncol // 2 + 7, ncol)]:
            T.checksum(X[:, :, lo:hi], out=parts)
        assert whole == int(parts.item()), solver
    del dl, d, du, B, X
    torch.cuda.empty_cache()

    x1, y1, x2 = li_fill(dt, ncol, NLEV, NLEV, 1, device=cuda)
    y2 = torch.full((ncol, NLEV), -777.0, dtype=dt, device=cuda)
    LinInterp(ncol, NLEV, NLEV, dt, 1, device=cuda).run(x1, x2, y1, y2)
    for c0 in slabs(ncol):
        s = slice(c0, c0 + W)
        want, _ = port.lin_interp(x1[s].cpu().numpy(), x2[s].cpu().numpy(), y1[s].cpu().numpy(), NLEV, NLEV, 1)
        assert bits_equal(y2[s].cpu().numpy(), want), c0
    # every target lies strictly inside its column's source range (synth.lin_interp_inputs), so on EVERY column y2 is
    # bracketed by the column's y1 range (to rounding: y1 is in [0, 100))
    lo, hi = y1.min(dim=1).values - 1e-9, y1.max(dim=1).values + 1e-9
    assert bool(((y2 >= lo[:, None]) & (y2 <= hi[:, None])).all())


def test_config4_full_size(cuda, port):
    """1,572,864 columns x 128 levels x 10 right-hand sides sharing one matrix, double: the one-kernel solve + remap, the
    10-rhs thomas on its own and the 10-rhs bfb."""
    ncol, nrhs, dt, npdt = 1572864, 10, torch.float64, np.float64
    dl, d, du, B = T.fill(dt, ncol, NLEV, 1, nrhs, device=cuda)
    x1, _, x2 = li_fill(dt, ncol, NLEV, NLEV, 1, device=cuda, lev_major=True)
    Y2 = torch.zeros((nrhs, NLEV, ncol), dtype=dt, device=cuda)
    X = torch.zeros_like(B)
    tridiag_lin_interp(dl[0], d[0], du[0], B, x1, x2, Y2, X)
    assert lib().ekat_b200_last_kernel().decode().startswith("solve_remap_tma")
    assert float(T.residual(dl, d, du, B, X).item()) <= 50 * eps(npdt)
    Xb = B.clone()
    T.bfb(dl, d, du, Xb)
    assert lib().ekat_b200_last_kernel().decode().startswith("tridiag_bfb_shared")
    assert float(T.residual(dl, d, du, B, Xb).item()) <= 50 * eps(npdt)
    Xt = B.clone()
    T.thomas(dl, d, du, Xt)
    assert float(T.residual(dl, d, du, B, Xt).item()) <= 50 * eps(npdt)
    for c0 in slabs(ncol):
        s = slice(c0, c0 + W)
        A_h = host_matrix(dl[0], d[0], du[0], c0)
        B_h = x_from_lev(B[:, :, s].cpu().numpy())                               # [W][nlev][nrhs]
        Xref = port.tridiag("thomas", A_h.copy(), B_h.copy())
        x1h, x2h = (np.ascontiguousarray(t[:, s].cpu().numpy().T) for t in (x1, x2))
        Yref, _ = port.lin_interp(x1h, x2h, np.ascontiguousarray(Xref.transpose(2, 0, 1)), NLEV, NLEV, 1)
        Xg = x_from_lev(X[:, :, s].cpu().numpy())
        Yg = np.ascontiguousarray(Y2[:, :, s].cpu().numpy().transpose(0, 2, 1))  # [nrhs][W][km2]
        assert rel_diff(Xref, Xg) <= tol(npdt), c0
        assert rel_diff(Yref, Yg) <= 10 * tol(npdt), c0
        # the remap half is bit-identical to LinInterp applied to the solutions the kernel itself produced
        Yk, _ = port.lin_interp(x1h, x2h, np.ascontiguousarray(Xg.transpose(2, 0, 1)), NLEV, NLEV, 1)
        assert bits_equal(Yk, Yg), c0
        assert rel_diff(Xref, x_from_lev(Xt[:, :, s].cpu().numpy())) <= tol(npdt), c0
        assert bits_equal(x_from_lev(Xb[:, :, s].cpu().numpy()), port.tridiag("bfb", A_h.copy(), B_h.copy())), c0
