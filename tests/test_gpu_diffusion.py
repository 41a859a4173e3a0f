"""The fused "difference -> (dl, d, du) -> solve" step (include/ekat_b200/ekat_b200.h
"diffusion_solve"): the matrix the kernel implies is checked bit for bit against numpy (every operation is one
separately rounded IEEE operation), and the fused solve against the oracle's thomas on that matrix, at the tolerance
the reference's own tridiag tests use (tests/algorithm/tridiag_tests.hpp:42-106: residual <= 50 eps)."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def stencil(z, kd):
    """numpy restatement, level-major [nlev, ncol]"""
    nlev = z.shape[0]
    w = np.zeros_like(z)
    w[:-1] = kd[:-1] / (z[1:] - z[:-1])
    wm = np.zeros_like(z)
    wm[1:] = w[:-1]
    one = z.dtype.type(1)
    return -wm, (one + w) + wm, -w


def operands(dt, ncol, nlev, pitch, seed):
    rng = np.random.default_rng(seed)
    z = np.zeros((nlev, pitch), dt)
    z[:, :ncol] = np.cumsum(rng.uniform(20.0, 400.0, (nlev, ncol)), axis=0)
    z[:, ncol:] = np.arange(nlev)[:, None]  # padding columns stay finite
    kd = rng.uniform(0.0, 3000.0, (nlev, pitch)).astype(dt)
    b = rng.uniform(-1.0, 1.0, (nlev, pitch)).astype(dt)
    return z.astype(dt), kd, b


@pytest.mark.parametrize("dt", [np.float64, np.float32])
@pytest.mark.parametrize("ncol,nlev,pitch", [(1, 2, 4), (33, 5, 36), (257, 72, 260), (2048, 128, 2048), (100, 250, 104), (64, 17, 64)])
def test_fused_difference_fill_solve(port, dt, ncol, nlev, pitch):
    from ekat_b200 import tridiag as T
    from helpers import bits_equal

    if dt == np.float32 and pitch % 4:
        pitch += 4 - pitch % 4
    z, kd, b = operands(dt, ncol, nlev, pitch, 11 * nlev + ncol)
    dev = torch.device("cuda:0")
    zt, kt, xt = (torch.from_numpy(a).to(dev) for a in (z, kd, b))
    dl, d, du = T.diffusion_matrix(zt, kt, ncol)
    edl, ed, edu = stencil(z, kd)
    for got, exp in ((dl, edl), (d, ed), (du, edu)):
        assert bits_equal(got.cpu().numpy()[:, :ncol], exp[:, :ncol])
    T.diffusion_solve(zt, kt, xt, ncol)
    x = xt.cpu().numpy()
    assert bits_equal(x[:, ncol:], b[:, ncol:])  # padding columns are not written
    # the oracle on the same matrix (reference layout A[col][3][lev][1], X[col][lev][1])
    A = np.ascontiguousarray(np.stack([edl[:, :ncol].T, ed[:, :ncol].T, edu[:, :ncol].T], axis=1)[..., None])
    B = np.ascontiguousarray(b[:, :ncol].T[..., None])
    Xo = port.tridiag("thomas", A.copy(), B.copy())
    Xg = np.ascontiguousarray(x[:, :ncol].T[..., None])
    eps = np.finfo(dt).eps
    # the reference's 50 eps bar is for its own fills (|A| ~ 10); these matrices reach |A| ~ 300, so the bar is the
    # oracle's own residual on the same matrix, with a factor for the different (equally valid) rounding sequence
    assert port.residual(A, B, Xg) <= max(50 * eps, 4 * port.residual(A, B, Xo))
    scale = np.abs(Xo).max(axis=1, keepdims=True)
    assert (np.abs(Xg - Xo) / scale).max() <= 200 * eps  # diagonally dominant: both are within a few ulps of the solution


def test_fused_matches_two_pass_path():
    """the fused kernel against the library's own two-pass form (diffusion_matrix + thomas) at a full-size tile count"""
    from ekat_b200 import tridiag as T

    dev = torch.device("cuda:0")
    ncol, nlev = 21600, 72
    z, kd, b = operands(np.float64, ncol, nlev, ncol, 5)
    zt, kt, xt = (torch.from_numpy(a).to(dev) for a in (z, kd, b))
    dl, d, du = T.diffusion_matrix(zt, kt)
    x2 = xt.clone()
    T.thomas(dl[None], d[None], du[None], x2[None])
    T.diffusion_solve(zt, kt, xt)
    err = ((xt - x2).abs().amax(0) / x2.abs().amax(0)).max().item()
    assert err <= 200 * np.finfo(np.float64).eps


def test_unsupported_shapes_fail_loudly():
    from ekat_b200 import tridiag as T
    from ekat_b200._lib import EkatB200Error

    dev = torch.device("cuda:0")
    z = torch.arange(400, dtype=torch.float64, device=dev)[:, None].repeat(1, 32).contiguous()
    with pytest.raises(EkatB200Error):
        T.diffusion_solve(z, z.clone(), z.clone())  # 3 planes x 400 levels x 32 columns do not fit on chip
