"""ekat_b200_lin_interp_static_map_*: LinInterp with the index map in the library's file-scope device array
(ekat_b200/csrc/no_alloc.cu) instead of caller-owned scratch memory. No counterpart in the reference; added
for callers that cannot call cudaMalloc.

Called through the C ABI (ctypes), checked bit for bit against the existing entry point ekat_b200_lin_interp_* (and the
CPU oracle) on several calls in a row that SHARE the array. Collected last on purpose (file name), so that under
`pytest -x` a failure here hides nothing else; passes alone in a fresh process and after the other tests."""
import ctypes

import numpy as np
import pytest
import torch

from helpers import bits_equal
from ekat_b200 import synth
import ekat_b200._lib as L

pytestmark = pytest.mark.gpu
DT = {"f64": (np.float64, torch.float64), "f32": (np.float32, torch.float32)}
MAX_ENTRIES = 4194304  # EKAT_B200_STATIC_MAP_MAX_ENTRIES (include/ekat_b200/ekat_b200.h)


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _case(dev, npdt, ncol, km1, km2, N, col0, nf):
    """Device operands in the reference layout ([ncol][pitch], Pack<.,N> padding) + the oracle's inputs."""
    x1, y1, x2 = synth.lin_interp_inputs(npdt, ncol, km1, km2, N, col0=col0)
    x2[0, 0] = x1[0, 0] - 0.5            # extrapolate below
    x2[0, km2 - 1] = x1[0, km1 - 1] + 2  # and above
    y1f = np.stack([(f + 1) * y1 - f for f in range(nf)]).astype(npdt)
    tx1, tx2, ty1 = (torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (x1, x2, y1f))
    return (x1, x2, y1f), (tx1, tx2, ty1)


def _call(name, dt, tx1, tx2, ty1, ty2, ncol, km1, km2, nf, with_map_arg):
    fn = getattr(L.lib(), f"{name}_{dt}")
    p1, p2 = tx1.shape[-1], tx2.shape[-1]
    args = [_ptr(tx1), _ptr(tx2), _ptr(ty1), _ptr(ty2)]
    if with_map_arg:
        args.append(None)  # the existing fused entry point: map stays on chip
    args += [ncol, km1, km2, p1, p2, nf, ncol * p1, ncol * p2, None]
    return fn(*args)


@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_static_map_matches_existing_entry_point_on_calls_in_a_row(cuda, port, dt):
    npdt, tdt = DT[dt]
    # big, then SMALLER (stale entries of the first call stay behind it in the array), then the first shape again
    shapes = [(4099, 72, 128, 1, 3), (257, 33, 70, 8, 2), (4099, 72, 128, 1, 3), (1000, 128, 128, 16, 1)]
    for rep, (ncol, km1, km2, N, nf) in enumerate(shapes):
        (x1, x2, y1f), (tx1, tx2, ty1) = _case(cuda, npdt, ncol, km1, km2, N, col0=100 * rep + 3, nf=nf)
        p2 = tx2.shape[-1]
        want_t = torch.full((nf, ncol, p2), -777.0, dtype=tdt, device=cuda)
        got_t = torch.full((nf, ncol, p2), -777.0, dtype=tdt, device=cuda)
        assert _call("ekat_b200_lin_interp", dt, tx1, tx2, ty1, want_t, ncol, km1, km2, nf, True) == 0
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        launches0 = L.lib().ekat_b200_launch_count()
        assert _call("ekat_b200_lin_interp_static_map", dt, tx1, tx2, ty1, got_t, ncol, km1, km2, nf, False) == 0
        torch.cuda.synchronize()
        assert L.lib().ekat_b200_launch_count() == launches0 + 2  # setup + lin_interp kernels, nothing else
        if rep == 2:  # this shape's kernels were loaded by call 0: device memory in use must not move at all
            assert torch.cuda.mem_get_info()[0] == free0
        got, want = got_t.cpu().numpy(), want_t.cpu().numpy()
        assert bits_equal(got, want), (rep, ncol, km1, km2, N)  # bit-exact, padding lanes (-777) included
        oracle_y2, _ = port.lin_interp(x1, x2, y1f, km1, km2, N)
        assert bits_equal(got[..., :km2], oracle_y2[..., :km2]), (rep, ncol, km1, km2, N)


def test_static_map_limit_and_argument_checks(cuda):
    lib = L.lib()
    fn = lib.ekat_b200_lin_interp_static_map_f32
    km1, km2 = 4, 64
    ncol_ok, ncol_big = MAX_ENTRIES // km2, MAX_ENTRIES // km2 + 1
    x1 = torch.arange(km1, dtype=torch.float32, device=cuda).repeat(ncol_big, 1).contiguous()
    x2 = torch.linspace(0, km1 - 1, km2, dtype=torch.float32, device=cuda).repeat(ncol_big, 1).contiguous()
    y1 = (2 * x1 + 1).contiguous()
    y2 = torch.zeros(ncol_big, km2, dtype=torch.float32, device=cuda)
    args = lambda n: (_ptr(x1), _ptr(x2), _ptr(y1), _ptr(y2), n, km1, km2, km1, km2, 1, 0, 0, None)
    assert fn(*args(ncol_big)) == -4  # EKAT_B200_ERR_TOO_LARGE
    assert b"EKAT_B200_STATIC_MAP_MAX_ENTRIES" in lib.ekat_b200_last_cuda_error()
    assert fn(*args(ncol_ok)) == 0    # exactly the maximum is served
    torch.cuda.synchronize()
    assert torch.allclose(y2[:ncol_ok], 2 * x2[:ncol_ok] + 1, rtol=1e-5, atol=1e-5)  # y = 2x+1 is reproduced
    assert (y2[ncol_ok:] == 0).all()
    assert fn(*args(0)) == 0 and fn(*args(-1)) == -1
    assert fn(_ptr(x1), _ptr(x2), None, _ptr(y2), 8, km1, km2, km1, km2, 1, 0, 0, None) == -1
    assert fn(_ptr(x1), _ptr(x2), _ptr(y1), _ptr(y2), 8, 1, km2, km1, km2, 1, 0, 0, None) == -1  # km1 < 2
