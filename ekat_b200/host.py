"""Host-buffer entry points (pinned numpy / torch CPU tensors in, results out) through the C ABI's
chunked upload -> solve -> download pipeline (csrc/host_pipeline.cu). This is the call a user of the
reference's CPU build makes; bench.py times it for the end-to-end number."""
import ctypes

import torch

from . import LEV_MAJOR
from ._lib import lib, check

_SFX = {torch.float64: "f64", torch.float32: "f32"}
SOLVERS = {"thomas": 0, "bfb": 1, "cr": 2}


class Pipeline:
    """Owns device staging + streams (ekat_b200_pipeline). device_bytes is split over nslots."""

    def __init__(self, device_bytes=3 << 30, nslots=3):
        self._h = ctypes.c_void_p()
        check("ekat_b200_pipeline_create", lib().ekat_b200_pipeline_create(ctypes.byref(self._h), device_bytes, nslots))

    def close(self):
        if self._h:
            lib().ekat_b200_pipeline_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_deferred(self, on=True):
        """Deferred mode: tridiag()/lin_interp() return once their chunks are enqueued; sync() waits for all of them."""
        check("ekat_b200_pipeline_set_deferred", lib().ekat_b200_pipeline_set_deferred(self._h, 1 if on else 0))

    def sync(self):
        check("ekat_b200_pipeline_sync", lib().ekat_b200_pipeline_sync(self._h))

    def stats(self):
        a, b, c = ctypes.c_uint64(), ctypes.c_uint64(), ctypes.c_uint64()
        check("ekat_b200_pipeline_stats", lib().ekat_b200_pipeline_stats(self._h, ctypes.byref(a), ctypes.byref(b), ctypes.byref(c)))
        return {"h2d_bytes": a.value, "d2h_bytes": b.value, "chunks": c.value}

    def tridiag(self, solver, dl, d, du, X, layout=LEV_MAJOR, ncol=None):
        """Host tensors, same shapes as ekat_b200.tridiag; X (host) is overwritten with the solution."""
        for t in (dl, d, du, X):
            if t.is_cuda:
                raise ValueError("Pipeline.tridiag takes host tensors")
        ncol, nlev, nA, nrhs, a_stride, x_stride = _geometry_host(dl, d, du, X, layout, ncol)
        fn = f"ekat_b200_tridiag_host_{_SFX[X.dtype]}"
        check(fn, getattr(lib(), fn)(self._h, SOLVERS[solver], dl.data_ptr(), d.data_ptr(), du.data_ptr(), X.data_ptr(),
                                     ncol, nlev, nA, nrhs, a_stride, x_stride, layout))
        return X

    def lin_interp(self, x1, x2, y1, y2, km1, km2):
        """Host tensors x1,[y1]: [ncol][pitch1] ([nfields][ncol][pitch1]); x2,[y2]: [ncol][pitch2]; y2 is overwritten."""
        if x1.dim() != 2 or x2.dim() != 2 or x1.shape[0] != x2.shape[0]:
            raise ValueError("x1 [ncol][pitch1] and x2 [ncol][pitch2] expected")
        ncol, pitch1 = x1.shape
        pitch2 = x2.shape[1]
        y1f = y1 if y1.dim() == 3 else y1.unsqueeze(0)
        y2f = y2 if y2.dim() == 3 else y2.unsqueeze(0)
        if y1f.shape[0] != y2f.shape[0] or tuple(y1f.shape[1:]) != (ncol, pitch1) or tuple(y2f.shape[1:]) != (ncol, pitch2):
            raise ValueError("y1 must be [nfields][ncol][pitch1] and y2 [nfields][ncol][pitch2]")
        if not (2 <= km1 <= pitch1 and 0 <= km2 <= pitch2):
            raise ValueError("km1/km2 outside the row pitch")
        for t in (x1, x2, y1f, y2f):
            if t.is_cuda:
                raise ValueError("Pipeline.lin_interp takes host tensors")
            if t.dtype != x1.dtype or t.dtype not in _SFX:
                raise ValueError("x1, x2, y1, y2 must share one dtype (float32 or float64)")
            if not t.is_contiguous():
                raise ValueError("Pipeline.lin_interp takes contiguous tensors")
        fn = f"ekat_b200_lin_interp_host_{_SFX[x1.dtype]}"
        check(fn, getattr(lib(), fn)(self._h, x1.data_ptr(), x2.data_ptr(), y1f.data_ptr(), y2f.data_ptr(), ncol, km1, km2,
                                     pitch1, pitch2, y1f.shape[0], ncol * pitch1, ncol * pitch2))
        return y2


def _geometry_host(dl, d, du, X, layout, ncol):
    from . import tridiag as _t

    class _View:
        """Presents a host tensor to tridiag._geometry as if it were a CUDA tensor (shape/stride checks only)."""

        def __init__(self, t):
            self._t = t
            self.is_cuda = True

        def __getattr__(self, n):
            return getattr(self._t, n)

        def dim(self):
            return self._t.dim()

        def unsqueeze(self, i):
            return _View(self._t.unsqueeze(i))

        def stride(self, *a):
            return self._t.stride(*a)

    return _t._geometry(_View(dl), _View(d), _View(du), _View(X), layout, ncol)
