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
        ncol, pitch1 = x1.shape
        pitch2 = x2.shape[1]
        y1f = y1 if y1.dim() == 3 else y1.unsqueeze(0)
        y2f = y2 if y2.dim() == 3 else y2.unsqueeze(0)
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
