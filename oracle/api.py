"""ctypes face of the two CPU checkers. TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Array conventions (the reference's batch layout,
tests/algorithm/tridiag_tests_performance.cpp:181-185):
  A : [ncol, 3, nrow, nA]   (dl, d, du)      X : [ncol, nrow, nrhs]
LinInterp (src/algorithm/ekat_lin_interp_impl.hpp:12-19):
  x1, y1 : [ncol, pitch1]   x2, y2 : [ncol, pitch2],  pitch = npack(km, N) * N
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SOLVERS = {"thomas": 0, "thomas_team": 1, "cr": 2, "bfb": 3}
_i64, _int = ctypes.c_int64, ctypes.c_int


def build(force=False, quiet=True):
    """Run oracle/Makefile (port always; _ref only where /root/reference exists)."""
    args = ["make", "-C", _HERE] + (["-B"] if force else [])
    r = subprocess.run(args, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + r.stdout + r.stderr)
    if not quiet:
        print(r.stdout)


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def _sfx(dtype):
    dtype = np.dtype(dtype)
    if dtype == np.float64:
        return "f64"
    if dtype == np.float32:
        return "f32"
    raise TypeError(f"unsupported dtype {dtype}")


def npack(n, packn):
    return (n + packn - 1) // packn


class CpuOracle:
    """Common face over libekat_oracle.so (kind='port') and _ref/libekat_ref.so (kind='reference')."""

    def __init__(self, path, kind):
        self.kind = kind
        self.path = path
        self.lib = ctypes.CDLL(path)
        self._pre = "ekat_oracle_" if kind == "port" else "ekat_ref_"
        for s in ("f64", "f32"):
            getattr(self.lib, self._pre + "residual_" + s).restype = ctypes.c_double
        if kind == "reference":
            self.lib.ekat_ref_build_info.restype = ctypes.c_char_p

    # ---- threads -------------------------------------------------------
    def max_threads(self):
        return int(getattr(self.lib, self._pre + "max_threads")())

    def set_threads(self, n):
        getattr(self.lib, self._pre + "set_threads")(_int(int(n)))

    def info(self):
        if self.kind == "reference":
            return self.lib.ekat_ref_build_info().decode()
        return "plain-C restatement (oracle/ekat_oracle.c), gcc -O2 -ffp-contract=off -fopenmp"

    # ---- tridiag -------------------------------------------------------
    def fill_tridiag(self, dtype, ncol, nrow, nA, nrhs, col0=0, seed_mod=32768):
        """Reference fills (tests/algorithm/tridiag_tests.hpp:11-40), seed = (col0+c) % seed_mod."""
        A = np.zeros((ncol, 3, nrow, nA), dtype)
        X = np.zeros((ncol, nrow, nrhs), dtype)
        getattr(self.lib, self._pre + "fill_tridiag_" + _sfx(dtype))(
            _p(A), _p(X), _i64(ncol), _i64(col0), _int(nrow), _int(nA), _int(nrhs), _int(seed_mod))
        return A, X

    def tridiag(self, solver, A, X):
        """In place: X <- A^-1 X, A clobbered as the reference clobbers it. Returns X."""
        assert A.flags.c_contiguous and X.flags.c_contiguous and A.dtype == X.dtype
        ncol, three, nrow, nA = A.shape
        assert three == 3 and X.shape[:2] == (ncol, nrow)
        nrhs = X.shape[2]
        rc = getattr(self.lib, self._pre + "tridiag_" + _sfx(A.dtype))(
            _int(SOLVERS[solver]), _p(A), _p(X), _i64(ncol), _int(nrow), _int(nA), _int(nrhs))
        if rc != 0:
            raise ValueError(f"{self.kind} oracle tridiag rc={rc}")
        return X

    def residual(self, A, B, X):
        """Worst column of max|B - A X| / max|B| (tests/algorithm/tridiag_tests.hpp:42-106)."""
        ncol, _, nrow, nA = A.shape
        nrhs = X.shape[2]
        return float(getattr(self.lib, self._pre + "residual_" + _sfx(A.dtype))(
            _p(np.ascontiguousarray(A)), _p(np.ascontiguousarray(B)), _p(np.ascontiguousarray(X)),
            _i64(ncol), _int(nrow), _int(nA), _int(nrhs)))

    # ---- LinInterp -----------------------------------------------------
    def lin_interp(self, x1, x2, y1, km1, km2, packn=1, want_indx=True):
        """setup + lin_interp for every column. y1 may be [ncol,pitch1] or [nfields,ncol,pitch1].

        Returns (y2, indx_map) with y2 shaped like y1 but pitch2 wide; padding lanes are zero."""
        x1 = np.ascontiguousarray(x1)
        x2 = np.ascontiguousarray(x2)
        y1 = np.ascontiguousarray(y1)
        ncol = x1.shape[0]
        pitch1, pitch2 = npack(km1, packn) * packn, npack(km2, packn) * packn
        assert x1.shape == (ncol, pitch1) and x2.shape == (ncol, pitch2), (x1.shape, x2.shape, pitch1, pitch2)
        single = y1.ndim == 2
        y1f = y1[None] if single else y1
        nfields = y1f.shape[0]
        assert y1f.shape == (nfields, ncol, pitch1)
        y2 = np.zeros((nfields, ncol, pitch2), x1.dtype)
        indx = np.zeros((ncol, km2), np.int32) if want_indx else None
        s = _sfx(x1.dtype)
        if self.kind == "port":
            rc = getattr(self.lib, "ekat_oracle_lin_interp_" + s)(
                _p(x1), _p(x2), _p(y1f), _p(y2), _p(indx), _i64(ncol), _int(km1), _int(km2),
                _i64(pitch1), _i64(pitch2), _int(nfields))
        else:
            rc = getattr(self.lib, "ekat_ref_lin_interp_" + s)(
                _int(packn), _p(x1), _p(x2), _p(y1f), _p(y2), _p(indx), _i64(ncol), _int(km1), _int(km2),
                _int(nfields))
            if pitch2 > km2:  # the reference computes garbage in padding lanes (impl.hpp:109-157); drop it
                y2[..., km2:] = 0
        if rc != 0:
            raise ValueError(f"{self.kind} oracle lin_interp rc={rc}")
        return (y2[0] if single else y2), indx


    def lin_interp_raw(self, x1, x2, y1f, y2, indx, km1, km2, packn=1):
        """The bare C call on caller-allocated, C-contiguous arrays (y1f/y2: [nfields,ncol,pitch]) -- for timing."""
        ncol, nfields = x1.shape[0], y1f.shape[0]
        s = _sfx(x1.dtype)
        if self.kind == "port":
            return getattr(self.lib, "ekat_oracle_lin_interp_" + s)(
                _p(x1), _p(x2), _p(y1f), _p(y2), _p(indx), _i64(ncol), _int(km1), _int(km2),
                _i64(x1.shape[1]), _i64(x2.shape[1]), _int(nfields))
        return getattr(self.lib, "ekat_ref_lin_interp_" + s)(
            _int(packn), _p(x1), _p(x2), _p(y1f), _p(y2), _p(indx), _i64(ncol), _int(km1), _int(km2), _int(nfields))


    # ---- ReductionUtils ------------------------------------------------
    def view_reduction(self, x, begin, end, packn=0):
        """Per-column sum over the scalar range [begin, end) of x [ncol, pitch].
        packn = 0: Serialize = true (in order); packn = N >= 1: the reference's packed order for Pack<Real,N>."""
        x = np.ascontiguousarray(x)
        ncol, pitch = x.shape
        out = np.zeros(ncol, x.dtype)
        s = _sfx(x.dtype)
        if self.kind == "port":
            rc = getattr(self.lib, "ekat_oracle_view_reduction_" + s)(
                _p(x), _p(out), _i64(ncol), _int(begin), _int(end), _i64(pitch), _int(packn))
        else:
            rc = getattr(self.lib, "ekat_ref_view_reduction_" + s)(
                _int(max(packn, 1)), _int(1 if packn == 0 else 0), _p(x), _p(out), _i64(ncol), _int(begin), _int(end),
                _i64(pitch))
        if rc != 0:
            raise ValueError(f"{self.kind} oracle view_reduction rc={rc}")
        return out


    def adj_diff(self, x, forward=True, packn=0):
        """ekat::adj_diff at every index of every column of x [ncol, pitch] (packn: 0 = scalar view, N = Pack<Real,N>)."""
        x = np.ascontiguousarray(x)
        ncol, pitch = x.shape
        out = np.zeros_like(x)
        s = _sfx(x.dtype)
        if self.kind == "port":
            rc = getattr(self.lib, "ekat_oracle_adj_diff_" + s)(_p(x), _p(out), _i64(ncol), _int(pitch), _i64(pitch),
                                                                _int(1 if forward else 0))
        else:
            rc = getattr(self.lib, "ekat_ref_adj_diff_" + s)(_int(packn), _int(1 if forward else 0), _p(x), _p(out),
                                                             _i64(ncol), _i64(pitch))
        if rc != 0:
            raise ValueError(f"{self.kind} oracle adj_diff rc={rc}")
        return out


    def shift(self, x, shift_by, fill=0.0, packn=1):
        """ekat::shift_right (shift_by=+1) / shift_left (-1) pack after pack along every column of x [ncol, pitch]."""
        x = np.ascontiguousarray(x)
        ncol, pitch = x.shape
        out = np.zeros_like(x)
        s = _sfx(x.dtype)
        cf = ctypes.c_double if x.dtype == np.float64 else ctypes.c_float
        if self.kind == "port":
            rc = getattr(self.lib, "ekat_oracle_shift_" + s)(_p(x), _p(out), _i64(ncol), _int(pitch), _i64(pitch),
                                                             _int(shift_by), cf(fill))
        else:
            if shift_by not in (1, -1):
                raise ValueError("the reference shifts by one place")
            rc = getattr(self.lib, "ekat_ref_shift_" + s)(_int(max(packn, 1)), _int(shift_by), cf(fill), _p(x), _p(out),
                                                          _i64(ncol), _i64(pitch))
        if rc != 0:
            raise ValueError(f"{self.kind} oracle shift rc={rc}")
        return out

    def index_and_shift(self, a, idx, shift_by, out0=None, packn=1):
        """ekat::index_and_shift<shift_by> for every index of every column: a [ncol, n], idx [ncol, m] int32.
        Returns (index, index_shift); lanes with an out-of-range (shifted) index keep out0 / get NaN."""
        a = np.ascontiguousarray(a)
        idx = np.ascontiguousarray(idx, dtype=np.int32)
        ncol, n = a.shape
        m = idx.shape[1]
        out = np.full((ncol, m), -777.0, a.dtype) if out0 is None else np.ascontiguousarray(out0).copy()
        outs = np.zeros((ncol, m), a.dtype)
        s = _sfx(a.dtype)
        if self.kind == "port":
            rc = getattr(self.lib, "ekat_oracle_index_and_shift_" + s)(_p(a), _p(idx), _p(out), _p(outs), _i64(ncol), _int(n),
                                                                       _int(m), _i64(n), _int(shift_by))
        else:
            rc = getattr(self.lib, "ekat_ref_index_and_shift_" + s)(_int(packn), _int(shift_by), _p(a), _p(idx), _p(out),
                                                                    _p(outs), _i64(ncol), _int(n), _int(m))
        if rc != 0:
            raise ValueError(f"{self.kind} oracle index_and_shift rc={rc}")
        return out, outs

    def transpose(self, sv, shape, c2f):
        """ekat::transpose<c2f|f2c> of a flat array: shape = (ni, nk) or (ni, nk, nj). Returns the flat result."""
        sv = np.ascontiguousarray(sv).ravel()
        dv = np.zeros_like(sv)
        ni, nk = shape[0], shape[1]
        nj = shape[2] if len(shape) == 3 else 0
        s = _sfx(sv.dtype)
        if self.kind == "port":
            rc = getattr(self.lib, "ekat_oracle_transpose_" + s)(_p(sv), _p(dv), _int(ni), _int(nk), _int(max(nj, 1)), _int(1 if c2f else 0))
        else:
            rc = getattr(self.lib, "ekat_ref_transpose_" + s)(_p(sv), _p(dv), _int(ni), _int(nk), _int(nj), _int(1 if c2f else 0))
        if rc != 0:
            raise ValueError(f"{self.kind} oracle transpose rc={rc}")
        return dv


_cache = {}


def port():
    """The plain-C restatement; built on demand (gcc is on every box)."""
    if "port" not in _cache:
        path = os.path.join(_HERE, "libekat_oracle.so")
        src_m = max(os.path.getmtime(os.path.join(_HERE, f)) for f in ("ekat_oracle.c", "ekat_oracle_impl.inc"))
        if not os.path.exists(path) or os.path.getmtime(path) < src_m:
            build()
        _cache["port"] = CpuOracle(path, "port")
    return _cache["port"]


def ref():
    """The compiled-in-place reference (oracle/_ref), or None where it was never built."""
    if "ref" not in _cache:
        path = os.path.join(_HERE, "_ref", "libekat_ref.so")
        _cache["ref"] = CpuOracle(path, "reference") if os.path.exists(path) else None
    return _cache["ref"]
