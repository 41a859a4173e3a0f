"""Shared test helpers: layout conversions between the reference batch layout and the native one, norms."""
import numpy as np

# nrow list of the reference's property test (tests/algorithm/tridiag_tests_correctness.cpp:410-414)
NROWS = [1, 2, 3, 4, 5, 8, 10, 16, 32, 43, 63, 64, 65, 111, 128, 129]
NROWS_BIG = NROWS + [2048]


def eps(dtype):
    return np.finfo(dtype).eps


def tol(dtype):
    """north_star tolerances: 1e-13 relative in double, 1e-5 in float."""
    return 1e-13 if np.dtype(dtype) == np.float64 else 1e-5


def rel_diff(a, b):
    """max|a-b| / max|a| -- the reference's norm (tests/algorithm/tridiag_tests.hpp:89-106)."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    if not (np.isfinite(a).all() and np.isfinite(b).all()):
        return np.inf
    den = np.abs(a).max()
    return np.abs(a - b).max() / den if den > 0 else np.abs(a - b).max()


def bits_equal(a, b):
    a = np.ascontiguousarray(a)
    b = np.ascontiguousarray(b)
    if a.shape != b.shape or a.dtype != b.dtype:
        return False
    u = np.uint64 if a.dtype == np.float64 else np.uint32
    return bool(np.array_equal(a.view(u), b.view(u)))


def a_to_lev(A):
    """A[ncol,3,nrow,nA] -> (dl,d,du) each [nA,nrow,ncol] C-contiguous."""
    return tuple(np.ascontiguousarray(A[:, k].transpose(2, 1, 0)) for k in range(3))


def x_to_lev(X):
    """X[ncol,nrow,nrhs] -> [nrhs,nrow,ncol]."""
    return np.ascontiguousarray(X.transpose(2, 1, 0))


def x_from_lev(Xl):
    return np.ascontiguousarray(Xl.transpose(2, 1, 0))


def u64sum(a):
    a = np.ascontiguousarray(a)
    if a.dtype == np.float64:
        return int(a.view(np.uint64).sum(dtype=np.uint64))
    return int(a.view(np.uint32).astype(np.uint64).sum(dtype=np.uint64))
