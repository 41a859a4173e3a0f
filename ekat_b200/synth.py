"""Synthetic inputs of SURVEY.md section 8d as plain numpy -- the host statement of the generators the
library also implements on the device (csrc/util.cu), used by tests and bench to feed identical bytes to
the CUDA path and to the CPU checkers. Input synthesis only: nothing here solves or interpolates."""
import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def _splitmix64(z):
    with np.errstate(over="ignore"):
        z = (z + np.uint64(0x9E3779B97F4A7C15)) & _M64
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
        return z ^ (z >> np.uint64(31))


def _unit(key, stream, k):
    """u in [0,1) for (column key, stream, level) -- csrc/util.cu unit_rand."""
    with np.errstate(over="ignore"):
        v = _splitmix64(key[:, None] + (np.uint64(stream) << np.uint64(32)) + k[None, :].astype(np.uint64))
    return (v >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def lin_interp_inputs(dtype, ncol, km1, km2, pack_size=1, col0=0, seed=20260411):
    """x1 (strictly increasing), y1, x2 (sorted, strictly inside range(x1)); reference layout with zero padding.
    Mirrors tests/algorithm/lin_interp_test.cpp:182-216 and :386-390 with a counter-based RNG."""
    p1 = (km1 + pack_size - 1) // pack_size * pack_size
    p2 = (km2 + pack_size - 1) // pack_size * pack_size
    cols = np.arange(col0, col0 + ncol, dtype=np.uint64)
    key = _splitmix64(np.uint64(seed) ^ _splitmix64(cols))
    k1 = np.arange(km1)
    u = _unit(key, 0, k1)
    inc = 0.1 + 0.9 * u
    inc[:, 0] = u[:, 0]
    x1d = np.cumsum(inc, axis=1)  # sequential adds, same order as the device kernel
    y1d = 100.0 * _unit(key, 1, k1)
    x1 = np.zeros((ncol, p1), dtype)
    y1 = np.zeros((ncol, p1), dtype)
    x1[:, :km1] = x1d
    y1[:, :km1] = y1d
    x2 = np.zeros((ncol, p2), dtype)
    if km2 > 0:
        first, last = x1d[:, 0], x1d[:, km1 - 1]
        delta = last - first
        lo, hi = first + delta / 1000.0, last - delta / 1000.0
        w = 0.05 + _unit(key, 2, np.arange(km2 + 1))
        W = np.cumsum(w, axis=1)[:, km2]
        ck = np.cumsum(w[:, :km2], axis=1)
        x2[:, :km2] = lo[:, None] + (hi - lo)[:, None] * (ck / W[:, None])
    return x1, y1, x2
