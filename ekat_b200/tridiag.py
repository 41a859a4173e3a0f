"""Batched tridiagonal solvers on torch CUDA tensors -- the Python face of
ekat::tridiag::{thomas, bfb, cr} (reference: src/algorithm/ekat_tridiag.hpp:51-88).

Where the reference is called once per column inside the caller's team kernel, these take the
whole batch. X is overwritten with the solution, as in the reference;
dl/d/du are left untouched unless ``clobber=True``.

Shapes
  LEV_MAJOR (native):  dl,d,du : [nlev, pitch] or [nA, nlev, pitch];  X : [nlev, pitch] or [nrhs, nlev, pitch]
  COL_MAJOR (reference, tests/algorithm/tridiag_tests_performance.cpp:181-185):
                       dl,d,du : [ncol, nlev] or [ncol, nlev, nA];    X : [ncol, nlev] or [ncol, nlev, nrhs]
                       (dim-0 may be strided, e.g. slices A[:,0], A[:,1], A[:,2] of A[ncol,3,nlev,nA])
"""
import torch

from . import COL_MAJOR, LEV_MAJOR, CLOBBER_DIAGS
from ._lib import lib, check

_SFX = {torch.float64: "f64", torch.float32: "f32"}


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _require(cond, msg):
    if not cond:
        raise ValueError(msg)


def _geometry(dl, d, du, X, layout, ncol):
    for t in (dl, d, du, X):
        _require(t.is_cuda, "ekat_b200.tridiag needs CUDA tensors (there is no CPU path)")
        _require(t.dtype == X.dtype and t.dtype in _SFX, "dl,d,du,X must share dtype float32/float64")
    _require(dl.shape == d.shape == du.shape and dl.stride() == d.stride() == du.stride(),
             "dl,d,du must have identical shape and strides")
    if layout == LEV_MAJOR:
        A3 = dl if dl.dim() == 3 else dl.unsqueeze(0)
        X3 = X if X.dim() == 3 else X.unsqueeze(0)
        _require(A3.dim() == 3 and X3.dim() == 3, "LEV_MAJOR expects [nlev,pitch] or [n,nlev,pitch]")
        nA, nlev, pitch_a = A3.shape
        nrhs, nlev_x, pitch_x = X3.shape
        _require(nlev == nlev_x, "level counts differ")
        for t3, n in ((A3, nA), (X3, nrhs)):
            _require(t3.stride(2) == 1 and t3.stride(1) >= t3.shape[2] and (n == 1 or t3.stride(0) == nlev * t3.stride(1)),
                     "LEV_MAJOR planes must be [plane][lev][pitch] with unit column stride")
        a_stride, x_stride = A3.stride(1), X3.stride(1)
        ncol = min(pitch_a, pitch_x) if ncol is None else ncol
        _require(ncol <= pitch_a and ncol <= pitch_x, "ncol exceeds the column pitch")
    elif layout == COL_MAJOR:
        A3 = dl if dl.dim() == 3 else dl.unsqueeze(-1)
        X3 = X if X.dim() == 3 else X.unsqueeze(-1)
        nc_a, nlev, nA = A3.shape
        nc_x, nlev_x, nrhs = X3.shape
        _require(nlev == nlev_x and nc_a == nc_x, "column / level counts differ")
        for t3 in (A3, X3):
            _require(t3.stride(2) == 1 and t3.stride(1) == t3.shape[2],
                     "COL_MAJOR columns must be contiguous [lev][n] blocks (LayoutRight; LayoutStride is "
                     "rejected as in ekat_tridiag.hpp:33-35)")
        a_stride = A3.stride(0) if nc_a > 1 else nlev * nA
        x_stride = X3.stride(0) if nc_x > 1 else nlev * nrhs
        ncol = nc_a if ncol is None else ncol
        _require(ncol <= nc_a, "ncol exceeds the batch")
    else:
        raise ValueError(f"unknown layout {layout}")
    _require(nA in (1, nrhs), "nA must be 1 or nrhs (ekat_tridiag.hpp:18-31)")
    return ncol, nlev, nA, nrhs, a_stride, x_stride


def _solve(name, dl, d, du, X, layout, clobber, ncol):
    ncol, nlev, nA, nrhs, a_stride, x_stride = _geometry(dl, d, du, X, layout, ncol)
    fn = f"ekat_b200_tridiag_{name}_{_SFX[X.dtype]}"
    rc = getattr(lib(), fn)(dl.data_ptr(), d.data_ptr(), du.data_ptr(), X.data_ptr(), ncol, nlev, nA, nrhs,
                            a_stride, x_stride, layout, CLOBBER_DIAGS if clobber else 0, _stream())
    check(fn, rc)
    return X


def thomas(dl, d, du, X, layout=LEV_MAJOR, clobber=False, ncol=None):
    """ekat::tridiag::thomas for every column (ekat_tridiag.hpp:323-381). Tolerance-level (FMA) arithmetic."""
    return _solve("thomas", dl, d, du, X, layout, clobber, ncol)


def bfb(dl, d, du, X, layout=LEV_MAJOR, clobber=False, ncol=None):
    """ekat::tridiag::bfb for every column (ekat_tridiag.hpp:632-678). Bit-for-bit with the reference."""
    return _solve("bfb", dl, d, du, X, layout, clobber, ncol)


def cr(dl, d, du, X, layout=LEV_MAJOR, clobber=False, ncol=None):
    """ekat::tridiag::cr for every column (ekat_tridiag.hpp:385-630). Bit-for-bit with the reference's CR."""
    return _solve("cr", dl, d, du, X, layout, clobber, ncol)


def residual(dl, d, du, B, X, ncol=None):
    """max over columns of max|B - A X| / max|B| (tests/algorithm/tridiag_tests.hpp:42-106), LEV_MAJOR operands.
    Returns a 1-element float64 CUDA tensor (ready for one all-reduce MAX)."""
    ncol, nlev, nA, nrhs, a_stride, x_stride = _geometry(dl, d, du, X, LEV_MAJOR, ncol)
    if a_stride != x_stride or B.stride() != X.stride():
        raise ValueError("residual needs one common column pitch")
    out = torch.zeros(1, dtype=torch.float64, device=X.device)
    fn = f"ekat_b200_tridiag_residual_{_SFX[X.dtype]}"
    check(fn, getattr(lib(), fn)(dl.data_ptr(), d.data_ptr(), du.data_ptr(), B.data_ptr(), X.data_ptr(), ncol, nlev,
                                 nA, nrhs, x_stride, out.data_ptr(), _stream()))
    return out


def checksum(a, ncol=None, out=None):
    """Wrap-around uint64 sum of the raw bit patterns of a[..., :ncol] (last dim = column pitch).
    Returns/accumulates into a 1-element int64 CUDA tensor (bit pattern of the uint64 sum)."""
    if a.dtype not in _SFX or not a.is_cuda or a.stride(-1) != 1:
        raise ValueError("checksum needs a float32/float64 CUDA tensor with unit last stride")
    pitch = a.stride(-2) if a.dim() > 1 else a.shape[-1]
    nrows = a.numel() // a.shape[-1]
    if a.dim() > 2 and not a.flatten(0, -2).stride(0) == pitch:
        raise ValueError("checksum needs uniformly pitched rows")
    ncol = a.shape[-1] if ncol is None else ncol
    if out is None:
        out = torch.zeros(1, dtype=torch.int64, device=a.device)
    fn = f"ekat_b200_checksum_{_SFX[a.dtype]}"
    check(fn, getattr(lib(), fn)(a.data_ptr(), nrows, ncol, pitch, out.data_ptr(), _stream()))
    return out


def col_to_lev(src, pitch=None):
    """[ncol, n] (dim-0 may be strided) -> [n, pitch] level-major copy."""
    ncol, n = src.shape
    pitch = ncol if pitch is None else pitch
    dst = torch.empty((n, pitch), dtype=src.dtype, device=src.device)
    fn = f"ekat_b200_col_to_lev_{_SFX[src.dtype]}"
    check(fn, getattr(lib(), fn)(src.data_ptr(), dst.data_ptr(), ncol, n, src.stride(0) if ncol > 1 else n, pitch,
                                 _stream()))
    return dst


def lev_to_col(src, ncol=None):
    """[n, pitch] -> [ncol, n] column-major copy."""
    n, pitch = src.shape
    ncol = pitch if ncol is None else ncol
    dst = torch.empty((ncol, n), dtype=src.dtype, device=src.device)
    fn = f"ekat_b200_lev_to_col_{_SFX[src.dtype]}"
    check(fn, getattr(lib(), fn)(src.data_ptr(), dst.data_ptr(), ncol, n, src.stride(0), n, _stream()))
    return dst


def fill(dtype, ncol, nlev, nA, nrhs, layout=LEV_MAJOR, col0=0, seed_mod=32768, device="cuda", pitch=None, out=None):
    """Device-side reference fills (tests/algorithm/tridiag_tests.hpp:11-40), seed = (col0+c) % seed_mod.
    Returns (dl, d, du, X) in the requested layout. out = (dl, d, du, X) from an earlier call with the same layout
    and a pitch >= ncol: refill those buffers in place (a resident chunk buffer streamed over a larger batch)."""
    if out is not None:
        dl, d, du, X = out
        _, _, _, _, a_stride, x_stride = _geometry(dl, d, du, X, layout, ncol)
    elif layout == LEV_MAJOR:
        pitch = ncol if pitch is None else pitch
        A = torch.empty((3, nA, nlev, pitch), dtype=dtype, device=device)
        X = torch.empty((nrhs, nlev, pitch), dtype=dtype, device=device)
        dl, d, du = A[0], A[1], A[2]
        a_stride = x_stride = pitch
    else:
        A = torch.empty((ncol, 3, nlev, nA), dtype=dtype, device=device)
        X = torch.empty((ncol, nlev, nrhs), dtype=dtype, device=device)
        dl, d, du = A[:, 0], A[:, 1], A[:, 2]
        a_stride, x_stride = 3 * nlev * nA, nlev * nrhs
    fn = f"ekat_b200_fill_tridiag_{_SFX[dtype]}"
    check(fn, getattr(lib(), fn)(dl.data_ptr(), d.data_ptr(), du.data_ptr(), X.data_ptr(), ncol, col0, nlev, nA, nrhs,
                                 a_stride, x_stride, layout, seed_mod, _stream()))
    return dl, d, du, X


def transpose3(src, c2f=True, f_pitch=None, ni=None, out=None):
    """ekat::transpose<c2f|f2c> for rank-3 arrays on the device (reference src/kokkos/ekat_math_utils.hpp:124-147).
    c2f: src [ni, nk, nj] (dim-0 may be strided, e.g. the reference batch X[col][lev][rhs]) -> [nj, nk, f_pitch]
    (the native X[rhs][lev][col]); f2c: src [nj, nk, pitch] -> [ni, nk, nj]."""
    fn = f"ekat_b200_transpose3_{_SFX[src.dtype]}"
    if c2f:
        n_i, nk, nj = src.shape
        _require(src.stride(2) == 1 and src.stride(1) == nj, "transpose3: C side must be [ni][nk*nj] rows")
        f_pitch = n_i if f_pitch is None else f_pitch
        dst = torch.empty((nj, nk, f_pitch), dtype=src.dtype, device=src.device) if out is None else out
        _require(dst.shape == (nj, nk, f_pitch) and dst.is_contiguous() and dst.dtype == src.dtype, "transpose3: out")
        check(fn, getattr(lib(), fn)(src.data_ptr(), dst.data_ptr(), n_i, nk, nj, src.stride(0) if n_i > 1 else nk * nj,
                                     f_pitch, 1, _stream()))
    else:
        nj, nk, pitch = src.shape
        _require(src.is_contiguous(), "transpose3: Fortran side must be contiguous [nj][nk][pitch]")
        n_i = pitch if ni is None else ni
        dst = torch.empty((n_i, nk, nj), dtype=src.dtype, device=src.device) if out is None else out
        _require(dst.shape == (n_i, nk, nj) and dst.is_contiguous() and dst.dtype == src.dtype, "transpose3: out")
        check(fn, getattr(lib(), fn)(src.data_ptr(), dst.data_ptr(), n_i, nk, nj, nk * nj, pitch, 0, _stream()))
    return dst


def repack(src, pitch, zero_pad=True):
    """Row repacking of host_to_device / device_to_host (reference src/pack/ekat_pack_kokkos.hpp:332-1035):
    [nrows, n] scalars -> [nrows, pitch] (pitch = npack*N >= n), padding lanes zero if zero_pad."""
    nrows, n = src.shape
    _require(src.stride(1) == 1 and pitch >= n, "repack: unit last stride and pitch >= n")
    dst = torch.empty((nrows, pitch), dtype=src.dtype, device=src.device)
    fn = f"ekat_b200_repack_{_SFX[src.dtype]}"
    check(fn, getattr(lib(), fn)(src.data_ptr(), dst.data_ptr(), nrows, n, src.stride(0) if nrows > 1 else n, pitch,
                                 1 if zero_pad else 0, _stream()))
    return dst


def diffusion_solve(z, kd, X, ncol=None):
    """Fused adj_diff -> (dl, d, du) -> thomas for the implicit vertical-diffusion stencil (ekat_b200.h): z, kd, X
    [nlev, pitch] level-major, X overwritten with the solution. The diagonals never touch HBM."""
    nlev, pitch = X.shape
    for t in (z, kd, X):
        _require(t.is_cuda and t.shape == X.shape and t.is_contiguous() and t.dtype == X.dtype and t.dtype in _SFX,
                 "diffusion_solve: contiguous [nlev, pitch] CUDA tensors of one dtype")
    ncol = pitch if ncol is None else ncol
    fn = f"ekat_b200_diffusion_solve_{_SFX[X.dtype]}"
    check(fn, getattr(lib(), fn)(z.data_ptr(), kd.data_ptr(), X.data_ptr(), ncol, nlev, pitch, _stream()))
    return X


def diffusion_matrix(z, kd, ncol=None):
    """The matrix diffusion_solve implies, written out: returns (dl, d, du), each [nlev, pitch]."""
    nlev, pitch = z.shape
    for t in (z, kd):
        _require(t.is_cuda and t.shape == z.shape and t.is_contiguous() and t.dtype == z.dtype and t.dtype in _SFX,
                 "diffusion_matrix: contiguous [nlev, pitch] CUDA tensors of one dtype")
    ncol = pitch if ncol is None else ncol
    A = torch.empty((3, nlev, pitch), dtype=z.dtype, device=z.device)
    fn = f"ekat_b200_diffusion_matrix_{_SFX[z.dtype]}"
    check(fn, getattr(lib(), fn)(z.data_ptr(), kd.data_ptr(), A[0].data_ptr(), A[1].data_ptr(), A[2].data_ptr(), ncol, nlev,
                                 pitch, _stream()))
    return A[0], A[1], A[2]
