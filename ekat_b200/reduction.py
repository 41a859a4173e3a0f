"""Per-column vertical sums on torch CUDA tensors -- the Python face of
ekat::ReductionUtils<ExeSpace,Serialize>::view_reduction (reference:
src/algorithm/ekat_reduction_utils.hpp:67-144, 156-179; call form tests/algorithm/reduction_tests.cpp:117-126).

Where the reference is called once per column inside the caller's team kernel, this takes the whole batch
(SURVEY.md section 8b / 8f-2). The summation order is the reference's, so results are bit-identical to its host build:
  serialize=True   (ekatBFB):  element after element over the scalar range [begin, end);
  serialize=False, pack=N:     leading partial pack, per-lane sums of the full packs added in lane order, trailing
                               partial pack (what a team computes for a view of Pack<Real,N>).
Shapes: COL_MAJOR x [ncol, pitch] (the reference's rows of npack*N scalars; dim-0 may be strided);
        LEV_MAJOR x [nlev, pitch] (columns fastest).
"""
import torch

from . import COL_MAJOR, LEV_MAJOR
from ._lib import lib, check

_SFX = {torch.float64: "f64", torch.float32: "f32"}


def view_reduction(x, begin=0, end=None, serialize=True, pack=1, layout=COL_MAJOR, ncol=None, out=None):
    if not x.is_cuda or x.dtype not in _SFX or x.dim() != 2 or (x.numel() > 0 and x.shape[1] > 1 and x.stride(1) != 1):
        raise ValueError("view_reduction needs a 2-D float32/float64 CUDA tensor with unit last stride (no CPU path)")
    if layout == COL_MAJOR:
        nc, width = x.shape
        pitch = x.stride(0) if nc > 1 else width
        end = width if end is None else end
    elif layout == LEV_MAJOR:
        nlev, width = x.shape
        nc = width if ncol is None else ncol
        pitch = x.stride(0) if nlev > 1 else width
        end = nlev if end is None else end
        need = -(-end // pack) * pack if (not serialize and pack > 1) else end
        if need > nlev or nc > width:
            raise ValueError("view_reduction: range exceeds the array")
    else:
        raise ValueError("unknown layout")
    if out is None:
        out = torch.empty(nc, dtype=x.dtype, device=x.device)
    fn = f"ekat_b200_view_reduction_{_SFX[x.dtype]}"
    check(fn, getattr(lib(), fn)(x.data_ptr(), out.data_ptr(), nc, int(begin), int(end), pitch, layout,
                                 0 if serialize else int(pack), torch.cuda.current_stream().cuda_stream))
    return out


def adj_diff(x, forward=True, layout=COL_MAJOR, ncol=None, out=None):
    """ekat::adj_diff<Forward> at every index of every column (reference src/pack/ekat_pack_kokkos.hpp:100-164):
    forward: out[k] = (k < n-1 ? x[k+1] : 0) - x[k]; backward: out[k] = x[k] - (k > 0 ? x[k-1] : 0), over the n scalars
    of the scalarized column (a Pack view's padding lanes take part, as in the reference).
    COL_MAJOR x [ncol, n] (dim-0 may be strided); LEV_MAJOR x [n, pitch]."""
    if not x.is_cuda or x.dtype not in _SFX or x.dim() != 2 or (x.numel() > 0 and x.shape[1] > 1 and x.stride(1) != 1):
        raise ValueError("adj_diff needs a 2-D float32/float64 CUDA tensor with unit last stride (no CPU path)")
    if out is None:
        out = torch.empty(x.shape, dtype=x.dtype, device=x.device)
    if out.shape != x.shape or out.dtype != x.dtype or (out.numel() > 0 and out.shape[1] > 1 and out.stride(1) != 1):
        raise ValueError("adj_diff: out must match x")
    if layout == COL_MAJOR:
        nc, n = x.shape
        pin, pout = (x.stride(0) if nc > 1 else n), (out.stride(0) if nc > 1 else n)
    else:
        n, width = x.shape
        nc = width if ncol is None else ncol
        pin, pout = (x.stride(0) if n > 1 else width), (out.stride(0) if n > 1 else width)
    fn = f"ekat_b200_adj_diff_{_SFX[x.dtype]}"
    check(fn, getattr(lib(), fn)(x.data_ptr(), out.data_ptr(), nc, n, pin, pout, layout, 1 if forward else 0,
                                 torch.cuda.current_stream().cuda_stream))
    return out


def _dims(x, layout, ncol):
    """(ncol, n, pitch) of a 2-D operand in either layout."""
    if layout == COL_MAJOR:
        nc, n = x.shape
        return nc, n, (x.stride(0) if nc > 1 else n)
    n, width = x.shape
    return (width if ncol is None else ncol), n, (x.stride(0) if n > 1 else width)


def shift(x, shift_by, fill=0.0, layout=COL_MAJOR, ncol=None, out=None):
    """ekat::shift_right (shift_by = +1) / ekat::shift_left (-1) applied pack after pack along every (scalarized) column
    (reference src/pack/ekat_pack.hpp:521-555): out[k] = x[k - shift_by], `fill` where that falls off the column.
    COL_MAJOR x [ncol, n] (dim-0 may be strided); LEV_MAJOR x [n, pitch]."""
    if not x.is_cuda or x.dtype not in _SFX or x.dim() != 2 or (x.numel() > 0 and x.shape[1] > 1 and x.stride(1) != 1):
        raise ValueError("shift needs a 2-D float32/float64 CUDA tensor with unit last stride (no CPU path)")
    if out is None:
        out = torch.empty(x.shape, dtype=x.dtype, device=x.device)
    if out.shape != x.shape or out.dtype != x.dtype or (out.numel() > 0 and out.shape[1] > 1 and out.stride(1) != 1):
        raise ValueError("shift: out must match x")
    nc, n, pin = _dims(x, layout, ncol)
    _, _, pout = _dims(out, layout, ncol)
    fn = f"ekat_b200_shift_{_SFX[x.dtype]}"
    check(fn, getattr(lib(), fn)(x.data_ptr(), out.data_ptr(), nc, n, pin, pout, layout, int(shift_by), float(fill),
                                 torch.cuda.current_stream().cuda_stream))
    return out


def index_and_shift(a, idx, shift_by, layout=COL_MAJOR, ncol=None, out=None, out_shift=None):
    """ekat::index_and_shift<shift_by> for every index of every column (reference src/pack/ekat_pack_kokkos.hpp:71-98):
    out[j] = a[idx[j]], out_shift[j] = a[idx[j] + shift_by]; where an index leaves the column, out[j] is left as it was
    and out_shift[j] is NaN (the reference's debug-build behaviour). a [ncol, n] / idx [ncol, m] int32 (COL_MAJOR) or
    a [n, pitch] / idx [m, pitch] (LEV_MAJOR). Returns (out, out_shift)."""
    if not a.is_cuda or a.dtype not in _SFX or a.dim() != 2 or not idx.is_cuda or idx.dtype != torch.int32 or idx.dim() != 2:
        raise ValueError("index_and_shift needs a 2-D float CUDA tensor and a 2-D int32 CUDA index tensor (no CPU path)")
    for t in (a, idx):
        if t.numel() > 0 and t.shape[1] > 1 and t.stride(1) != 1:
            raise ValueError("index_and_shift: unit last stride required")
    if out is None:
        out = torch.zeros(idx.shape, dtype=a.dtype, device=a.device)
    if out_shift is None:
        out_shift = torch.zeros(idx.shape, dtype=a.dtype, device=a.device)
    for t in (out, out_shift):
        if t.shape != idx.shape or t.dtype != a.dtype or not t.is_contiguous():
            raise ValueError("index_and_shift: outputs must be contiguous and shaped like idx")
    nc, n, pa = _dims(a, layout, ncol)
    nci, m, pi = _dims(idx, layout, ncol)
    if nci != nc:
        raise ValueError("a and idx disagree on the number of columns")
    _, _, po = _dims(out, layout, ncol)
    fn = f"ekat_b200_index_and_shift_{_SFX[a.dtype]}"
    check(fn, getattr(lib(), fn)(a.data_ptr(), idx.data_ptr(), out.data_ptr(), out_shift.data_ptr(), nc, n, m, pa, pi, po,
                                 layout, int(shift_by), torch.cuda.current_stream().cuda_stream))
    return out, out_shift
