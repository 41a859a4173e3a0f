"""ekat_b200 -- B200 (sm_100a) implementation of EKAT's Algorithm-package hot path.

Host-side mirror of the reference interface on top of the C ABI in
include/ekat_b200/ekat_b200.h (libekat_b200.so, hand-written CUDA):

  ekat_b200.tridiag.thomas / bfb / cr      <- ekat::tridiag::{thomas,bfb,cr}  (src/algorithm/ekat_tridiag.hpp)
  ekat_b200.lin_interp.LinInterp           <- ekat::LinInterp                (src/algorithm/ekat_lin_interp.hpp)
  ekat_b200.reduction.view_reduction       <- ekat::ReductionUtils::view_reduction (src/algorithm/ekat_reduction_utils.hpp)
  ekat_b200.comm.Comm                      <- ekat::Comm                     (src/core/ekat_comm.hpp)

PyTorch is used for device memory, streams and torch.distributed only. There is no CPU or
PyTorch fallback: every operation goes through the CUDA library or raises.
"""
from . import _lib  # noqa: F401
from ._lib import EkatB200Error  # noqa: F401

COL_MAJOR = 0
LEV_MAJOR = 1
CLOBBER_DIAGS = 1

__all__ = ["COL_MAJOR", "LEV_MAJOR", "CLOBBER_DIAGS", "EkatB200Error"]
