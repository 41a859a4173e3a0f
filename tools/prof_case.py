"""Run a few launches of selected kernels at a fixed shape (for ncu captures):
    python tools/prof_case.py [case ...]   cases: thomas bfb cr li shared fused col  (default: thomas bfb li)
Each case launches its kernel 3 times at BASELINE config 3 size (345,600 columns x 128 levels, f64) unless
PROF_NCOL / PROF_NLEV / PROF_DT say otherwise (config 1: PROF_NCOL=21600 PROF_NLEV=72)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ekat_b200  # noqa: E402
from ekat_b200 import tridiag as T  # noqa: E402
from ekat_b200.lin_interp import LinInterp, fill as li_fill, tridiag_lin_interp  # noqa: E402

dev = torch.device("cuda:0")
NCOL, NLEV = int(os.environ.get("PROF_NCOL", 345600)), int(os.environ.get("PROF_NLEV", 128))
dt = torch.float32 if os.environ.get("PROF_DT") == "f32" else torch.float64
cases = sys.argv[1:] or ["thomas", "bfb", "li"]
REPS = 3

for c in cases:
    if c in ("thomas", "bfb", "cr"):
        dl, d, du, B = T.fill(dt, NCOL, NLEV, 1, 1, device=dev)
        X = B.clone()
        for _ in range(REPS):
            X.copy_(B)
            getattr(T, c)(dl, d, du, X)
    elif c == "shared":
        dl, d, du, B = T.fill(dt, NCOL, NLEV, 1, 10, device=dev)
        X = B.clone()
        for s in ("thomas", "bfb"):
            for _ in range(REPS):
                X.copy_(B)
                getattr(T, s)(dl, d, du, X)
    elif c == "col":
        dl, d, du, B = T.fill(dt, NCOL, NLEV, 1, 1, layout=ekat_b200.COL_MAJOR, device=dev)
        X = B.clone()
        for s in ("thomas", "bfb", "cr"):
            for _ in range(REPS):
                X.copy_(B)
                getattr(T, s)(dl, d, du, X, layout=ekat_b200.COL_MAJOR)
    elif c in ("shared_thomas", "shared_bfb"):
        dl, d, du, B = T.fill(dt, NCOL, NLEV, 1, 10, device=dev)
        X = B.clone()
        for _ in range(REPS):
            X.copy_(B)
            getattr(T, c.split("_")[1])(dl, d, du, X)
    elif c in ("col_thomas", "col_bfb"):
        dl, d, du, B = T.fill(dt, NCOL, NLEV, 1, 1, layout=ekat_b200.COL_MAJOR, device=dev)
        X = B.clone()
        for _ in range(REPS):
            X.copy_(B)
            getattr(T, c.split("_")[1])(dl, d, du, X, layout=ekat_b200.COL_MAJOR)
    elif c in ("cm10_thomas", "cm10_bfb"):  # reference layout X[col][lev][10], one matrix per column
        dl, d, du, B = T.fill(dt, NCOL, NLEV, 1, 10, layout=ekat_b200.COL_MAJOR, device=dev)
        X = B.clone()
        for _ in range(REPS):
            X.copy_(B)
            getattr(T, c.split("_")[1])(dl, d, du, X, layout=ekat_b200.COL_MAJOR)
    elif c == "li72":
        x1, y1, x2 = li_fill(dt, NCOL, 72, 128, 16, device=dev)
        li = LinInterp(NCOL, 72, 128, dt, 16, device=dev)
        y2 = torch.zeros((NCOL, 128), dtype=dt, device=dev)
        for _ in range(REPS):
            li.run(x1, x2, y1, y2)
    elif c == "li":
        for km1 in (72, 128):
            x1, y1, x2 = li_fill(dt, NCOL, km1, 128, 1, device=dev)
            li = LinInterp(NCOL, km1, 128, dt, 1, device=dev)
            y2 = torch.zeros((NCOL, 128), dtype=dt, device=dev)
            for _ in range(REPS):
                li.run(x1, x2, y1, y2)
    elif c == "fused":
        dl, d, du, B = T.fill(dt, NCOL, NLEV, 1, 10, device=dev)
        x1, _, x2 = li_fill(dt, NCOL, NLEV, 128, 1, device=dev, lev_major=True)
        Y2 = torch.zeros((10, 128, NCOL), dtype=dt, device=dev)
        for _ in range(REPS):
            tridiag_lin_interp(dl[0], d[0], du[0], B, x1, x2, Y2)
    torch.cuda.synchronize()
print("ok", cases)
