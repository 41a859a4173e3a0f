"""Phase trace of the persistent shared-matrix kernel (tuning variant 7): python tools/trace_shared.py [ncol] [nlev] [nrhs]
Prints, per solver and traced CTA, mean cycles per work item spent in each phase / wait (items 8..63)."""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ekat_b200 import tridiag as T  # noqa: E402
from ekat_b200._lib import lib  # noqa: E402

dev = torch.device("cuda:0")
ncol = int(sys.argv[1]) if len(sys.argv) > 1 else 345600
nlev = int(sys.argv[2]) if len(sys.argv) > 2 else 128
nrhs = int(sys.argv[3]) if len(sys.argv) > 3 else 10
NAMES = ["f.waitB", "f.waitFactor", "f.lds", "f.math+sts", "b.waitDu", "b.lds", "b.math+sts", "b.fence", "b.stg", "b.commit",
         "b.waitread", "b.refill", "t_start", "factor", "factor.waitRing", "factor.waitPlane"]
for dt in (torch.float64,):
    dl, d, du, B = T.fill(dt, ncol, nlev, 1, nrhs, device=dev)
    X = B.clone()
    for solver in ("thomas", "bfb"):
        for tu in [(0, 0, 7)] + [tuple(int(v) for v in a.split(",")) for a in sys.argv[4:]]:
            lib().ekat_b200_set_tuning(*tu)
            for _ in range(2):
                X.copy_(B)
                getattr(T, solver)(dl, d, du, X)
            torch.cuda.synchronize()
            lib().ekat_b200_set_tuning(0, 0, 0)
            buf = np.zeros(2 * 64 * 16, dtype=np.int64)
            rc = lib().ekat_b200_debug_trace(buf.ctypes.data_as(ctypes.c_void_p), buf.size)
            assert rc == 0
            tr = buf.reshape(2, 64, 16)
            for c in range(2):
                sel = tr[c, 8:60]
                period = np.diff(sel[:, 12]).mean()
                print(f"{solver} tuning={tu} cta{c}: item period {period:.0f} cyc | " +
                      " ".join(f"{n}={sel[:, i].mean():.0f}" for i, n in enumerate(NAMES) if not n.endswith("t_start")))
