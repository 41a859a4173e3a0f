"""config 4 in one kernel: python tools/time_fused.py [ncol]"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ekat_b200 import tridiag as T
from ekat_b200._lib import lib
from ekat_b200.lin_interp import fill as li_fill, tridiag_lin_interp
dev = torch.device("cuda:0")
ncol = int(sys.argv[1]) if len(sys.argv) > 1 else 1572864
for dt in (torch.float64, torch.float32):
    for nrhs in (10, 6):
        dl, d, du, B = T.fill(dt, ncol, 128, 1, nrhs, device=dev)
        x1, _, x2 = li_fill(dt, ncol, 128, 128, 1, device=dev, lev_major=True)
        Y2 = torch.zeros((nrhs, 128, ncol), dtype=dt, device=dev)
        ts = []
        for _ in range(5):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); tridiag_lin_interp(dl[0], d[0], du[0], B, x1, x2, Y2); b.record(); torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        ms = min(ts[1:]); w = 8 if dt == torch.float64 else 4
        alg = (3 + nrhs + 2 + nrhs) * 128 * w * ncol
        print(f"fused {dt} nrhs={nrhs}: {ms:.3f} ms  {alg/ms/1e6:.0f} GB/s = {alg/ms/1e6/6551:.3f}  {lib().ekat_b200_last_kernel().decode()}")
        del dl, d, du, B, x1, x2, Y2
