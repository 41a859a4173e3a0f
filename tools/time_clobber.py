"""CLOBBER_DIAGS timing at config 3 (level-major): python tools/time_clobber.py"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ekat_b200
from ekat_b200 import tridiag as T
from ekat_b200._lib import lib
dev = torch.device("cuda:0")
for dt in (torch.float64, torch.float32):
    dl, d, du, B = T.fill(dt, 345600, 128, 1, 1, device=dev)
    X = B.clone(); l2, d2 = dl.clone(), d.clone()
    for solver in ("thomas", "bfb"):
        for clob in (False, True):
            ts = []
            for _ in range(6):
                X.copy_(B); l2.copy_(dl); d2.copy_(d)
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(); getattr(T, solver)(l2, d2, du, X, clobber=clob); b.record(); torch.cuda.synchronize()
                ts.append(a.elapsed_time(b))
            ms = min(ts[2:]); w = 8 if dt == torch.float64 else 4
            alg = 5 * w * 345600 * 128
            print(f"{solver} {dt} clobber={clob}: {ms*1e3:.1f} us  {alg/ms/1e6:.0f} GB/s (algorithmic 5 words)  {lib().ekat_b200_last_kernel().decode()}")
