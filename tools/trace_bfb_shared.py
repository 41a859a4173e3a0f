"""Where a work item of k_tridiag_bfb_shared spends its cycles (needs the library built with -DEKB_TRACE_BFB:
    EKAT_B200_NVCC_FLAGS=-DEKB_TRACE_BFB python -m ekat_b200.build --force   # and a plain --force rebuild afterwards
    python tools/trace_bfb_shared.py [ncol nlev nrhs]"""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ekat_b200 import tridiag as T  # noqa: E402
from ekat_b200._lib import lib  # noqa: E402

ncol, nlev, nrhs = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (1572864, 128, 10)
dev = torch.device("cuda:0")
dl, d, du, B = T.fill(torch.float64, ncol, nlev, 1, nrhs, device=dev)
X = B.clone()
for _ in range(2):
    X.copy_(B)
    T.bfb(dl, d, du, X)
torch.cuda.synchronize()
print(lib().ekat_b200_last_kernel().decode())
buf = np.zeros(2 * 64 * 16, dtype=np.int64)
fn = lib().ekat_b200_debug_trace_bfb
fn.argtypes = [ctypes.c_void_p, ctypes.c_int]
assert fn(buf.ctypes.data_as(ctypes.c_void_p), buf.size) == 0
tr = buf.reshape(2, 64, 16)
names = ["fac_t0", "fac_t1", "fac_w_raw", "fac_w_plc", "fwd_t0", "fwd_t1", "fwd_w_plx", "fwd_w_ready", "bwd_t0", "bwd_t1",
         "bwd_w_fdone", "bwd_w_uraw", "bwd_w_store", "bwd_sec_load", "bwd_sec_compute", "bwd_sec_store"]
for cs in range(2):
    t = tr[cs]
    base = t[0, 0]
    print(f"--- CTA {cs}: item | factor start..end (len, wait raw, wait plc) | fwd0 start..end (wait plx, wait ready) | bwd0 start..end "
          f"(wait fdone, uraw, store) | bwd0 sections: load+arrive, rcp+chain, sts+fence+tma")
    for it in range(4, 14):
        r = t[it]
        print(f"{it:3d} | {r[0]-base:8d}..{r[1]-base:8d} ({r[1]-r[0]:6d}, {r[2]:6d}, {r[3]:6d}) | {r[4]-base:8d}..{r[5]-base:8d} "
              f"({r[6]:6d}, {r[7]:6d}) | {r[8]-base:8d}..{r[9]-base:8d} ({r[10]:6d}, {r[11]:6d}, {r[12]:6d}) | {r[13]:6d} {r[14]:6d} {r[15]:6d}")
    per = (t[23, 0] - t[4, 0]) / 19
    print(f"cycles per item (factor start to factor start): {per:.0f}")
