"""Timing of every bfb route (bit-exact solver) at the bench's shapes, GB/s of algorithmic bytes and fraction of the
measured HBM peak:  python tools/time_bfb.py  -- plus the flag rate / mismatch count of the certified quotient."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ekat_b200  # noqa: E402
from ekat_b200 import tridiag as T  # noqa: E402
from ekat_b200._lib import lib  # noqa: E402

dev = torch.device("cuda:0")
L = lib()
PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6551.0
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timed(fn, reps=7, cold=True):
    ms = []
    for _ in range(reps + 2):
        if cold:
            flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    ms = sorted(ms[2:])
    return ms[len(ms) // 2]


out = torch.zeros(1, dtype=torch.int64, device=dev)
for dt in ("f64", "f32"):
    fn = getattr(L, f"ekat_b200_selftest_div_{dt}")
    res = {}
    for mode in (2, 3, 4, 5):
        out.zero_()
        fn(1 << 26, 777 + mode, mode, out.data_ptr(), None)
        torch.cuda.synchronize()
        res[mode] = int(out.item())
    print(f"selftest {dt}: mismatches random-bits {res[2]}, well-scaled {res[3]}; flagged of 2^27: random-bits {res[4]}, well-scaled {res[5]}", flush=True)

LM, CM = ekat_b200.LEV_MAJOR, ekat_b200.COL_MAJOR
cases = [("config 5 shard bfb f64", torch.float64, 3145728, 128, 1, 1, LM, True),
         ("config 3 bfb f64", torch.float64, 345600, 128, 1, 1, LM, True),
         ("config 3 bfb f32", torch.float32, 345600, 128, 1, 1, LM, True),
         ("config 3 bfb f64 reference layout", torch.float64, 345600, 128, 1, 1, CM, True),
         ("config 1 bfb f64 (warm)", torch.float64, 21600, 72, 1, 1, LM, False),
         ("config 4 bfb f64 10 rhs", torch.float64, 1572864, 128, 1, 10, LM, True),
         ("config 4 bfb f32 10 rhs", torch.float32, 1572864, 128, 1, 10, LM, True),
         ("72 levels x 10 rhs bfb f64", torch.float64, 1572864, 72, 1, 10, LM, True),
         ("config 4 shape reference layout bfb f64 10 rhs", torch.float64, 786432, 128, 1, 10, CM, True),
         ("config 3 shape bfb f64 nA == nrhs == 4", torch.float64, 345600, 128, 4, 4, LM, True)]
for name, dt, ncol, nlev, nA, nrhs, layout, cold in cases:
    dl, d, du, B = T.fill(dt, ncol, nlev, nA, nrhs, layout=layout, device=dev)
    X = B.clone()
    T.bfb(dl, d, du, X, layout=layout)
    kern = L.ekat_b200_last_kernel().decode()
    ms = timed(lambda: T.bfb(dl, d, du, X, layout=layout), cold=cold)
    w = 8 if dt == torch.float64 else 4
    alg = (3 * nA + 2 * nrhs) * nlev * w * ncol
    print(f"{name:50s} {kern:30s} {ms:8.4f} ms {alg / ms / 1e6:8.1f} GB/s  {alg / ms / 1e6 / PEAK:5.3f}", flush=True)
    del dl, d, du, B, X
    torch.cuda.empty_cache()
