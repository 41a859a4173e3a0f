"""A/B timing of tridiag_bfb_shared (bfb, several right-hand sides sharing one matrix) with and without the spare plane
slot:  python tools/time_bfb_shared.py   -> GB/s (algorithmic bytes) per shape and variant; checks bit-equality of both."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ekat_b200 import tridiag as T  # noqa: E402
from ekat_b200._lib import lib  # noqa: E402

dev = torch.device("cuda:0")
L = lib()
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timed(fn, reps=5):
    ms = []
    for _ in range(reps + 2):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    ms = sorted(ms[2:])
    return ms[len(ms) // 2]


for dt, ncol, nlev, nrhs in [(torch.float64, 1572864, 128, 10), (torch.float64, 786432, 256, 10), (torch.float64, 1572864, 128, 4),
                             (torch.float32, 1572864, 128, 10), (torch.float64, 1572864, 100, 6)]:
    dl, d, du, B = T.fill(dt, ncol, nlev, 1, nrhs, device=dev)
    w = 8 if dt == torch.float64 else 4
    alg = (3 + 2 * nrhs) * nlev * w * ncol
    outs = {}
    for variant in (15, 0):
        L.ekat_b200_set_tuning(0, 0, variant)
        X = B.clone()
        T.bfb(dl, d, du, X)
        name = L.ekat_b200_last_kernel().decode()
        ms = timed(lambda: T.bfb(dl, d, du, X))  # (re-solving the solution: same arithmetic cost)
        X.copy_(B)
        T.bfb(dl, d, du, X)
        outs[variant] = X.clone()
        print(f"{str(dt)[6:]} {ncol}x{nlev}x{nrhs} variant {variant:2d} {name:28s} {ms:8.3f} ms {alg / ms / 1e6:8.1f} GB/s", flush=True)
    L.ekat_b200_set_tuning(0, 0, 0)
    print("   bit-equal:", bool(torch.equal(outs[0].view(torch.uint8), outs[15].view(torch.uint8))), flush=True)
    del dl, d, du, B, X, outs
    torch.cuda.empty_cache()
