"""Latency experiments at small batch sizes (BASELINE config 1 and below): kernel duration by CUDA events with the
device kept busy in front of the launch, mean of 20, for a list of (cols_per_cta, prefetch, variant) tunings.
    python tools/exp_small.py solver nlev ncol[,ncol...] [c,p,v ...]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ekat_b200  # noqa: E402
from ekat_b200 import tridiag as T  # noqa: E402
from ekat_b200._lib import lib  # noqa: E402

dev = torch.device("cuda:0")
solver, nlev = sys.argv[1], int(sys.argv[2])
ncols = [int(v) for v in sys.argv[3].split(",")]
tunings = [tuple(int(v) for v in a.split(",")) for a in sys.argv[4:]] or [(0, 0, 0)]
layout = ekat_b200.COL_MAJOR if os.environ.get("EXP_CM") else ekat_b200.LEV_MAJOR
dt = torch.float32 if os.environ.get("EXP_DT") == "f32" else torch.float64
nrhs = int(os.environ.get("EXP_NRHS", "1"))
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev) if os.environ.get("EXP_COLD") else None
# floor of the timing method: the same event pair around a near-empty kernel of the library
tiny = torch.ones((1, 4), dtype=torch.float64, device=dev)
acc = torch.zeros(1, dtype=torch.int64, device=dev)
for _ in range(3):
    T.checksum(tiny, out=acc)
tot = 0.0
for _ in range(20):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda._sleep(200_000)
    a.record()
    T.checksum(tiny, out=acc)
    b.record()
    torch.cuda.synchronize()
    tot += a.elapsed_time(b)
print(f"floor (one tiny kernel between the events): {tot / 20 * 1e3:.2f} us", flush=True)
for ncol in ncols:
    dl, d, du, B = T.fill(dt, ncol, nlev, 1, nrhs, layout=layout, device=dev)
    X = B.clone()
    for tu in tunings:
        lib().ekat_b200_set_tuning(*tu)
        try:
            for _ in range(3):
                X.copy_(B)
                getattr(T, solver)(dl, d, du, X, layout=layout)
            torch.cuda.synchronize()
            tot, n = 0.0, 20
            for _ in range(n):
                X.copy_(B)
                if flush is not None:
                    flush.fill_(1)
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda._sleep(200_000)
                a.record()
                getattr(T, solver)(dl, d, du, X, layout=layout)
                b.record()
                torch.cuda.synchronize()
                tot += a.elapsed_time(b)
            # back to back on the device: 50 launches captured in a CUDA graph, replayed between one event pair
            # (no restore: every launch solves the previous solution again; no host launch path in the interval)
            st = torch.cuda.Stream()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.stream(st):
                getattr(T, solver)(dl, d, du, X, layout=layout)
                torch.cuda.synchronize()
                with torch.cuda.graph(g, stream=st):
                    for _ in range(50):
                        getattr(T, solver)(dl, d, du, X, layout=layout)
            g.replay()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            g.replay()
            b.record()
            torch.cuda.synchronize()
            b2b = a.elapsed_time(b) / 50
            # host path of the C ABI: the raw ctypes call with prebuilt arguments, 200 calls, wall clock
            import time as _t
            from ekat_b200.tridiag import _geometry, _SFX
            nc, nl, nA_, nr, as_, xs_ = _geometry(dl, d, du, X, layout, None)
            fn = getattr(lib(), f"ekat_b200_tridiag_{solver}_{_SFX[X.dtype]}")
            args = (dl.data_ptr(), d.data_ptr(), du.data_ptr(), X.data_ptr(), nc, nl, nA_, nr, as_, xs_, layout, 0,
                    torch.cuda.current_stream().cuda_stream)
            for _ in range(10):
                fn(*args)
            torch.cuda.synchronize()
            t0 = _t.perf_counter()
            for _ in range(200):
                fn(*args)
            t_issue = (_t.perf_counter() - t0) / 200 * 1e6
            torch.cuda.synchronize()
            t_wall = (_t.perf_counter() - t0) / 200 * 1e6
            hostpath = f" | C-ABI call: issue {t_issue:.1f} us, wall {t_wall:.1f} us"
            k = lib().ekat_b200_last_kernel().decode()
            w = 8 if dt == torch.float64 else 4
            alg = (3 + 2 * nrhs) * nlev * w * ncol
            print(f"{solver} {ncol:>8} x {nlev} rhs={nrhs} tuning={tu}: single {tot / n * 1e3:8.2f} us | back-to-back {b2b * 1e3:8.2f} us "
                  f"{alg / b2b / 1e6:7.0f} GB/s  {k}{hostpath}", flush=True)
        except Exception as e:
            print(f"{solver} {ncol} tuning={tu}: {type(e).__name__} {e}")
        finally:
            lib().ekat_b200_set_tuning(0, 0, 0)
