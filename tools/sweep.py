"""Tuning / measurement sweep for the kernels (run on the GPU box):  python tools/sweep.py [what ...]
Prints one JSON line per measurement: kernel, shape, tuning, ms, col-lev/s, algorithmic GB/s, fraction of the
measured HBM peak. CUDA-event timing, 3 warm-ups, inputs restored outside the timed region; L2 is flushed
between iterations for working sets that would fit in it."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ekat_b200  # noqa: E402
from ekat_b200 import tridiag as T  # noqa: E402
from ekat_b200._lib import lib  # noqa: E402
from ekat_b200.lin_interp import LinInterp, fill as li_fill, tridiag_lin_interp  # noqa: E402

dev = torch.device("cuda:0")
PEAK = 6551.0
try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass
_flush = None


def flush_l2():
    global _flush
    if _flush is None:
        _flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    _flush.zero_()


def timeit(fn, restore=None, iters=10, flush=False):
    for _ in range(3):
        if restore:
            restore()
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        if restore:
            restore()
        if flush:
            flush_l2()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda._sleep(200_000)  # the host enqueues while the device is busy: the events then bracket the kernel only
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        tot += a.elapsed_time(b)
    return tot / iters


def time_graph(fn, n=50, reps=5, restore=None):
    """Warm (L2-resident) timing as bench.py's small rows: n back-to-back launches in a CUDA graph, ms per launch.
    restore: put in-place operands back before every timed replay (solving a solution again and again shrinks the values
    until the bit-exact solvers leave their fast division path -- the first version of the c1bfb sweep measured that)."""
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            for _ in range(n):
                fn()
        best = 1e9
        for _ in range(reps):
            if restore:
                restore()
                torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            g.replay()
            b.record()
            torch.cuda.synchronize()
            best = min(best, a.elapsed_time(b) / n)
    return best


def report(**kw):
    print(json.dumps(kw), flush=True)


def tridiag_case(solver, dt, ncol, nlev, nA=1, nrhs=1, layout=ekat_b200.LEV_MAJOR, tunings=((0, 0, 0),), tag=""):
    tdt = torch.float64 if dt == "f64" else torch.float32
    w = 8 if dt == "f64" else 4
    dl, d, du, B = T.fill(tdt, ncol, nlev, nA, nrhs, layout=layout, device=dev)
    X = B.clone()
    alg = (3 * nA + 2 * nrhs) * nlev * w * ncol
    small = alg < (200 << 20)
    for tu in tunings:
        lib().ekat_b200_set_tuning(*tu)
        try:
            ms = timeit(lambda: getattr(T, solver)(dl, d, du, X, layout=layout), restore=lambda: X.copy_(B), flush=small)
        except Exception as e:  # noqa: BLE001
            report(kernel=solver, dt=dt, ncol=ncol, nlev=nlev, nA=nA, nrhs=nrhs, layout=layout, tuning=tu, error=str(e))
            continue
        finally:
            lib().ekat_b200_set_tuning(0, 0, 0)
        report(kernel=solver, variant=lib().ekat_b200_last_kernel().decode(), dt=dt, ncol=ncol, nlev=nlev, nA=nA, nrhs=nrhs,
               layout="lev" if layout else "col", tuning=tu, ms=round(ms, 4), col_lev_per_s=ncol * nlev / ms * 1e3,
               gbs=round(alg / ms / 1e6, 1), frac=round(alg / ms / 1e6 / PEAK, 3), l2_flushed=small, tag=tag)


def li_case(dt, ncol, km1, km2, N=1, nf=1, tunings=((0, 0, 0),)):
    tdt = torch.float64 if dt == "f64" else torch.float32
    w = 8 if dt == "f64" else 4
    x1, y1, x2 = li_fill(tdt, ncol, km1, km2, N, device=dev)
    li = LinInterp(ncol, km1, km2, tdt, N, device=dev)
    y1f = y1.unsqueeze(0).repeat(nf, 1, 1).contiguous()
    y2 = torch.zeros((nf, ncol, li.pitch2), dtype=tdt, device=dev)
    alg = (km1 + km2) * (1 + nf) * w * ncol
    small = alg < (200 << 20)
    for tu in tunings:
        lib().ekat_b200_set_tuning(*tu)
        ms = timeit(lambda: li.run(x1, x2, y1f, y2), flush=small)
        lib().ekat_b200_set_tuning(0, 0, 0)
        report(kernel="lin_interp", dt=dt, ncol=ncol, km1=km1, km2=km2, pack=N, nfields=nf, tuning=tu, ms=round(ms, 4),
               col_lev_per_s=ncol * km2 / ms * 1e3, gbs=round(alg / ms / 1e6, 1), frac=round(alg / ms / 1e6 / PEAK, 3),
               l2_flushed=small)
    ms = timeit(lambda: (li.setup(x1, x2), li.lin_interp(x1, x2, y1f, y2)), flush=small)
    report(kernel="lin_interp setup+apply", dt=dt, ncol=ncol, km1=km1, km2=km2, pack=N, nfields=nf, ms=round(ms, 4),
           gbs=round(alg / ms / 1e6, 1), frac=round(alg / ms / 1e6 / PEAK, 3))


def fused_case(dt, ncol, nlev, km2, nrhs, tunings=((0, 0, 0),)):
    tdt = torch.float64 if dt == "f64" else torch.float32
    w = 8 if dt == "f64" else 4
    dl, d, du, B = T.fill(tdt, ncol, nlev, 1, nrhs, device=dev)
    x1, _, x2 = li_fill(tdt, ncol, nlev, km2, 1, device=dev, lev_major=True)
    Y2 = torch.zeros((nrhs, km2, ncol), dtype=tdt, device=dev)
    alg = ((3 + nrhs + 1) * nlev + (1 + nrhs) * km2) * w * ncol
    for tu in tunings:
        lib().ekat_b200_set_tuning(*tu)
        ms = timeit(lambda: tridiag_lin_interp(dl[0], d[0], du[0], B, x1, x2, Y2))
        lib().ekat_b200_set_tuning(0, 0, 0)
        report(kernel="tridiag_lin_interp", dt=dt, ncol=ncol, nlev=nlev, km2=km2, nrhs=nrhs, tuning=tu, ms=round(ms, 4),
               col_lev_per_s=ncol * nlev / ms * 1e3, gbs=round(alg / ms / 1e6, 1), frac=round(alg / ms / 1e6 / PEAK, 3))


def li_lm_case(dt, ncol, km1, km2, nf, tunings=((0, 0, 0),)):
    from ekat_b200.lin_interp import lin_interp_lm
    tdt = torch.float64 if dt == "f64" else torch.float32
    w = 8 if dt == "f64" else 4
    x1, _, x2 = li_fill(tdt, ncol, km1, km2, 1, device=dev, lev_major=True)
    y1 = torch.rand((nf, km1, ncol), dtype=tdt, device=dev)
    y2 = torch.empty((nf, km2, ncol), dtype=tdt, device=dev)
    alg = (km1 + km2) * (1 + nf) * w * ncol
    for tu in tunings:
        lib().ekat_b200_set_tuning(*tu)
        ms = timeit(lambda: lin_interp_lm(x1, x2, y1, y2))
        lib().ekat_b200_set_tuning(0, 0, 0)
        report(kernel="lin_interp_lm", dt=dt, ncol=ncol, km1=km1, km2=km2, nfields=nf, tuning=tu, ms=round(ms, 4),
               gbs=round(alg / ms / 1e6, 1), frac=round(alg / ms / 1e6 / PEAK, 3))


def reduction_case(dt, ncol, nlev, layout=ekat_b200.COL_MAJOR, serialize=True, pack=16):
    from ekat_b200.reduction import view_reduction
    tdt = torch.float64 if dt == "f64" else torch.float32
    w = 8 if dt == "f64" else 4
    x = torch.rand((ncol, nlev) if layout == ekat_b200.COL_MAJOR else (nlev, ncol), dtype=tdt, device=dev)
    out = torch.empty(ncol, dtype=tdt, device=dev)
    alg = nlev * w * ncol
    small = alg < (200 << 20)
    ms = timeit(lambda: view_reduction(x, 0, nlev, serialize=serialize, pack=pack, layout=layout, out=out), flush=small)
    report(kernel="view_reduction", variant=lib().ekat_b200_last_kernel().decode(), dt=dt, ncol=ncol, nlev=nlev,
           layout="lev" if layout else "col", serialize=serialize, pack=pack, ms=round(ms, 4), col_lev_per_s=ncol * nlev / ms * 1e3,
           gbs=round(alg / ms / 1e6, 1), frac=round(alg / ms / 1e6 / PEAK, 3), l2_flushed=small)


def adj_diff_case(dt, ncol, nlev, layout=ekat_b200.COL_MAJOR, forward=True):
    from ekat_b200.reduction import adj_diff
    tdt = torch.float64 if dt == "f64" else torch.float32
    w = 8 if dt == "f64" else 4
    x = torch.rand((ncol, nlev) if layout == ekat_b200.COL_MAJOR else (nlev, ncol), dtype=tdt, device=dev)
    out = torch.empty_like(x)
    alg = 2 * nlev * w * ncol
    ms = timeit(lambda: adj_diff(x, forward, layout=layout, out=out), flush=alg < (200 << 20))
    report(kernel="adj_diff", variant=lib().ekat_b200_last_kernel().decode(), dt=dt, ncol=ncol, nlev=nlev,
           layout="lev" if layout else "col", forward=forward, ms=round(ms, 4), gbs=round(alg / ms / 1e6, 1),
           frac=round(alg / ms / 1e6 / PEAK, 3))


def copy_peak():
    n = 1 << 30
    a = torch.empty(n, dtype=torch.uint8, device=dev)
    b = torch.empty(n, dtype=torch.uint8, device=dev)
    ms = timeit(lambda: b.copy_(a))
    report(kernel="torch copy 1 GiB", ms=round(ms, 4), gbs=round(2 * n / ms / 1e6, 1))


if __name__ == "__main__":
    what = sys.argv[1:] or ["peak", "c3", "c1", "li", "multi", "col", "fused"]
    if "peak" in what:
        copy_peak()
    if "c3" in what:
        grid = [(t, p, 0) for t in (16, 32, 48, 64, 96) for p in (8, 16, 32)]
        for dt in ("f64", "f32"):
            for s in ("thomas", "bfb"):
                tridiag_case(s, dt, 345600, 128, tunings=grid)
            tridiag_case("cr", dt, 345600, 128, tunings=[(0, 0, 0), (8, 0, 0), (16, 0, 0), (32, 0, 0)])
    if "tma" in what:
        grid = [(0, 0, 0), (2, 2, 0), (2, 5, 0), (2, 8, 0), (2, 16, 0), (4, 4, 0)]
        for s in ("thomas", "bfb"):
            tridiag_case(s, "f64", 345600, 128, tunings=grid)
            tridiag_case(s, "f32", 345600, 128)
            tridiag_case(s, "f64", 3145728, 128)
            tridiag_case(s, "f64", 21600, 72)
    if "ring" in what:  # the bench's dominant kernel: ring stages 2..8 (fewer stages = less shared memory = a third CTA per SM), io warps off
        tridiag_case("thomas", "f64", 3145728, 128, tunings=[(0, p, 0) for p in (0, 2, 4, 6, 8, 9, 10, 11, 12, 14, 16)] + [(0, 0, 8)])
        tridiag_case("thomas", "f64", 345600, 128, tunings=[(0, p, 0) for p in (0, 8, 10, 11, 12, 16)])
        tridiag_case("thomas", "f32", 3145728, 128, tunings=[(0, p, 0) for p in (0, 8, 10, 12, 16)])
        tridiag_case("thomas", "f64", 1572864, 128, nrhs=10, tunings=[(0, p, 0) for p in (0, 8, 10, 12)])
    if "bfbring" in what:  # bfb_stream: ring stages against tiles per SM
        tridiag_case("bfb", "f64", 3145728, 128, tunings=[(c, 0, 0) for c in (0, 2, 3, 4, 5, 6)] + [(0, p, 0) for p in (2, 6, 8)])
        tridiag_case("thomas", "f64", 3145728, 128)
        tridiag_case("thomas", "f64", 345600, 128)
        tridiag_case("thomas", "f64", 345600, 72)
        tridiag_case("thomas", "f64", 21600, 72)
    if "f32split" in what:  # f32, one right-hand side: one-warp-per-tile kernel (default) against the warp-specialised one (variant 9)
        for ncol in (3145728, 345600):
            tridiag_case("thomas", "f32", ncol, 128, tunings=[(0, 0, 0), (0, 0, 9), (0, 8, 9), (0, 12, 9), (0, 16, 9)])
            tridiag_case("bfb", "f32", ncol, 128, tunings=[(0, 0, 0), (0, 0, 9)])
    if "li21k" in what:  # BASELINE config 2 at its stated 21,600 columns, warm: columns per tile
        for N in (1, 16):
            x1, y1, x2 = li_fill(torch.float64, 21600, 72, 128, N, device=dev)
            li = LinInterp(21600, 72, 128, torch.float64, N, device=dev)
            y2 = torch.zeros((21600, li.pitch2), dtype=torch.float64, device=dev)
            alg = (72 + 128) * 2 * 8 * 21600
            for tu in [(c, 0, v) for v in (0, 21, 22) for c in (0, 4, 8, 12, 16, 24, 32)]:
                lib().ekat_b200_set_tuning(*tu)
                ms = time_graph(lambda: li.run(x1, x2, y1, y2))
                lib().ekat_b200_set_tuning(0, 0, 0)
                report(kernel="lin_interp warm", pack=N, tuning=tu, us=round(ms * 1e3, 2), frac=round(alg / ms / 1e6 / PEAK, 3))
    if "c1bfb" in what:  # BASELINE config 1, bfb, warm: ring stages, plane prefetch distance, sweeping-warp forms, other kernels
        dl, d, du, B = T.fill(torch.float64, 21600, 72, 1, 1, device=dev)
        X = B.clone()
        alg = 5 * 72 * 8 * 21600
        for tu in [(0, 0, 0)] + [(c, p, v) for v in (0, 5, 11) for c in (0, 2, 3, 4) for p in (0, 2, 5, 9)] + [(0, 0, 4), (0, 0, 6), (0, 0, 9), (0, 0, 10)]:
            lib().ekat_b200_set_tuning(*tu)
            try:
                ms = time_graph(lambda: T.bfb(dl, d, du, X), restore=lambda: X.copy_(B))
                name = lib().ekat_b200_last_kernel().decode()
            except Exception as e:  # a pinned form may not support the shape
                ms, name = float("nan"), str(e)[:40]
            lib().ekat_b200_set_tuning(0, 0, 0)
            report(kernel=name, tuning=tu, us=round(ms * 1e3, 2), frac=round(alg / ms / 1e6 / PEAK, 3))
    if "split" in what:  # one right-hand side: warp-specialised kernel (default for thomas) against the one-warp kernel (variant 4)
        for s in ("bfb", "thomas"):
            for dt in ("f64", "f32"):
                tridiag_case(s, dt, 345600, 128, tunings=[(0, 0, 0), (0, 0, 4), (0, 0, 9)])
            tridiag_case(s, "f64", 3145728, 128, tunings=[(0, 0, 0), (0, 0, 4), (0, 0, 9)])
            tridiag_case(s, "f64", 21600, 72, tunings=[(0, 0, 0), (0, 0, 4), (0, 0, 9)])
            tridiag_case(s, "f64", 345600, 72, tunings=[(0, 0, 0), (0, 0, 4), (0, 0, 9)])
    if "bfb" in what:
        for dt in ("f64", "f32"):
            tridiag_case("bfb", dt, 345600, 128, tunings=[(0, 0, 0), (2, 0, 0), (3, 0, 0), (4, 0, 0), (2, 1, 0), (2, 5, 0), (0, 0, 4)])
        tridiag_case("bfb", "f64", 3145728, 128, tunings=[(0, 0, 0), (2, 0, 0), (3, 0, 0), (0, 0, 4)])
        tridiag_case("bfb", "f64", 345600, 72, tunings=[(0, 0, 0), (2, 0, 0), (3, 0, 0), (4, 0, 0)])
        tridiag_case("bfb", "f64", 345600, 128, layout=ekat_b200.COL_MAJOR, tunings=[(0, 0, 0), (2, 0, 0), (3, 0, 0), (0, 0, 4)])
        tridiag_case("bfb", "f64", 345600, 128, 10, 10)
    if "licb" in what:
        for km1 in (72, 128):
            li_case("f64", 345600, km1, 128, 1, tunings=[(8, 0, 0), (12, 0, 0), (16, 0, 0), (20, 0, 0), (24, 0, 0)])
            li_case("f32", 345600, km1, 128, 1, tunings=[(8, 0, 0), (16, 0, 0), (24, 0, 0), (32, 0, 0), (48, 0, 0)])
        li_case("f64", 3145728, 128, 128, 1, tunings=[(8, 0, 0), (16, 0, 0)])
        li_case("f64", 345600, 72, 128, 16, tunings=[(8, 0, 0), (16, 0, 0)])
    if "lilm" in what:
        from ekat_b200.lin_interp import lin_interp_lm
        for dt, tdt, w in (("f64", torch.float64, 8), ("f32", torch.float32, 4)):
            for ncol, km1, nf in ((345600, 128, 1), (345600, 128, 10), (345600, 72, 1), (1572864, 128, 10)):
                if dt == "f32" and ncol > 345600:
                    continue
                x1, y1, x2 = li_fill(tdt, ncol, km1, 128, 1, device=dev, lev_major=True)
                y1f = y1.unsqueeze(0).repeat(nf, 1, 1).contiguous()
                y2 = torch.zeros((nf, 128, ncol), dtype=tdt, device=dev)
                alg = (km1 + 128) * (1 + nf) * w * ncol
                for tu in ((0, 0, 0), (16, 1, 0), (16, 3, 0), (32, 1, 0)):
                    lib().ekat_b200_set_tuning(*tu)
                    ms = timeit(lambda: lin_interp_lm(x1, x2, y1f, y2))
                    lib().ekat_b200_set_tuning(0, 0, 0)
                    report(kernel="lin_interp_lm", dt=dt, ncol=ncol, km1=km1, km2=128, nfields=nf, tuning=tu, ms=round(ms, 4),
                           gbs=round(alg / ms / 1e6, 1), frac=round(alg / ms / 1e6 / PEAK, 3))
                del x1, y1, x2, y1f, y2
    if "c1" in what:
        for s in ("thomas", "bfb", "cr"):
            tridiag_case(s, "f64", 21600, 72, tunings=[(0, 0, 0), (16, 8, 0), (16, 16, 0), (32, 8, 0), (64, 16, 0)])
    if "li" in what:
        li_case("f64", 21600, 72, 128, 1, tunings=[(0, 0, 0), (0, 0, 1), (8, 0, 0), (32, 0, 0)])
        li_case("f64", 21600, 72, 128, 16)
        li_case("f64", 345600, 72, 128, 1, tunings=[(0, 0, 0), (0, 0, 1), (4, 0, 0), (8, 0, 0), (16, 0, 0), (32, 0, 0), (64, 0, 0)])
        li_case("f64", 345600, 128, 128, 1)
        li_case("f32", 345600, 72, 128, 1)
        li_case("f64", 345600, 72, 128, 1, nf=4)
    if "multi" in what:
        for s in ("thomas", "bfb"):
            tridiag_case(s, "f64", 345600, 128, 1, 10, tunings=[(0, 0, 0), (5, 3, 0), (2, 8, 0), (1, 8, 0), (0, 0, 1)])
            tridiag_case(s, "f32", 345600, 128, 1, 10)
            tridiag_case(s, "f64", 1572864, 128, 1, 10, tag="config 4")
            tridiag_case(s, "f64", 345600, 72, 1, 10)
            tridiag_case(s, "f64", 345600, 128, 10, 10)
        tridiag_case("cr", "f64", 100000, 128, 1, 10)
    if "col" in what:
        for s in ("thomas", "bfb", "cr"):
            tridiag_case(s, "f64", 345600, 128, layout=ekat_b200.COL_MAJOR, tunings=[(0, 0, 0), (8, 0, 0), (32, 0, 0)])
        tridiag_case("cr", "f64", 64, 2048, layout=ekat_b200.COL_MAJOR)
        tridiag_case("cr", "f64", 2048, 128, nrhs=43, nA=43, layout=ekat_b200.COL_MAJOR, tag="reference perf default")
    if "fused" in what:
        fused_case("f64", 393216, 128, 128, 10, tunings=[(0, 0, 0), (0, 0, 1)])
        fused_case("f64", 1572864, 128, 128, 10)
