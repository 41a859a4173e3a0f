#!/usr/bin/env python
"""bench.py -- tridiag + LinInterp hot path on N B200s of one node (contract: see the task statement).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--cols-per-gpu C]
                    [--scaling weak|strong] [--global-cols G]

Headline workload (`config.workload`): BASELINE config 5 (ne1024pg2, 25,165,824 columns x 128 levels, f64)
column-sharded over 8 GPUs -> 3,145,728 columns per GPU, resident in HBM; weak scaling (N GPUs = N shards; N = 8 is
config 5 itself). One step = one pass of the hot path over the rank's shard:
    ekat::tridiag::thomas   (1 rhs, level-major)              40 B / column-level algorithmic
    ekat::LinInterp         (128 -> 128 levels, fused)        32 B / column-level algorithmic
metric = column-levels per second, summed over both kernels and all ranks.
Inputs (3.2 GB per array) are far larger than the 126 MB L2, so no flush is needed between steps.
The solvers work in place, so X is restored from a pristine copy between steps, outside the timed events
(the reference's perf driver does the same, tests/algorithm/tridiag_tests_performance.cpp:231-259).

Beside the headline the same JSON line carries
  strong   : config 5 AS STATED -- the fixed 25,165,824-column problem at this N: rank r owns columns
             [r*G/P, (r+1)*G/P) (ekat::Comm decomposition, src/core/ekat_comm.hpp:78-79) and streams them in chunks of
             <= 3,145,728 columns through the resident buffers (device-side fills per chunk, outside the events); one
             all-reduce at the end. Its checksum is the same number at N = 1, 2, 4, 8 (--scaling strong makes this
             leg the headline `value`).
  e2e      : the headline step through the host-buffer C ABI (pinned host memory, H2D + D2H inside the timed region),
             whole shard, the two host calls overlapped on one pipeline.
  other_kernels (N = 1): every other cell of BASELINE configs 1-4 and the SURVEY 8(f) rows.
  cpu_baseline (N = 1, rank 0): the reference's own headers under OpenMP on a bounded sample.
`--impl reference` times that CPU implementation on the headline workload itself (all host threads).
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

NLEV = 128
KM2 = 128
COLS_PER_GPU = 3_145_728          # 25,165,824 / 8
GLOBAL_COLS = 25_165_824          # BASELINE config 5 (ne1024pg2)
ALG_BYTES_TRIDIAG = 40            # (3 + 2*1) * 8 B per column-level          
ALG_BYTES_LI = (2 * NLEV + 2 * KM2) * 8 / KM2  # 32 B per target column-level  (3,200 B/col at 72->128; 4,096 here)
METRIC = "column-levels/s (tridiag thomas + LinInterp)"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cols-per-gpu", type=int, default=COLS_PER_GPU)
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="which leg is the headline `value` (the other is still reported beside it)")
    ap.add_argument("--global-cols", type=int, default=GLOBAL_COLS, help="column count of the strong-scaling problem")
    ap.add_argument("--no-strong", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the other kernels of the path (N = 1 only)")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons of one GPU while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, line in self.rows:
            if ts < t0 - 0.05 or ts > t1 + 0.05:
                continue
            f = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------ reference arm / cpu baseline
def host_cores():
    """Host threads this process may use. torch.distributed.run exports OMP_NUM_THREADS=1 to its children; the CPU
    arm must not inherit that (round 1's N >= 2 reference lines ran on one thread), so the count comes from the
    affinity mask and is handed to OpenMP explicitly."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except (AttributeError, OSError):
        return max(1, os.cpu_count() or 1)


def host_available_bytes():
    try:
        import psutil
        return int(psutil.virtual_memory().available)
    except Exception:
        return 64 << 30


def cpu_oracle():
    """oracle/_ref (the unchanged reference headers, OpenMP league loop) when it was built, else the plain-C port;
    thread count = every host core this process may use."""
    cores = host_cores()  # before libgomp loads: OMP_PROC_BIND then pins this thread to one place
    import oracle
    orc = oracle.ref() or oracle.port()
    orc.set_threads(cores)
    return orc, ("reference" if orc.kind == "reference" else "port"), orc.max_threads()


def cpu_path(orc, ncol, steps, warmup):
    """The headline step on the host: thomas (serial form = the reference's CPU production path,
    tests/algorithm/tridiag_tests_performance.cpp:260-303) on ncol x 128 f64, 1 rhs, + LinInterp 128 -> 128 on the
    same ncol. Inputs restored between passes outside the timed region, as the reference driver does (:231-259);
    serial thomas modifies only d and X (ekat_tridiag.hpp:216-227), so only those are restored.
    Returns per-pass (tridiag seconds, LinInterp seconds)."""
    import numpy as np
    from ekat_b200 import synth
    A, B = orc.fill_tridiag(np.float64, ncol, NLEV, 1, 1, col0=0)
    d0 = A[:, 1].copy()
    x = B.copy()
    # LinInterp inputs: the numpy statement of the generator is single-threaded, so columns beyond the first 65,536
    # repeat those (same distribution of brackets per column; this arm's output is timed, not compared)
    blk = min(ncol, 65536)
    bx1, by1, bx2 = synth.lin_interp_inputs(np.float64, blk, NLEV, KM2, 1)
    reps = (ncol + blk - 1) // blk
    x1, y1, x2 = (np.ascontiguousarray(np.tile(a, (reps, 1))[:ncol]) for a in (bx1, by1, bx2))
    y1f, y2 = y1[None], np.zeros((1, ncol, KM2))

    def one():
        np.copyto(A[:, 1], d0)  # restore (untimed)
        np.copyto(x, B)
        t0 = time.perf_counter()
        orc.tridiag("thomas", A, x)
        t1 = time.perf_counter()
        orc.lin_interp_raw(x1, x2, y1f, y2, None, NLEV, KM2, 1)
        t2 = time.perf_counter()
        return t1 - t0, t2 - t1

    for _ in range(warmup):
        one()
    return [one() for _ in range(steps)]


def cpu_summary(passes, ncol):
    best = min(passes, key=lambda p: p[0] + p[1])
    tt, tl = sum(p[0] for p in passes), sum(p[1] for p in passes)
    n = len(passes)
    units = ncol * (NLEV + KM2)
    return {"best_value": units / (best[0] + best[1]), "mean_value": n * units / (tt + tl),
            "best_ms": 1e3 * (best[0] + best[1]), "mean_ms": 1e3 * (tt + tl) / n,
            "tridiag_col_lev_per_s": ncol * NLEV / best[0], "lin_interp_col_lev_per_s": ncol * KM2 / best[1]}


def cpu_other_cells(orc, cores):
    """The reference's other solvers / Pack sizes on a small sample (best of 3): what BASELINE.md section 4 lists
    beside serial thomas -- bfb, cr, team thomas, and LinInterp with Pack<double,16> against scalar."""
    import numpy as np
    from ekat_b200 import synth
    ncol = 8192 * cores
    out = {"sample_columns": ncol}
    A, B = orc.fill_tridiag(np.float64, ncol, NLEV, 1, 1, col0=0)
    for solver in ("bfb", "cr", "thomas_team"):
        best = 1e30
        for _ in range(3):
            a, x = A.copy(), B.copy()
            t0 = time.perf_counter()
            orc.tridiag(solver, a, x)
            best = min(best, time.perf_counter() - t0)
        out[f"tridiag_{solver}_col_lev_per_s"] = ncol * NLEV / best
    if orc.kind == "reference":
        for N in (16, 1):
            x1, y1, x2 = synth.lin_interp_inputs(np.float64, ncol, 72, 128, N)
            y1f, y2 = np.ascontiguousarray(y1[None]), np.zeros((1, ncol, x2.shape[1]))
            best = 1e30
            for _ in range(3):
                t0 = time.perf_counter()
                orc.lin_interp_raw(x1, x2, y1f, y2, None, 72, 128, N)
                best = min(best, time.perf_counter() - t0)
            out[f"lin_interp_72_128_pack{N}_col_lev_per_s"] = ncol * 128 / best
    return out


def run_reference(args):
    """`--impl reference`: the reference's own CPU implementation of the headline step on the headline workload
    (same column count as the GPU arm's shard unless host memory forbids it -- said in the line), all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    os.environ.setdefault("OMP_PROC_BIND", "spread")  # bin/test-launcher:103-107
    os.environ.setdefault("OMP_PLACES", "threads")
    os.environ["OMP_NUM_THREADS"] = str(host_cores())  # torch.distributed.run exports 1
    orc, kind, cores = cpu_oracle()
    ncol = args.cols_per_gpu
    per_col = (3 + 1 + 1 + 1 + 4) * NLEV * 8 + 4096  # A, B, d0, x, LinInterp arrays (+ slack)
    avail = host_available_bytes()
    same = True
    while ncol * per_col > 0.5 * avail and ncol > 65536:
        ncol //= 2
        same = False
    # bound the run: a pass is ~1.5 s of compute + ~2 s of restore at the full shard
    probe = cpu_path(orc, min(ncol, 262_144), 1, 1)[0]
    est = (probe[0] + probe[1]) * ncol / min(ncol, 262_144) * 2.5
    steps = max(1, min(args.steps, int(240.0 / max(est, 1e-3)) - args.warmup))
    t_all = time.perf_counter()
    passes = cpu_path(orc, ncol, steps, args.warmup)
    s = cpu_summary(passes, ncol)
    sample = (f"{ncol} columns x 128 levels f64 ({'the whole shard' if same else 'reduced: host memory'}): thomas (serial form) + "
              f"LinInterp 128->128, best of {steps} timed passes after {args.warmup} warm-up, OpenMP over {cores} threads; {orc.info()}")
    line = {"impl": "reference", "metric": METRIC, "value": s["best_value"], "unit": "column-levels/s",
            "n_gpus": args.gpus, "steps": steps, "warmup": args.warmup, "ms_per_step": s["best_ms"],
            "mean_value": s["mean_value"], "mean_ms_per_step": s["mean_ms"],
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args.cols_per_gpu), "cpu_columns": ncol, "same_columns_as_gpu_arm": same,
                       "note": "CPU arm: the reference headers (Kokkos stand-in, OpenMP league loop); value = best pass"},
            "cpu_baseline": {"value": s["best_value"], "unit": "column-levels/s", "cores": cores, "kind": kind, "sample": sample,
                             "tridiag_col_lev_per_s": s["tridiag_col_lev_per_s"],
                             "lin_interp_col_lev_per_s": s["lin_interp_col_lev_per_s"]},
            "e2e": {"value": s["best_value"], "unit": "column-levels/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "wall_s": time.perf_counter() - t_all}
    print(json.dumps(line))


def workload_name(cols):
    return (f"BASELINE config 5 shard (ne1024pg2 / 8): {cols} columns x {NLEV} levels per GPU, f64: "
            f"tridiag thomas 1 rhs (level-major) + LinInterp {NLEV}->{KM2} (fused setup+interp)")


# ------------------------------------------------------------------------------------------------ host placement
def bind_near_gpu(local):
    """Bind this rank's threads and its future (pinned) allocations to the NUMA node of its GPU
    (bin/test-launcher:103-107 binds the reference's ranks the same way). Best effort; reports what it did."""
    info = {"numa_node": None, "cpus": None, "mempolicy": None}
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        dom, rest = bus.split(":", 1)
        path = f"/sys/bus/pci/devices/{dom[-4:].lower()}:{rest.lower()}/numa_node"
        node = int(open(path).read().strip())
        info["numa_node"] = node
        if node < 0:
            return info
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus |= set(range(int(lo), int(hi or lo) + 1))
        mine = cpus & set(os.sched_getaffinity(0))
        if mine:
            os.sched_setaffinity(0, mine)
            info["cpus"] = len(mine)
        libc = ctypes.CDLL(None, use_errno=True)
        mask = ctypes.c_ulong(1 << node)
        MPOL_PREFERRED, SYS_set_mempolicy = 1, 238  # x86-64
        rc = libc.syscall(SYS_set_mempolicy, MPOL_PREFERRED, ctypes.byref(mask), ctypes.c_ulong(64))
        info["mempolicy"] = "preferred" if rc == 0 else f"errno {ctypes.get_errno()}"
    except Exception as e:  # no sysfs / nvml in this container: leave placement to the OS
        info["error"] = f"{type(e).__name__}: {e}"[:120]
    return info


# ------------------------------------------------------------------------------------------------ our arm
def run_ours(args):
    import torch
    import ekat_b200
    from ekat_b200 import tridiag as T
    from ekat_b200._lib import lib
    from ekat_b200.comm import Comm, reduce_verification
    from ekat_b200.lin_interp import LinInterp, fill as li_fill

    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: ekat_b200 has no CPU path (use --impl reference for the CPU arm)")
    placement = bind_near_gpu(local) if world > 1 else {"numa_node": None, "note": "single rank: not bound"}
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    comm = Comm.from_env(device=dev)
    rank, size = comm.rank(), comm.size()
    ncol = args.cols_per_gpu
    col0 = rank * ncol  # this rank's contiguous column range of the global batch (ekat::Comm-style decomposition)
    dt = torch.float64
    L = lib()
    steps, warmup = args.steps, max(args.warmup, 3)

    # ---- inputs resident in HBM ---------------------------------------------------------------------------
    dl, d, du, B = T.fill(dt, ncol, NLEV, 1, 1, layout=ekat_b200.LEV_MAJOR, col0=col0, device=dev)
    X = B.clone()
    x1, y1, x2 = li_fill(dt, ncol, NLEV, KM2, 1, col0=col0, device=dev)
    y2 = torch.zeros((ncol, KM2), dtype=dt, device=dev)
    li = LinInterp(ncol, NLEV, KM2, dt, 1, device=dev)

    def step(e=None, n=None):
        if e:
            e[0].record()
        T.thomas(dl, d, du, X, ncol=n)
        if e:
            e[1].record()
        if n is None or n == ncol:
            li.run(x1, x2, y1, y2)
        else:  # partial last chunk of the strong-scaling leg
            LinInterp(n, NLEV, KM2, dt, 1, device=dev).run(x1[:n], x2[:n], y1[:n], y2[:n])
        if e:
            e[2].record()

    X.copy_(B)
    T.thomas(dl, d, du, X)
    k_tri = L.ekat_b200_last_kernel().decode()
    li.run(x1, x2, y1, y2)
    k_li = L.ekat_b200_last_kernel().decode()

    # ---- weak leg: the resident shard, K timed steps ---------------------------------------------------------
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(steps)]
    for _ in range(warmup):
        X.copy_(B)
        step()
    torch.cuda.synchronize()
    comm.barrier()
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.25)
    n0 = L.ekat_b200_launch_count()
    torch.cuda.synchronize()
    t0 = time.time()
    for k in range(steps):
        X.copy_(B)  # restore, outside the timed events
        step(ev[k])
    torch.cuda.synchronize()
    comm.barrier()
    t1 = time.time()
    launches = int(L.ekat_b200_launch_count() - n0)
    clocks = sampler.stop(t0, t1)
    ms_tri = sum(e[0].elapsed_time(e[1]) for e in ev)
    ms_li = sum(e[1].elapsed_time(e[2]) for e in ev)
    t_ms = torch.tensor([ms_tri + ms_li, ms_tri, ms_li], dtype=torch.float64, device=dev)
    comm.all_reduce(t_ms, "max")  # max over ranks
    ms_total, ms_tri_max, ms_li_max = (float(v) for v in t_ms.tolist())

    # verification: the run's one real collective (checksum SUM + residual MAX over ranks)
    cks = T.checksum(X)
    T.checksum(y2, out=cks)
    res = T.residual(dl, d, du, B, X)
    checksum, residual = reduce_verification(comm, cks, res)
    weak_value = size * steps * ncol * (NLEV + KM2) / (ms_total * 1e-3)

    # ---- roofline of the dominant kernel (thomas) ----------------------------------------------------------
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    else:
        peak, peak_src = 6650.0, "B200_PROFILING.md fallback 6.65 TB/s (of fallback)"
    tri_gbs = ALG_BYTES_TRIDIAG * ncol * NLEV * steps / (ms_tri_max * 1e-3) / 1e9
    li_gbs = ALG_BYTES_LI * ncol * KM2 * steps / (ms_li_max * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": k_tri, "achieved": tri_gbs, "peak": peak, "unit": "GB/s",
                "frac": tri_gbs / peak, "frac_of_nominal_8TBs": tri_gbs / 8000.0, "peak_source": peak_src,
                "traffic": None, "algorithmic_bytes_per_launch": ALG_BYTES_TRIDIAG * ncol * NLEV,
                "avg_launch_ms": ms_tri_max / steps,
                "second_kernel": {"kernel": k_li, "achieved": li_gbs, "frac": li_gbs / peak,
                                  "algorithmic_bytes_per_launch": int(ALG_BYTES_LI * ncol * KM2),
                                  "avg_launch_ms": ms_li_max / steps}}
    attach_traffic(roofline, ncol, k_tri, k_li)

    # ---- end to end: host (pinned) buffers through the C-ABI host pipeline ---------------------------------
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(torch, dev, comm, ncol, dl, d, du, B, x1, y1, x2, min(steps, 3), world)
        e2e["placement"] = placement

    # ---- strong leg: the fixed global problem, chunk-streamed through the same resident buffers -------------
    strong = None
    if not args.no_strong:
        strong = run_strong(torch, comm, dev, args.global_cols, ncol, steps, (dl, d, du, B, X), (x1, y1, x2, y2), step,
                            T, li_fill, reduce_verification)

    extras = None
    if size == 1 and not args.no_extras:
        del dl, d, du, B, X, x1, y1, x2, y2
        torch.cuda.empty_cache()
        extras = run_extras(torch, dev, peak)

    if rank == 0:
        cpu = None
        if size == 1 and not args.no_cpu:
            cpu = run_cpu_baseline()
        headline_strong = args.scaling == "strong" and strong is not None
        value = strong["value"] if headline_strong else weak_value
        ms_step = strong["ms_per_step"] if headline_strong else ms_total / steps
        line = {"metric": METRIC, "value": value, "unit": "column-levels/s",
                "n_gpus": size, "steps": steps, "warmup": warmup, "ms_per_step": ms_step,
                "higher_is_better": True, "scaling": "strong" if headline_strong else "weak", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic",
                "config": {"workload": workload_name(ncol), "global_columns": ncol * size, "levels": NLEV,
                           "l2": "inputs (3.2 GB/array) >> 126 MB L2, no flush needed",
                           "restore": "X <- B device copy between steps, outside the timed CUDA events",
                           "timing": "CUDA events on the launching stream, summed over steps, max over ranks",
                           "parallelism": f"column ranges over {size} rank(s), no data-path collective"},
                "weak": {"value": weak_value, "ms_per_step": ms_total / steps, "columns_per_gpu": ncol,
                         "checksum_u64": hex(checksum)},
                "strong": strong,
                "hbm_gbs": {"tridiag": tri_gbs, "lin_interp": li_gbs,
                            "step": (ALG_BYTES_TRIDIAG * NLEV + ALG_BYTES_LI * KM2) * ncol * steps / (ms_total * 1e-3) / 1e9},
                "roofline": roofline, "other_kernels": extras, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
                "verification": {"checksum_u64": hex(checksum), "residual_max": residual,
                                 "residual_ok": residual <= 50 * 2.220446049250313e-16},
                "wall_s_timed_region": t1 - t0}
        print(json.dumps(line))
    if size > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


def attach_traffic(roofline, ncol, k_tri, k_li):
    """dram bytes per launch from the committed `ncu --set full` capture -- only if that capture was taken of the same
    kernels at the same size from the same kernel sources (hash of ekat_b200/csrc), else null: a stale number is
    worse than none."""
    prof = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(prof):
        return
    tj = json.load(open(prof))
    src = csrc_hash()
    if tj.get("ncol") != ncol:
        roofline["traffic_note"] = "profiles/traffic.json was captured at another size: not used"
        return
    ok_src = tj.get("csrc_sha16") == src
    for tgt, kern, key in ((roofline, k_tri, "tridiag_thomas_f64"), (roofline["second_kernel"], k_li, "lin_interp_f64")):
        ent = tj.get("kernels", {}).get(key)
        if ent and ent.get("kernel") == kern and (ok_src or ent.get("file_sha16") == file_hash(ent.get("file", ""))):
            tgt["traffic"] = ent.get("dram_bytes")
    roofline["traffic_source"] = tj.get("source")
    if roofline["traffic"] is None:
        roofline["traffic_note"] = "profiles/traffic.json does not match the kernels/sources of this run: not used"


def file_hash(rel):
    import hashlib
    p = os.path.join(ROOT, rel)
    if not rel or not os.path.exists(p):
        return None
    return hashlib.sha256(open(p, "rb").read()).hexdigest()[:16]


def csrc_hash():
    import hashlib
    h = hashlib.sha256()
    d = os.path.join(ROOT, "ekat_b200", "csrc")
    for fn in sorted(os.listdir(d)):
        h.update(fn.encode())
        h.update(open(os.path.join(d, fn), "rb").read())
    return h.hexdigest()[:16]


def run_strong(torch, comm, dev, gcols, chunk, steps, tri_bufs, li_bufs, step, T, li_fill, reduce_verification):
    """BASELINE config 5 as stated: the fixed `gcols`-column problem on this many ranks. Rank r owns
    [r*G/P, (r+1)*G/P) and streams it through the resident shard buffers in chunks of <= `chunk` columns: fill the
    chunk on the device (untimed), run the step (timed with events), fold the chunk into the running checksum /
    residual. One all-reduce at the end -- so the checksum is a function of the problem, not of N."""
    import ekat_b200
    rank, size = comm.rank(), comm.size()
    dl, d, du, B, X = tri_bufs
    x1, y1, x2, y2 = li_bufs
    lo, hi = comm.column_range(gcols)
    chunks = [(c, min(chunk, hi - c)) for c in range(lo, hi, chunk)]
    nsteps = max(1, min(steps, 3))
    cks = torch.zeros(1, dtype=torch.int64, device=dev)
    res = torch.zeros(1, dtype=torch.float64, device=dev)
    ms = 0.0
    for s in range(nsteps + 1):  # pass 0: warm-up + verification, then the timed passes
        evs = []
        for c0, n in chunks:
            T.fill(torch.float64, n, NLEV, 1, 1, layout=ekat_b200.LEV_MAJOR, col0=c0, out=(dl, d, du, B))
            X.copy_(B)
            li_fill(torch.float64, n, NLEV, KM2, 1, col0=c0, out=(x1, y1, x2))
            e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            step(e, n)
            evs.append(e)
            if s == 0:
                T.checksum(X, ncol=n, out=cks)
                T.checksum(y2[:n], out=cks)
                res = torch.maximum(res, T.residual(dl, d, du, B, X, ncol=n))
        torch.cuda.synchronize()
        if s > 0:
            ms += sum(e[0].elapsed_time(e[2]) for e in evs)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    comm.all_reduce(t, "max")
    checksum, residual = reduce_verification(comm, cks, res)
    ms_step = float(t.item()) / nsteps
    return {"global_columns": gcols, "levels": NLEV, "n_gpus": size, "columns_per_rank": hi - lo, "chunks_per_rank": len(chunks),
            "chunk_columns": chunk, "steps": nsteps, "ms_per_step": ms_step,
            "value": gcols * (NLEV + KM2) / (ms_step * 1e-3), "unit": "column-levels/s",
            "checksum_u64": hex(checksum), "residual_max": residual,
            "note": "fixed problem: the checksum must be identical at every N; device-side fills between chunks are outside the events"}


def run_cpu_baseline():
    """cpu_baseline of the contract: a bounded sample (10-30 s of CPU work) of the headline step on all host threads."""
    os.environ.setdefault("OMP_PROC_BIND", "spread")
    os.environ.setdefault("OMP_PLACES", "threads")
    orc, kind, cores = cpu_oracle()
    ncol = max(4096, min(COLS_PER_GPU, 16384 * cores))
    passes = cpu_path(orc, ncol, 5, 1)
    s = cpu_summary(passes, ncol)
    return {"value": s["best_value"], "unit": "column-levels/s", "cores": cores, "kind": kind,
            "sample": f"{ncol} columns x 128 levels f64: thomas (serial form) + LinInterp 128->128, best of 5 passes "
                      f"(mean {s['mean_value']:.3e}), OpenMP over {cores} threads; {orc.info()}",
            "tridiag_col_lev_per_s": s["tridiag_col_lev_per_s"], "lin_interp_col_lev_per_s": s["lin_interp_col_lev_per_s"],
            "other_cells": cpu_other_cells(orc, cores)}


def run_extras(torch, dev, peak):
    """The other kernels of the hot path at BASELINE sizes, each timed alone with CUDA events: algorithmic GB/s and
    fraction of the measured peak. Not part of `value`; reported so that every cell of BASELINE configs 1-4 and every
    built row of SURVEY.md section 8 has a measured roofline fraction.
      ms        : kernel duration. Large cases: an event pair around one call while the device is kept busy by a short
                  spin kernel (the host's launch path is not inside the interval); mean of 3 after 2 warm-ups;
                  operands >> L2. Small cases (tens of microseconds): 50 launches captured in a CUDA graph and
                  replayed between one event pair, per launch (`timing` says so);
      cold_ms   : (small, L2-resident cases) the same with a 512 MB write between the launches, i.e. operands from
                  HBM (the fill's own time, measured the same way, subtracted);
      wall_us   : (small cases) host wall clock per call over 200 back-to-back calls through the Python face on a
                  warm stream, host launch path included;
      single_launch_event_pair_ms : (small cases) the large-case method applied to the small case, for continuity
                  with round 1's numbers (it carries the ~6.5 us floor of an event pair)."""
    import ekat_b200
    from ekat_b200 import tridiag as T
    from ekat_b200._lib import lib
    from ekat_b200.lin_interp import LinInterp, fill as li_fill
    out = []
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)

    def timed(fn, restore=None, reps=3, cold=False):
        for _ in range(2):
            if restore:
                restore()
            fn()
        torch.cuda.synchronize()
        tot = 0.0
        for _ in range(reps):
            if restore:
                restore()
            if cold:
                flush.fill_(1)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda._sleep(200_000)  # ~0.1 ms of device work: the launch below is queued behind it
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            tot += a.elapsed_time(b)
        return tot / reps

    def wall(fn, n=200):
        for _ in range(10):
            fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(n):
            fn()
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / n * 1e6

    def replay(fn, n=50, cold=False, restore=None):
        """Kernel duration of a small case: n launches captured in a CUDA graph and replayed between ONE event pair
        (launch gaps between consecutive kernels included, no host launch path, no per-event overhead -- an event
        pair around a single near-empty kernel reads ~6.5 us on this box, more than these kernels take). The solvers
        run in place, so launch i solves the solution of launch i-1 again: same work, no restore inside the graph;
        `restore` puts the right-hand sides back before the warm-up and again before the timed replay, so that the timed
        launches are solves 1..n of fresh data (a hundred solves on end shrink the values towards the range where the
        bit-exact solvers leave their fast division path: that would be timed, too).
        cold: a 512 MB fill between the launches (outside the kernel's own time it is subtracted as measured)."""
        st = torch.cuda.Stream()
        g = torch.cuda.CUDAGraph()
        if restore:
            restore()
            torch.cuda.synchronize()
        with torch.cuda.stream(st):
            fn()
            torch.cuda.synchronize()
            with torch.cuda.graph(g, stream=st):
                for _ in range(n):
                    if cold:
                        flush.fill_(1)
                    fn()
        g.replay()
        torch.cuda.synchronize()
        if restore:
            restore()
            torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        g.replay()
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / n
        if cold:
            g2 = torch.cuda.CUDAGraph()
            with torch.cuda.stream(st):
                with torch.cuda.graph(g2, stream=st):
                    for _ in range(n):
                        flush.fill_(1)
            g2.replay()
            torch.cuda.synchronize()
            a.record()
            g2.replay()
            b.record()
            torch.cuda.synchronize()
            ms -= a.elapsed_time(b) / n
        return ms

    def row(label, alg, units, ms, **kw):
        r = {"case": label, "kernel": lib().ekat_b200_last_kernel().decode(), "ms": ms, "gbs": alg / ms / 1e6,
             "frac": alg / ms / 1e6 / peak, "col_lev_per_s": units / ms * 1e3}
        r.update(kw)
        out.append(r)

    def tri(label, solver, ncol, nlev, nA, nrhs, layout, dt=torch.float64, clobber=False, small=False):
        w = 8 if dt == torch.float64 else 4
        dl, d, du, B = T.fill(dt, ncol, nlev, nA, nrhs, layout=layout, device=dev)
        X = B.clone()
        fn = lambda: getattr(T, solver)(dl, d, du, X, layout=layout, clobber=clobber)
        if clobber:
            A0 = (dl.clone(), d.clone(), du.clone())

            def restore():
                X.copy_(B)
                dl.copy_(A0[0])
                d.copy_(A0[1])
        else:
            restore = lambda: X.copy_(B)
        alg = (3 * nA + 2 * nrhs) * nlev * w * ncol
        kw = {}
        if small and not clobber:
            ms = replay(fn, restore=restore)
            kw["timing"] = "graph replay of 50 back-to-back launches"
            kw["cold_ms"] = replay(fn, cold=True, restore=restore)
            kw["cold_frac"] = alg / kw["cold_ms"] / 1e6 / peak
            kw["wall_us"] = wall(fn)
            kw["single_launch_event_pair_ms"] = timed(fn, restore=restore)
        else:
            ms = timed(fn, restore=restore)
        row(label, alg, ncol * nlev, ms, **kw)
        del dl, d, du, B, X
        torch.cuda.empty_cache()

    def lin(label, ncol, km1, km2, N, dt=torch.float64, small=False):
        w = 8 if dt == torch.float64 else 4
        x1, y1, x2 = li_fill(dt, ncol, km1, km2, N, device=dev)
        li = LinInterp(ncol, km1, km2, dt, N, device=dev)
        y2 = torch.zeros((ncol, li.pitch2), dtype=dt, device=dev)
        fn = lambda: li.run(x1, x2, y1, y2)
        alg = (2 * km1 + 2 * km2) * w * ncol
        kw = {}
        if small:
            ms = replay(fn)
            kw["timing"] = "graph replay of 50 back-to-back launches"
            kw["cold_ms"] = replay(fn, cold=True)
            kw["cold_frac"] = alg / kw["cold_ms"] / 1e6 / peak
            kw["wall_us"] = wall(fn)
            kw["single_launch_event_pair_ms"] = timed(fn)
        else:
            ms = timed(fn)
        row(label, alg, ncol * km2, ms, **kw)
        del x1, y1, x2, y2
        torch.cuda.empty_cache()

    LM, CMJ = ekat_b200.LEV_MAJOR, ekat_b200.COL_MAJOR
    f32 = torch.float32
    tri("config 5 shard, bfb (bit-exact), level-major", "bfb", 3145728, 128, 1, 1, LM)
    # config 4: 10 tracers sharing one matrix
    tri("config 4, thomas, 10 rhs sharing one matrix, level-major", "thomas", 1572864, 128, 1, 10, LM)
    tri("config 4, bfb, 10 rhs sharing one matrix, level-major", "bfb", 1572864, 128, 1, 10, LM)
    tri("config 4 shape in the reference layout X[col][lev][rhs], thomas, 10 rhs, 786,432 columns", "thomas", 786432, 128, 1, 10, CMJ)
    tri("config 4 shape in the reference layout, bfb, 10 rhs, 786,432 columns", "bfb", 786432, 128, 1, 10, CMJ)
    # config 3: thomas / cr / bfb, double and float, level-major; reference layout beside it
    for dt, nm in ((torch.float64, "f64"), (f32, "f32")):
        for solver in ("thomas", "bfb", "cr"):
            tri(f"config 3, {solver} {nm}, level-major", solver, 345600, 128, 1, 1, LM, dt)
    tri("config 3, thomas, reference (column-major) layout", "thomas", 345600, 128, 1, 1, CMJ)
    tri("config 3, bfb, reference (column-major) layout", "bfb", 345600, 128, 1, 1, CMJ)
    tri("config 3, cr, reference (column-major) layout", "cr", 345600, 128, 1, 1, CMJ)
    tri("config 3 shape, thomas, 4 rhs with one matrix each (nA == nrhs), level-major", "thomas", 345600, 128, 4, 4, LM)
    tri("config 3 shape, bfb, 4 rhs with one matrix each (nA == nrhs), level-major", "bfb", 345600, 128, 4, 4, LM)
    tri("config 3, thomas + CLOBBER_DIAGS (pivots written back to d), level-major", "thomas", 345600, 128, 1, 1, LM, clobber=True)
    tri("config 3, bfb + CLOBBER_DIAGS (multipliers, pivots written back), level-major", "bfb", 345600, 128, 1, 1, LM, clobber=True)
    # the reference perf driver's own default shape (tridiag_tests_performance.cpp:33-37): 2,048 problems x 128 rows x 43 rhs
    tri("reference perf-driver default: cr, 2,048 x 128 x 43 rhs, reference layout (190 MB)", "cr", 2048, 128, 1, 43, CMJ, small=True)
    tri("reference perf-driver default shape: thomas, 2,048 x 128 x 43 rhs, reference layout", "thomas", 2048, 128, 1, 43, CMJ, small=True)
    # cyclic reduction's own regime: few columns, many levels (tridiag_tests_correctness.cpp:410-414)
    tri("few columns / many levels: cr, 2,048 x 2,048 f64, level-major (335 MB)", "cr", 2048, 2048, 1, 1, LM, small=True)
    tri("few columns / many levels: thomas, 2,048 x 2,048 f64, level-major", "thomas", 2048, 2048, 1, 1, LM, small=True)
    tri("few columns / many levels: cr, 256 x 2,048 f64, level-major", "cr", 256, 2048, 1, 1, LM, small=True)
    tri("few columns / many levels: thomas, 256 x 2,048 f64, level-major", "thomas", 256, 2048, 1, 1, LM, small=True)
    # config 1: 21,600 x 72 (62 MB: L2-resident when warm)
    tri("config 1, thomas, 21,600 x 72 f64 (warm = L2-resident)", "thomas", 21600, 72, 1, 1, LM, small=True)
    tri("config 1, bfb (bit-exact), 21,600 x 72 f64 (warm = L2-resident)", "bfb", 21600, 72, 1, 1, LM, small=True)
    tri("config 1, thomas, 21,600 x 72 f64, reference layout", "thomas", 21600, 72, 1, 1, CMJ, small=True)
    # config 2: LinInterp 72 -> 128, Pack<double,16> against scalar pitches, at the stated 21,600 columns and at 345,600
    for N, nm in ((16, "Pack<double,16> pitch"), (1, "scalar pitch")):
        lin(f"config 2, LinInterp 72->128, {nm}, 21,600 columns (69 MB: warm = L2-resident)", 21600, 72, 128, N, small=True)
        lin(f"config 2 shape, LinInterp 72->128, {nm}, 345,600 columns", 345600, 72, 128, N)
    lin("LinInterp 128->128 f32, 3,145,728 columns", 3145728, 128, 128, 1, f32)
    lin("LinInterp 72->128 f32, 345,600 columns", 345600, 72, 128, 1, f32)

    # config 4 whole: 10-rhs solve + remap of the 10 solutions, one kernel (fused_tma.cu), f64 and f32
    from ekat_b200.lin_interp import tridiag_lin_interp
    ncol, nrhs = 1572864, 10
    for tdt, w in ((torch.float64, 8), (f32, 4)):
        dl, d, du, Bm = T.fill(tdt, ncol, 128, 1, nrhs, device=dev)
        x1, _, x2 = li_fill(tdt, ncol, 128, 128, 1, device=dev, lev_major=True)
        Y2 = torch.zeros((nrhs, 128, ncol), dtype=tdt, device=dev)
        ms = timed(lambda: tridiag_lin_interp(dl[0], d[0], du[0], Bm, x1, x2, Y2))
        alg = (3 + nrhs + 2 + nrhs) * 128 * w * ncol
        row(f"config 4, 10-rhs thomas + LinInterp 128->128 of the 10 solutions, f{8 * w} ({alg // ncol:,} B/column algorithmic)",
            alg, ncol * 128, ms)
        if not out[-1]["kernel"].startswith("solve_remap"):
            out[-1]["kernel"] = "tridiag_tma_shared + " + out[-1]["kernel"]
    del dl, d, du, Bm, x1, x2, Y2
    torch.cuda.empty_cache()
    # SURVEY.md section 8(f) rows, config 5 shard size
    torch.cuda.synchronize()
    time.sleep(0.5)  # let the board settle after the 10-rhs kernels: the first row here read 15% low right behind them
    from ekat_b200.reduction import view_reduction, adj_diff
    ncol = 3145728
    x = torch.rand((ncol, 128), dtype=torch.float64, device=dev)
    res = torch.empty(ncol, dtype=torch.float64, device=dev)
    for ser, label in ((True, "Serialize (bit-for-bit order)"), (False, "Pack<double,16> order")):
        ms = timed(lambda: view_reduction(x, 0, 128, serialize=ser, pack=16, out=res))
        row(f"8f-2 view_reduction, {ncol} x 128 f64, reference layout, {label}", 128 * 8 * ncol, ncol * 128, ms)
    dxo = torch.empty_like(x)
    ms = timed(lambda: adj_diff(x, True, out=dxo))
    row(f"8f-3 adj_diff<Forward>, {ncol} x 128 f64, reference layout (read + write)", 2 * 128 * 8 * ncol, ncol * 128, ms)
    del dxo
    ms = timed(lambda: T.col_to_lev(x))
    row(f"8f-1 marshalling [col][lev] -> [lev][col], {ncol} x 128 f64 (read + write)", 2 * 128 * 8 * ncol, ncol * 128, ms)
    del x, res
    torch.cuda.empty_cache()
    for fn in EXTRA_ROWS:  # rows registered by later additions (3-D marshalling, shift ops, fused difference + solve)
        try:
            fn(torch, dev, peak, timed, wall, row, replay)
        except Exception as e:  # a missing optional row must not cost the bench line
            out.append({"case": getattr(fn, "__name__", "extra"), "error": f"{type(e).__name__}: {e}"[:200]})
    return out


def rows_round2(torch, dev, peak, timed, wall, row, replay):
    """SURVEY.md section 8f-3 / 8f-1 rows added in round 2: the fused difference -> diagonals -> solve step beside its
    two-pass form, the shift / index_and_shift column ops and the rank-3 marshalling."""
    from ekat_b200 import tridiag as T
    from ekat_b200.reduction import shift, index_and_shift
    # ---- fused difference + fill + solve. Algorithmic traffic: z, kd, X in, X out = 4 words per column-level; the
    # two-pass form moves 10 (diffusion_matrix reads 2 + writes 3, thomas reads 4 + writes 1).
    for ncol, nlev, small in ((21600, 72, True), (345600, 72, False), (345600, 128, False)):
        z = torch.cumsum(torch.rand((nlev, ncol), dtype=torch.float64, device=dev) * 300 + 20, 0)
        kd = torch.rand((nlev, ncol), dtype=torch.float64, device=dev) * 3000
        B = torch.rand((nlev, ncol), dtype=torch.float64, device=dev)
        X = B.clone()
        A = torch.empty((3, 1, nlev, ncol), dtype=torch.float64, device=dev)
        L = T.lib()
        zp, kp, xp, s0 = z.data_ptr(), kd.data_ptr(), X.data_ptr(), torch.cuda.current_stream().cuda_stream

        def fused():
            T.diffusion_solve(z, kd, X)

        def two_pass():
            L.ekat_b200_diffusion_matrix_f64(zp, kp, A[0].data_ptr(), A[1].data_ptr(), A[2].data_ptr(), ncol, nlev, ncol,
                                             torch.cuda.current_stream().cuda_stream)
            T.thomas(A[0], A[1], A[2], X[None])

        alg = 4 * nlev * 8 * ncol
        for fn, nm, words in ((fused, "fused (one kernel, diagonals never in HBM)", 4), (two_pass, "two passes (diffusion_matrix + thomas)", 10)):
            if small:
                ms = replay(fn, restore=lambda: X.copy_(B))
                kw = {"timing": "graph replay of 50 back-to-back launches"}
            else:
                ms = timed(fn, restore=lambda: X.copy_(B))
                kw = {}
            row(f"8f-3 difference -> (dl,d,du) -> solve, {ncol:,} x {nlev} f64, {nm}", alg, ncol * nlev, ms,
                hbm_words_per_column_level=words, frac_of_own_traffic=words * nlev * 8 * ncol / ms / 1e6 / peak, **kw)
        del z, kd, B, X, A
        torch.cuda.empty_cache()
    # ---- shift / index_and_shift, reference layout, config 5 shard
    ncol = 3145728
    x = torch.rand((ncol, 128), dtype=torch.float64, device=dev)
    o = torch.empty_like(x)
    ms = timed(lambda: shift(x, 1, 0.0, out=o))
    row(f"8f-3 shift_right, {ncol:,} x 128 f64, reference layout (read + write)", 2 * 128 * 8 * ncol, ncol * 128, ms)
    idx = torch.randint(0, 127, (ncol, 128), dtype=torch.int32, device=dev)
    o2 = torch.empty_like(x)
    ms = timed(lambda: index_and_shift(x, idx, 1, out=o, out_shift=o2))
    row(f"8f-3 index_and_shift<1>, {ncol:,} x 128 f64 (read a + idx, write 2)", (3 * 8 + 4) * 128 * ncol, ncol * 128, ms)
    del idx, o, o2, x
    torch.cuda.empty_cache()
    # ---- rank-3 marshalling: the reference's X[col][lev][rhs] -> rhs planes [rhs][lev][col]
    ncol, nrhs = 786432, 10
    xc = torch.rand((ncol, 128, nrhs), dtype=torch.float64, device=dev)
    of = torch.empty((nrhs, 128, ncol), dtype=torch.float64, device=dev)
    ms = timed(lambda: T.transpose3(xc, True, out=of))
    row(f"8f-1 marshalling X[col][lev][rhs] -> [rhs][lev][col], {ncol:,} x 128 x {nrhs} f64 (read + write)", 2 * 128 * nrhs * 8 * ncol,
        ncol * 128, ms)


EXTRA_ROWS = [rows_round2]


def run_e2e(torch, dev, comm, ncol, dl, d, du, B, x1, y1, x2, steps, world):
    """Same step through ekat_b200_tridiag_host_* / ekat_b200_lin_interp_host_*: inputs in pinned host memory,
    H2D + solve + D2H inside the timed region (wall clock from the first call to ekat_b200_pipeline_sync, max over
    ranks). The whole shard unless pinning it would take more than a third of the host memory left for this rank.
    The two calls share one pipeline in deferred mode: LinInterp's uploads run under the solver's downloads."""
    import ekat_b200
    from ekat_b200.host import Pipeline
    per_col = 9 * NLEV * 8  # dl, d, du, B, X, x1, y1, x2, y2
    budget = host_available_bytes() / 3 / max(1, world)
    ne = ncol
    while ne * per_col > budget and ne > 65536:
        ne //= 2
    t_pin = time.perf_counter()

    def pinned_from(t):  # straight device -> pinned host copy (no pageable staging)
        h = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
        h.copy_(t)
        return h

    sl = (lambda t: t[0, :, :ne]) if ne != ncol else (lambda t: t[0])
    if ne != ncol:
        hdl, hd, hdu, hB = (pinned_from(sl(t).contiguous()) for t in (dl, d, du, B))
    else:
        hdl, hd, hdu, hB = (pinned_from(t[0]) for t in (dl, d, du, B))
    hX = torch.empty(hB.shape, dtype=hB.dtype, pin_memory=True)
    hx1, hy1, hx2 = (pinned_from(t[:ne]) for t in (x1, y1, x2))
    hy2 = torch.empty((ne, KM2), dtype=B.dtype, pin_memory=True)
    torch.cuda.synchronize()
    t_pin = time.perf_counter() - t_pin
    pipe = Pipeline(device_bytes=3 << 30, nslots=3)
    pipe.set_deferred(True)
    try:
        Bsrc = B[0] if ne == ncol else B[0, :, :ne]

        def one():
            hX.copy_(Bsrc)  # restore the right-hand sides from the resident copy (a DMA, untimed)
            torch.cuda.synchronize()
            comm.barrier()  # every rank's step starts together: the ranks share the host's memory bandwidth
            t0 = time.perf_counter()
            pipe.tridiag("thomas", hdl, hd, hdu, hX, layout=ekat_b200.LEV_MAJOR)
            pipe.lin_interp(hx1, hx2, hy1, hy2, NLEV, KM2)
            pipe.sync()
            return time.perf_counter() - t0
        one()
        s0 = pipe.stats()
        per_step = [one() for _ in range(steps)]
        tt = sum(per_step)
        s1 = pipe.stats()
    finally:
        pipe.close()
    if os.environ.get("EKAT_B200_BENCH_DEBUG"):
        print(f"[e2e rank {comm.rank()}] step ms: {[round(x * 1e3, 1) for x in per_step]}", file=sys.stderr, flush=True)
    t = torch.tensor([tt], dtype=torch.float64, device=dev)
    comm.all_reduce(t, "max")
    h2d = (s1["h2d_bytes"] - s0["h2d_bytes"]) // steps
    d2h = (s1["d2h_bytes"] - s0["d2h_bytes"]) // steps
    sec = float(t.item()) / steps
    return {"value": comm.size() * ne * (NLEV + KM2) / sec, "unit": "column-levels/s",
            "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": sec * 1e3,
            "host_link_gbs_per_gpu": {"h2d": h2d / sec / 1e9, "d2h": d2h / sec / 1e9},
            "columns_per_gpu": ne, "whole_shard": ne == ncol, "steps": steps, "pin_seconds": t_pin,
            "api": "ekat_b200_tridiag_host_f64 + ekat_b200_lin_interp_host_f64 on one deferred pipeline "
                   "(pinned host buffers, 3-slot chunk pipeline), ekat_b200_pipeline_sync"}


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
