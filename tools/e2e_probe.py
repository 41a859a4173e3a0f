"""Where does the host-buffer (e2e) leg lose bandwidth when several ranks run at once?

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/e2e_probe.py

Every rank pins the SAME buffers bench.py's run_e2e pins (a whole config-5 shard) and then, all ranks in lock step
(barrier before every case), times
  flat_h2d      plain contiguous cudaMemcpyAsync of those buffers into a device slot, 256 MiB pieces, 2 streams
  flat_duplex   the same with downloads into hX running on a third stream
  tri           Pipeline.tridiag alone            (nslots 3 / 3 GB, bench.py's setting)
  li            Pipeline.lin_interp alone
  both          bench.py's e2e step
  both_s6       the e2e step on 6 slots / 6 GB
  both_q        the e2e step on a quarter of the shard
Rank 0 prints, per case, every rank's GB/s up and down. Nothing here is a bench value."""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ekat_b200  # noqa: E402
from ekat_b200 import tridiag as T  # noqa: E402
from ekat_b200.host import Pipeline  # noqa: E402
from ekat_b200.lin_interp import fill as li_fill  # noqa: E402

NLEV, KM2 = 128, 128


def main():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    ncol = int(os.environ.get("PROBE_COLS", 3145728))
    dt = torch.float64

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def pinned_from(t):
        h = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
        h.copy_(t)
        return h

    t0 = time.perf_counter()
    dl, d, du, B = T.fill(dt, ncol, NLEV, 1, 1, layout=ekat_b200.LEV_MAJOR, col0=rank * ncol, device=dev)
    hdl, hd, hdu, hB = (pinned_from(t[0]) for t in (dl, d, du, B))
    hX = torch.empty(hB.shape, dtype=dt, pin_memory=True)
    hX.copy_(hB)
    del dl, d, du, B
    x1, y1, x2 = li_fill(dt, ncol, NLEV, KM2, 1, col0=rank * ncol, device=dev)
    hx1, hy1, hx2 = (pinned_from(t) for t in (x1, y1, x2))
    hy2 = torch.empty((ncol, KM2), dtype=dt, pin_memory=True)
    del x1, y1, x2
    torch.cuda.synchronize()
    pin_s = time.perf_counter() - t0
    results = {"pin_seconds": pin_s}

    def report(name, up_bytes, down_bytes, sec):
        v = torch.tensor([up_bytes / sec / 1e9, down_bytes / sec / 1e9, sec * 1e3], dtype=torch.float64, device=dev)
        if world > 1:
            out = [torch.zeros_like(v) for _ in range(world)]
            dist.all_gather(out, v)
        else:
            out = [v]
        results[name] = {"up_gbs": [round(float(o[0]), 1) for o in out], "down_gbs": [round(float(o[1]), 1) for o in out],
                         "ms": [round(float(o[2]), 1) for o in out]}
        if rank == 0:
            print(name, json.dumps(results[name]), flush=True)

    # ---- flat copies of the very same pinned buffers
    piece = 32 * 1024 * 1024  # doubles: 256 MiB
    slot = [torch.zeros(piece, dtype=dt, device=dev) for _ in range(3)]
    st = [torch.cuda.Stream(device=dev) for _ in range(3)]
    srcs = [h.view(-1) for h in (hdl, hd, hdu, hB)]

    def flat(duplex):
        barrier()
        t = time.perf_counter()
        up = down = 0
        k = 0
        for s in srcs:
            for o in range(0, s.numel(), piece):
                n = min(piece, s.numel() - o)
                with torch.cuda.stream(st[k & 1]):
                    slot[k & 1][:n].copy_(s[o:o + n], non_blocking=True)
                up += n * 8
                if duplex and (k % 3) == 0:  # ~ the e2e step's 22.5 : 6.4 up : down ratio
                    with torch.cuda.stream(st[2]):
                        hX.view(-1)[o:o + n].copy_(slot[2][:n], non_blocking=True)
                    down += n * 8
                k += 1
        torch.cuda.synchronize()
        return up, down, time.perf_counter() - t

    flat(False)
    report("flat_h2d", *flat(False))
    report("flat_duplex", *flat(True))
    del slot

    # ---- the pipeline
    def pipe_case(name, nslots, dev_bytes, ne, do_tri, do_li):
        a = (hdl, hd, hdu, hX) if ne == ncol else tuple(pinned_from(t[:, :ne].contiguous()) for t in (hdl, hd, hdu, hX))
        b = (hx1, hx2, hy1, hy2) if ne == ncol else (hx1[:ne], hx2[:ne], hy1[:ne], hy2[:ne])
        pipe = Pipeline(device_bytes=dev_bytes, nslots=nslots)
        pipe.set_deferred(True)
        try:
            def one():
                barrier()
                s0 = pipe.stats()
                t = time.perf_counter()
                if do_tri:
                    pipe.tridiag("thomas", a[0], a[1], a[2], a[3], layout=ekat_b200.LEV_MAJOR)
                if do_li:
                    pipe.lin_interp(b[0], b[1], b[2], b[3], NLEV, KM2)
                pipe.sync()
                sec = time.perf_counter() - t
                s1 = pipe.stats()
                return s1["h2d_bytes"] - s0["h2d_bytes"], s1["d2h_bytes"] - s0["d2h_bytes"], sec
            one()
            report(name, *one())
        finally:
            pipe.close()

    pipe_case("tri", 3, 3 << 30, ncol, True, False)
    pipe_case("li", 3, 3 << 30, ncol, False, True)
    pipe_case("both", 3, 3 << 30, ncol, True, True)
    pipe_case("both_s6", 6, 6 << 30, ncol, True, True)
    pipe_case("both_s6_small", 6, 3 << 29, ncol, True, True)
    pipe_case("both_q", 3, 3 << 30, ncol // 4, True, True)
    if rank == 0:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", f"e2e_probe_n{world}.json"), "w") as f:
            json.dump(results, f, indent=1)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
