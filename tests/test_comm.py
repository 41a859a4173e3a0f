"""CPU suite, part 3: the multi-GPU host logic on world_size-2 gloo (no GPU needed): ekat::Comm-style
decomposition, the checksum SUM / residual MAX reduction, and partition invariance of the checksum -- the only
collective step the column-sharded path has."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from helpers import u64sum


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, ncol, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import oracle
    from ekat_b200.comm import Comm, reduce_verification
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        comm = Comm()
        assert comm.rank() == rank and comm.size() == world and comm.am_i_root() == (rank == 0)
        lo, hi = comm.column_range(ncol)
        # each rank solves only its column range (CPU oracle stands in for the kernel in this host-logic test)
        orc = oracle.port()
        A, B = orc.fill_tridiag(np.float64, hi - lo, 16, 1, 2, col0=lo)
        X = orc.tridiag("bfb", A.copy(), B.copy())
        cks = torch.tensor([np.int64(np.uint64(u64sum(X)))], dtype=torch.int64)
        res = torch.tensor([orc.residual(A, B, X)], dtype=torch.float64)
        local_res = float(res.item())
        checksum, residual = reduce_verification(comm, cks, res)
        # collectives of the Comm surface
        t = torch.tensor([float(rank + 1)])
        assert comm.all_reduce(t.clone(), "sum").item() == world * (world + 1) / 2
        assert comm.all_reduce(t.clone(), "max").item() == world
        assert comm.broadcast(torch.tensor([rank]), root=1).item() == 1
        assert [int(x.item()) for x in comm.all_gather(torch.tensor([rank]))] == list(range(world))
        assert comm.scan(torch.tensor([rank + 1]), "sum").item() == (rank + 1) * (rank + 2) // 2
        comm.barrier()
        q.put((rank, lo, hi, checksum, residual, local_res))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("ncol", [101, 64])
def test_two_rank_decomposition_and_verification_reduce(ncol):
    import oracle
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, ncol, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # ranges tile [0, ncol) contiguously
    assert out[0][1] == 0 and out[0][2] == out[1][1] and out[1][2] == ncol
    # the reduced checksum equals the single-rank checksum of the whole batch (associative uint64 sum)
    orc = oracle.port()
    A, B = orc.fill_tridiag(np.float64, ncol, 16, 1, 2, col0=0)
    X = orc.tridiag("bfb", A.copy(), B.copy())
    assert out[0][3] == out[1][3] == u64sum(X)
    assert out[0][4] == out[1][4] == max(out[0][5], out[1][5]) == orc.residual(A, B, X)


def test_serial_comm_stub():
    """Without an initialised process group Comm is the 1-rank stub (ekat_comm_serial.cpp:12-28)."""
    from ekat_b200.comm import Comm
    c = Comm()
    assert c.rank() == 0 and c.size() == 1 and c.am_i_root()
    t = torch.tensor([3.0])
    assert c.all_reduce(t, "sum") is t and t.item() == 3.0
    assert c.column_range(10) == (0, 10)
    assert c.column_range(10, rank=1, size=3) == (3, 6)
    assert [c.column_range(25_165_824, r, 8) for r in (0, 7)] == [(0, 3_145_728), (22_020_096, 25_165_824)]
    c.barrier()
