"""ekat::Comm for the column-sharded multi-GPU driver (reference: src/core/ekat_comm.hpp:54-217).

One process per GPU; torch.distributed (NCCL on GPUs, gloo in the CPU tests) is the transport.
Columns are independent, so the data path has NO collective: Comm is used for the decomposition
(rank/size -> contiguous column range) and for one final all-reduce of a checksum (uint64 SUM,
associative => identical for any partition) and the residual norm (float64 MAX) --
Comm::all_reduce(..., MPI_SUM / MPI_MAX) in the reference (ekat_comm.hpp:152-162).
"""
import os

import torch
import torch.distributed as dist


class Comm:
    """rank()/size()/am_i_root()/barrier()/all_reduce()/broadcast()/all_gather()/scan(): the subset of
    ekat::Comm's surface (ekat_comm.hpp:78-127) that has a meaning for this path. A default-constructed Comm on
    an uninitialised process group is the serial stub (ekat_comm_serial.cpp:12-28): 1 rank, collectives are copies."""

    def __init__(self, group=None):
        self._on = dist.is_available() and dist.is_initialized()
        self._group = group

    @staticmethod
    def from_env(backend=None, device=None):
        """Initialise torch.distributed from RANK/WORLD_SIZE/MASTER_* if a launcher set them."""
        if int(os.environ.get("WORLD_SIZE", "1")) > 1 and not dist.is_initialized():
            if backend is None:
                backend = "nccl" if torch.cuda.is_available() else "gloo"
            kw = {}
            if backend == "nccl" and device is not None:
                kw["device_id"] = device
            dist.init_process_group(backend=backend, **kw)
        return Comm()

    def rank(self):
        return dist.get_rank(self._group) if self._on else 0

    def size(self):
        return dist.get_world_size(self._group) if self._on else 1

    def am_i_root(self):
        return self.rank() == 0

    def root_rank(self):
        return 0

    def barrier(self):
        if self._on:
            dist.barrier(self._group)

    def all_reduce(self, t, op="sum"):
        """In-place all-reduce of tensor t; op in {'sum','max','min'} (MPI_SUM/MPI_MAX/MPI_MIN)."""
        if self._on:
            rop = {"sum": dist.ReduceOp.SUM, "max": dist.ReduceOp.MAX, "min": dist.ReduceOp.MIN}[op]
            dist.all_reduce(t, op=rop, group=self._group)
        return t

    def broadcast(self, t, root=0):
        if self._on:
            dist.broadcast(t, src=root, group=self._group)
        return t

    def all_gather(self, t):
        if not self._on:
            return [t.clone()]
        out = [torch.empty_like(t) for _ in range(self.size())]
        dist.all_gather(out, t, group=self._group)
        return out

    def scan(self, t, op="sum"):
        """Inclusive prefix reduction over ranks (MPI_Scan, ekat_comm.hpp:139-150)."""
        parts = self.all_gather(t)
        acc = parts[0].clone()
        for r in range(1, self.rank() + 1):
            acc = acc + parts[r] if op == "sum" else (torch.maximum(acc, parts[r]) if op == "max" else torch.minimum(acc, parts[r]))
        return acc

    # ---- decomposition ---------------------------------------------------------------------
    def column_range(self, ncol, rank=None, size=None):
        """Contiguous shard [r*ncol/P, (r+1)*ncol/P) of a global column count."""
        r = self.rank() if rank is None else rank
        p = self.size() if size is None else size
        return (r * ncol) // p, ((r + 1) * ncol) // p


def reduce_verification(comm, checksum_i64, residual_f64):
    """The run's single collective step: checksum (wrap-around uint64 carried in an int64 tensor -- two's
    complement addition is the same operation) SUM and residual MAX over ranks. Returns Python (int, float)."""
    comm.all_reduce(checksum_i64, "sum")
    comm.all_reduce(residual_f64, "max")
    return int(checksum_i64.item()) & 0xFFFFFFFFFFFFFFFF, float(residual_f64.item())
