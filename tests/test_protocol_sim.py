"""The mbarrier / TMA hand-off protocol of the persistent shared-matrix kernel (csrc/tridiag_tma_shared.cu), run as a
randomised-schedule simulation (tools/sim_shared_protocol.py): no deadlock, no early overwrite, no stale read, every
chunk of every right-hand side stored exactly once -- for thomas and bfb rings, 1..7 items per CTA, ragged chunks."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
import sim_shared_protocol as sim  # noqa: E402


def test_shared_kernel_protocol_random_schedules():
    for seed in (1, 2, 3):
        assert sim.main(trials=150, seed=seed) == 150
