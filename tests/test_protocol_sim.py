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


def test_bfb_shared_kernel_protocol_random_schedules():
    """The same for k_tridiag_bfb_shared (csrc/tridiag_bfb_shared.cu): forward / backward / factor warps and three producer
    threads over two rings and mirrored plane slots, 1..7 items per CTA, ragged top chunks."""
    import sim_bfb_shared_protocol as simb
    for seed in (1, 2):
        assert simb.main(trials=120, seed=seed) == 120


def test_bfb_shared_protocol_model_notices_broken_protocols():
    """The model is only worth something if it fails on a wrong protocol: a refill that does not wait for the tensor store
    to have read its slot, a ring producer and a plane producer that do not wait for their consumers."""
    import pytest
    import sim_bfb_shared_protocol as simb
    for broken in ("refill_without_wait", "ring_without_wait", "plane_without_wait"):
        with pytest.raises(AssertionError):
            simb.main(trials=200, seed=3, break_it=broken)


def test_bfb_stream_kernel_protocol_random_schedules():
    """k_tridiag_bfb_stream (csrc/tridiag_bfb_stream.cu), two sweeping warps: io thread, input ring, two output buffers, the
    read-back of x' after the forward stores have completed, the named-barrier pair of the chain hand-over -- and the model
    fails on a read-back that does not wait for the stores and on an output buffer freed before its store has read it."""
    import pytest
    import sim_bfb_stream_protocol as sims
    for seed in (1, 2):
        assert sims.main(trials=150, seed=seed) == 150
    for broken in ("readback_without_flush", "ob_free_without_wait"):
        with pytest.raises(AssertionError):
            sims.main(trials=300, seed=3, break_it=broken)


def test_fused_kernel_protocol_random_schedules():
    """k_solve_remap_roles (csrc/fused_tma.cu, config 4 in one kernel): planes handed from tile to tile by the last remap
    warp out (atomic counters) and by the factor warp once every solver warp has arrived -- and the model fails when the
    factor warp does not wait for the solvers or when the FIRST warp out requests the next tile's plane."""
    import pytest
    import sim_fused_protocol as simf
    for seed in (1, 2):
        assert simf.main(trials=150, seed=seed) == 150
    for broken in ("factor_without_abar", "first_one_out"):
        with pytest.raises(AssertionError):
            simf.main(trials=300, seed=3, break_it=broken)
