"""GPU parity tests proper: every tridiagonal entry point of the C ABI (called through ekat_b200.tridiag's ctypes
face) against the CPU oracle on identical inputs.

  bfb, cr : bit-exact against the reference arithmetic (oracle port; == oracle/_ref, see test_oracle.py)
  thomas  : max|a-b|/max|a| <= 1e-13 (f64) / 1e-5 (f32)           [north_star tolerances]
Shapes follow the reference's property test (tests/algorithm/tridiag_tests_correctness.cpp:403-484): all its nrow
values, single / shared-A multi / per-rhs-A right-hand sides, plus ragged column counts, padded pitches and
strided batches for both layouts.
"""
import os

import numpy as np
import pytest
import torch

from helpers import NROWS, a_to_lev, bits_equal, eps, rel_diff, tol, x_from_lev, x_to_lev
import ekat_b200
from ekat_b200 import tridiag as T

pytestmark = pytest.mark.gpu
DT = {"f64": (np.float64, torch.float64), "f32": (np.float32, torch.float32)}
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
REFNAME = {"thomas": "thomas", "bfb": "bfb", "cr": "cr"}


def run_gpu(solver, A, X, layout, dev, clobber=False, pad=0):
    """A[ncol,3,nrow,nA], X[ncol,nrow,nrhs] (numpy, reference layout) -> solution in the same layout (+ clobbered A)."""
    fn = getattr(T, solver)
    ncol = A.shape[0]
    if layout == ekat_b200.LEV_MAJOR:
        dl, d, du = (torch.from_numpy(np.pad(a, ((0, 0), (0, 0), (0, pad)))).to(dev) for a in a_to_lev(A))
        x = torch.from_numpy(np.pad(x_to_lev(X), ((0, 0), (0, 0), (0, pad)))).to(dev)
        fn(dl, d, du, x, layout=layout, clobber=clobber, ncol=ncol)
        torch.cuda.synchronize()
        Aout = np.stack([t.cpu().numpy()[:, :, :ncol].transpose(2, 1, 0) for t in (dl, d, du)], axis=1)
        return x_from_lev(x.cpu().numpy()[:, :, :ncol]), np.ascontiguousarray(Aout)
    At = torch.from_numpy(A).to(dev)
    Xt = torch.from_numpy(X).to(dev)
    fn(At[:, 0], At[:, 1], At[:, 2], Xt, layout=layout, clobber=clobber)  # strided slices of A[ncol,3,nrow,nA]
    torch.cuda.synchronize()
    return Xt.cpu().numpy(), At.cpu().numpy()


def check(solver, got, want, dtype, ctx):
    if solver in ("bfb", "cr"):
        assert bits_equal(got, want), f"{solver} not bit-exact: {ctx} rel={rel_diff(want, got):.3e}"
    else:
        assert rel_diff(want, got) <= tol(dtype), f"thomas off by {rel_diff(want, got):.3e}: {ctx}"


@pytest.mark.parametrize("dt", ["f64", "f32"])
@pytest.mark.parametrize("solver", ["thomas", "bfb", "cr"])
@pytest.mark.parametrize("layout", [ekat_b200.LEV_MAJOR, ekat_b200.COL_MAJOR])
def test_parity_reference_shapes(cuda, port, dt, solver, layout):
    npdt, _ = DT[dt]
    for nrow in NROWS:
        for nrhs, nA in [(1, 1), (5, 1), (5, 5), (13, 1)]:
            ncol = 37  # ragged against every tile width
            A, B = port.fill_tridiag(npdt, ncol, nrow, nA, nrhs, col0=3)
            want = port.tridiag(REFNAME[solver], A.copy(), B.copy())
            got, Aout = run_gpu(solver, A, B, layout, cuda, pad=3 if layout == ekat_b200.LEV_MAJOR else 0)
            check(solver, got, want, npdt, (dt, layout, nrow, nrhs, nA))
            assert bits_equal(Aout, A), "dl/d/du must be untouched without CLOBBER_DIAGS"
            assert port.residual(A, B, got) <= 50 * eps(npdt)  # the reference's own acceptance test


@pytest.mark.parametrize("dt", ["f64", "f32"])
@pytest.mark.parametrize("layout", [ekat_b200.LEV_MAJOR, ekat_b200.COL_MAJOR])
def test_long_columns_2048(cuda, port, dt, layout):
    """nrow = 2048 (the reference's NDEBUG-only case, :410-414): few columns, many levels."""
    npdt, _ = DT[dt]
    for solver in ("thomas", "bfb", "cr"):
        for nrhs, nA in [(1, 1), (3, 1)]:
            A, B = port.fill_tridiag(npdt, 5, 2048, nA, nrhs, col0=1)
            want = port.tridiag(solver, A.copy(), B.copy())
            got, _ = run_gpu(solver, A, B, layout, cuda)
            check(solver, got, want, npdt, (dt, layout, solver, nrhs))


@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_golden_fixture(cuda, dt):
    """Against the committed outputs of the unchanged reference headers (tests/golden, oracle/gen_golden.py)."""
    npdt, tdt = DT[dt]
    g = np.load(os.path.join(GOLD, "tridiag.npz"), allow_pickle=True)
    for gdt, nrow, nrhs, nA, ncol, col0 in g["meta"]:
        if gdt != dt:
            continue
        for layout in (ekat_b200.LEV_MAJOR, ekat_b200.COL_MAJOR):
            dl, d, du, X = T.fill(tdt, ncol, nrow, nA, nrhs, layout=layout, col0=col0, device=cuda)
            for solver in ("bfb", "cr", "thomas"):
                x = X.clone()
                getattr(T, solver)(dl, d, du, x, layout=layout)
                got = x.cpu().numpy()
                got = x_from_lev(got) if layout == ekat_b200.LEV_MAJOR else got
                check(solver, got, g[f"{dt}_{solver}_n{nrow}_r{nrhs}_a{nA}"], npdt, ("golden", layout, nrow, nrhs, nA))


@pytest.mark.parametrize("dt", ["f64", "f32"])
@pytest.mark.parametrize("layout", [ekat_b200.LEV_MAJOR, ekat_b200.COL_MAJOR])
def test_clobber_diags_matches_reference_side_effect(cuda, port, dt, layout):
    """With CLOBBER_DIAGS bfb leaves multipliers in dl and pivots in d, bit for bit as the reference does
    (ekat_tridiag.hpp:299-302); du is unchanged."""
    npdt, _ = DT[dt]
    for nrhs, nA in [(1, 1), (4, 1), (4, 4)]:
        A, B = port.fill_tridiag(npdt, 45, 72, nA, nrhs, col0=9)
        Aref, Xref = A.copy(), B.copy()
        port.tridiag("bfb", Aref, Xref)
        got, Aout = run_gpu("bfb", A, B, layout, cuda, clobber=True)
        assert bits_equal(got, Xref)
        assert bits_equal(Aout[:, 0, 1:], Aref[:, 0, 1:]), "dl multipliers"
        assert bits_equal(Aout[:, 1], Aref[:, 1]), "d pivots"
        assert bits_equal(Aout[:, 2], A[:, 2]), "du untouched"
        # thomas with clobber: pivots agree to tolerance with the serial reference (:221)
        Aref2, Xref2 = A.copy(), B.copy()
        port.tridiag("thomas", Aref2, Xref2)
        got2, Aout2 = run_gpu("thomas", A, B, layout, cuda, clobber=True)
        assert rel_diff(Xref2, got2) <= tol(npdt)
        assert rel_diff(Aref2[:, 1], Aout2[:, 1]) <= tol(npdt), "thomas: pivots in d (:221), both layouts"
        assert bits_equal(Aout2[:, 2], A[:, 2]), "thomas never modifies du"


@pytest.mark.parametrize("dt", ["f64", "f32"])
@pytest.mark.parametrize("solver", ["bfb", "thomas"])
def test_clobber_diags_many_rhs_groups(cuda, port, dt, solver):
    """CLOBBER_DIAGS with more right-hand sides than one CTA holds (several rhs groups per column tile, e.g. the
    reference perf driver's default 128 rows x 43 rhs, tridiag_tests_performance.cpp:33-37): the matrix must be
    factorised from the ORIGINAL dl/d exactly once per column whatever the number of groups -- X and the
    clobbered dl/d bit-identical to the reference (bfb) / within tolerance (thomas), many columns."""
    npdt, _ = DT[dt]
    for nrhs, nrow, ncol in [(12, 128, 1500), (43, 128, 700), (24, 72, 1111)]:
        A, B = port.fill_tridiag(npdt, ncol, nrow, 1, nrhs, col0=5)
        Aref, Xref = A.copy(), B.copy()
        port.tridiag("bfb", Aref, Xref)
        got, Aout = run_gpu(solver, A, B, ekat_b200.LEV_MAJOR, cuda, clobber=True)
        if solver == "bfb":
            assert bits_equal(got, Xref), (nrhs, nrow, rel_diff(Xref, got))
            assert bits_equal(Aout[:, 0, 1:], Aref[:, 0, 1:]), "dl multipliers"
            assert bits_equal(Aout[:, 1], Aref[:, 1]), "d pivots"
        else:
            assert rel_diff(Xref, got) <= tol(npdt), (nrhs, nrow, rel_diff(Xref, got))
            assert rel_diff(Aref[:, 1], Aout[:, 1]) <= tol(npdt)
        assert bits_equal(Aout[:, 2], A[:, 2]), "du untouched"


def test_against_the_unchanged_reference_headers(cuda, ref):
    """The GPU path against oracle/_ref itself (the reference's own headers compiled where they lie), not only
    against the plain-C port: bfb and cr bit for bit, thomas to tolerance, single / shared / per-rhs matrices."""
    if ref is None:
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    for npdt in (np.float64, np.float32):
        for nrow, nrhs, nA in [(72, 1, 1), (128, 10, 1), (128, 10, 10), (129, 3, 1)]:
            A, B = ref.fill_tridiag(npdt, 333, nrow, nA, nrhs, col0=7)
            for solver in ("bfb", "cr", "thomas"):
                want = ref.tridiag(solver, A.copy(), B.copy())
                for layout in (ekat_b200.LEV_MAJOR, ekat_b200.COL_MAJOR):
                    got, _ = run_gpu(solver, A, B, layout, cuda, pad=(-333) % 4 if layout == ekat_b200.LEV_MAJOR else 0)
                    check(solver, got, want, npdt, ("_ref", solver, layout, nrow, nrhs, nA))


def test_tuning_variants_agree(cuda, port):
    """Every launch geometry / prefetch depth gives the same bits (bfb) -- results are independent of the
    'team' shape, as the reference's are of its (thread, vector) parameterisation (:375-390)."""
    from ekat_b200._lib import lib
    A, B = port.fill_tridiag(np.float64, 333, 72, 1, 1, col0=0)
    want = port.tridiag("bfb", A.copy(), B.copy())
    try:
        for cols in (8, 16, 32, 64, 96):
            for pf in (4, 8, 16):
                lib().ekat_b200_set_tuning(cols, pf, 0)
                got, _ = run_gpu("bfb", A, B, ekat_b200.LEV_MAJOR, cuda)
                assert bits_equal(got, want), (cols, pf)
        for nA, nrhs in [(1, 1), (1, 3), (3, 3)]:  # scalar (unvectorised) cp.async path == vector path
            A2, B2 = port.fill_tridiag(np.float64, 97, 33, nA, nrhs, col0=1)
            lib().ekat_b200_set_tuning(0, 0, 3)
            got, _ = run_gpu("bfb", A2, B2, ekat_b200.LEV_MAJOR, cuda)
            assert bits_equal(got, port.tridiag("bfb", A2.copy(), B2.copy()))
        lib().ekat_b200_set_tuning(0, 0, 2)  # force the tile kernel on level-major data
        got, _ = run_gpu("bfb", A, B, ekat_b200.LEV_MAJOR, cuda)
        assert bits_equal(got, want)
        for cols in (1, 3, 8, 21):
            lib().ekat_b200_set_tuning(cols, 0, 0)
            for layout in (ekat_b200.LEV_MAJOR, ekat_b200.COL_MAJOR):
                got, _ = run_gpu("cr", A, B, layout, cuda)
                assert bits_equal(got, port.tridiag("cr", A.copy(), B.copy())), cols
    finally:
        lib().ekat_b200_set_tuning(0, 0, 0)


@pytest.mark.parametrize("dt", ["f64", "f32"])
@pytest.mark.parametrize("solver", ["bfb", "cr"])
def test_signed_zeros_bit_exact(cuda, port, dt, solver):
    """+0 / -0 sub- and super-diagonal entries and right-hand sides: the quotients 0/d' and the products m*0 carry the
    sign of the zero, and the speculative divisions must hand every case their fast path is not proven for (a -0
    numerator ends on it with the wrong zero) to the full-range division. Bit for bit on every route, also the
    multipliers and pivots written back under CLOBBER_DIAGS."""
    npdt, _ = DT[dt]
    rng = np.random.default_rng(5)
    for layout, nrow, nrhs, nA, ncol, clobber in [(ekat_b200.LEV_MAJOR, 128, 1, 1, 5000, False), (ekat_b200.LEV_MAJOR, 72, 1, 1, 700, False),
                                                  (ekat_b200.LEV_MAJOR, 128, 10, 1, 4800, False), (ekat_b200.LEV_MAJOR, 72, 10, 1, 4800, False),
                                                  (ekat_b200.LEV_MAJOR, 128, 4, 4, 2000, True), (ekat_b200.COL_MAJOR, 128, 1, 1, 3000, False),
                                                  (ekat_b200.COL_MAJOR, 128, 10, 1, 2048, False), (ekat_b200.COL_MAJOR, 64, 2, 2, 999, True)]:
        if solver == "cr" and (clobber or nA > 1):
            continue
        A, B = port.fill_tridiag(npdt, ncol, nrow, nA, nrhs, col0=9)
        for arr, k in ((A[:, 0], 0.15), (A[:, 2], 0.15), (B, 0.2)):  # dl, du, right-hand sides
            m = rng.random(arr.shape)
            arr[m < k / 2] = 0.0
            arr[(m >= k / 2) & (m < k)] = -0.0
        B[1] = -0.0  # a whole column of -0 right-hand sides, one of +0
        B[2] = 0.0
        Aref, Xref = A.copy(), B.copy()
        port.tridiag(REFNAME[solver], Aref, Xref)
        got, Aout = run_gpu(solver, A, B, layout, cuda, clobber=clobber, pad=(-ncol) % 4 if layout == ekat_b200.LEV_MAJOR else 0)
        assert bits_equal(got, Xref), (solver, dt, layout, nrow, nrhs, nA, lib_last())
        if clobber:
            assert bits_equal(Aout[:, 0, 1:], Aref[:, 0, 1:]) and bits_equal(Aout[:, 1], Aref[:, 1]), (solver, dt, layout, nrow)


def lib_last():
    from ekat_b200._lib import lib
    return lib().ekat_b200_last_kernel().decode()


@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_pipelined_division_is_ieee(cuda, dt):
    """The divide used on the exact paths (reciprocal hoisted off the chain) equals __ddiv_rn / __fdiv_rn bit for
    bit on 2^28 random bit patterns (all exponents, denormals, infinities) and 2^28 well-scaled operands."""
    from ekat_b200._lib import lib, check
    out = torch.zeros(1, dtype=torch.int64, device=cuda)
    fn = getattr(lib(), f"ekat_b200_selftest_div_{dt}")
    for mode in (0, 1):
        check("selftest_div", fn(1 << 28, 12345 + mode, mode, out.data_ptr(), None))
    torch.cuda.synchronize()
    assert int(out.item()) == 0
    # the certified quotient of the bfb factorisation (reciprocal of a PREDICTED pivot, common.cuh bfb_pivot_step): right
    # whenever its certificate does not send the chunk to the full-range division, for reciprocals up to 2^-22 off
    for mode in (2, 3):
        check("selftest_div", fn(1 << 27, 777 + mode, mode, out.data_ptr(), None))
    torch.cuda.synchronize()
    assert int(out.item()) == 0
    # ... and with a prediction as close as the recurrence's (2^-52), well-scaled operands (almost) never need the redo
    check("selftest_div", fn(1 << 26, 4242, 7, out.data_ptr(), None))
    torch.cuda.synchronize()
    assert int(out.item()) <= 64, int(out.item())  # of 2^27 quotients (f32: MUFU.RCP's last bit; f64: none seen)


def test_odd_pitch_and_unaligned_views_use_the_scalar_path(cuda, port):
    """Pitches / base pointers that are not 16-byte multiples cannot use 16-byte cp.async: same bits through the
    8-byte path."""
    A, B = port.fill_tridiag(np.float64, 37, 72, 1, 1, col0=2)
    want = port.tridiag("bfb", A.copy(), B.copy())
    for pad in (0, 2, 3):  # pitch 37 (odd), 39 (odd), 40 (even)
        got, _ = run_gpu("bfb", A, B, ekat_b200.LEV_MAJOR, cuda, pad=pad)
        assert bits_equal(got, want), pad
    dl, d, du = (torch.from_numpy(np.pad(a, ((0, 0), (0, 0), (1, 2)))).to(cuda)[:, :, 1:] for a in a_to_lev(A))
    x = torch.from_numpy(np.pad(x_to_lev(B), ((0, 0), (0, 0), (1, 2)))).to(cuda)[:, :, 1:]  # base pointer off by 8 bytes
    T.bfb(dl, d, du, x, ncol=37)
    assert bits_equal(x_from_lev(x.cpu().numpy()[:, :, :37]), want)


def test_edge_cases(cuda, port):
    dev = cuda
    # ncol == 0 is a silent no-op; shapes are still validated
    e = torch.empty((72, 0), dtype=torch.float64, device=dev)
    T.bfb(e, e, e, e.clone())
    # one column, one level
    A, B = port.fill_tridiag(np.float64, 1, 1, 1, 1)
    for s in ("thomas", "bfb", "cr"):
        for layout in (ekat_b200.LEV_MAJOR, ekat_b200.COL_MAJOR):
            got, _ = run_gpu(s, A, B, layout, dev)
            if s == "thomas":
                assert abs(got[0, 0, 0] - B[0, 0, 0] / A[0, 1, 0, 0]) <= 4 * eps(np.float64) * abs(got[0, 0, 0])
            else:
                assert got[0, 0, 0] == B[0, 0, 0] / A[0, 1, 0, 0]
    # wrong shapes raise before any launch
    a = torch.zeros((72, 8), dtype=torch.float64, device=dev)
    with pytest.raises(ValueError):
        T.bfb(a, a, a, torch.zeros((71, 8), dtype=torch.float64, device=dev))
    with pytest.raises(ValueError):
        T.bfb(a, a, a, torch.zeros((72, 8), dtype=torch.float32, device=dev))
    with pytest.raises(ValueError):
        T.bfb(a.cpu(), a.cpu(), a.cpu(), a.cpu())  # no CPU path
    # a column too long for on-chip staging is reported, not silently mis-solved
    big = torch.ones((3, 40000), dtype=torch.float64, device=dev)
    with pytest.raises(ekat_b200.EkatB200Error) as ei:
        T.cr(big, big, big, big.clone(), layout=ekat_b200.COL_MAJOR)
    assert ei.value.rc == -4


@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_full_size_configs_properties(cuda, port, dt):
    """BASELINE sizes, checked through size-independent properties: the reference's residual test on every
    column (device reduction), bit-equality of a sampled slab with the oracle, and partition-invariance of the
    checksum (sum of per-shard checksums == whole-batch checksum)."""
    npdt, tdt = DT[dt]
    for ncol, nlev in [(21600, 72), (345600, 128)]:  # C1 and C3
        dl, d, du, B = T.fill(tdt, ncol, nlev, 1, 1, device=cuda)
        for solver in ("thomas", "bfb", "cr"):
            X = B.clone()
            getattr(T, solver)(dl, d, du, X)
            res = float(T.residual(dl, d, du, B, X).item())
            assert res <= 50 * eps(npdt), (solver, ncol, res)
            # sampled slab vs oracle (columns straddling the 32768 seed wrap for the big case)
            c0 = min(ncol - 64, 32768 - 32)
            A_h, B_h = port.fill_tridiag(npdt, 64, nlev, 1, 1, col0=c0)
            want = port.tridiag(solver, A_h.copy(), B_h.copy())
            got = x_from_lev(X[:, :, c0:c0 + 64].cpu().numpy())
            check(solver, got, want, npdt, (solver, ncol))
            # checksum is associative over any column partition
            whole = int(T.checksum(X).item())
            parts = torch.zeros(1, dtype=torch.int64, device=cuda)
            for lo, hi in [(0, ncol // 3), (ncol // 3, ncol // 2 + 7), (ncol // 2 + 7, ncol)]:
                T.checksum(X[:, :, lo:hi], out=parts)
            assert whole == int(parts.item())


@pytest.mark.parametrize("dt", ["f64", "f32"])
@pytest.mark.parametrize("solver", ["thomas", "bfb"])
def test_shared_matrix_persistent_many_items(cuda, port, dt, solver):
    """One matrix, many right-hand sides (BASELINE config 4's shape) at a column count that gives every persistent
    CTA of the TMA kernel several work items, so the cross-item prefetch (head slots, ring wrap-around, refills
    behind the tensor stores) is exercised; plus ragged level counts and right-hand-side counts with 1..3 groups."""
    npdt, _ = DT[dt]
    W = 32 if dt == "f64" else 64
    for nrow, nrhs, ncol in [(128, 10, 148 * W * 3 + 5), (72, 10, 148 * W * 2 + W + 1), (43, 4, 148 * W * 4 + 7),
                             (129, 6, 148 * W * 2 + 3), (5, 3, 148 * W * 5 + 9), (8, 2, 148 * W * 3)]:
        A, B = port.fill_tridiag(npdt, ncol, nrow, 1, nrhs, col0=11)
        want = port.tridiag(solver, A.copy(), B.copy())
        from ekat_b200._lib import lib
        # default route (bfb: the split forward / backward kernel where it applies) and the persistent kernel with one
        # warp per right-hand side pinned (variant 9)
        for variant in (0, 9):
            try:
                lib().ekat_b200_set_tuning(0, 0, variant)
                got, Aout = run_gpu(solver, A, B, ekat_b200.LEV_MAJOR, cuda, pad=(-ncol) % 4)  # TMA needs 16-byte pitches
            finally:
                lib().ekat_b200_set_tuning(0, 0, 0)
            name = lib().ekat_b200_last_kernel().decode()
            assert name.startswith("tridiag_tma_shared" if variant == 9 or solver == "thomas" else ("tridiag_bfb_shared", "tridiag_tma_shared")), name
            check(solver, got, want, npdt, (dt, nrow, nrhs, ncol, variant))
            assert bits_equal(Aout, A)


@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_cr_many_columns_bit_exact(cuda, port, dt):
    """Cyclic reduction at column counts that take the TMA kernels: bit-identical to the reference's CR for
    power-of-two, odd and tiny level counts, ragged last tile (tridiag_cr_tma.cu; tuning variant 14 pins it too)."""
    npdt, _ = DT[dt]
    from ekat_b200._lib import lib
    cases = [(128, 9472 + 5), (72, 9600), (43, 9473), (129, 9480), (5, 9999), (2, 9500), (1, 9476), (3, 9481), (65, 9700)]
    for variant in (0, 14):
        try:
            lib().ekat_b200_set_tuning(0, 0, variant)
            for nrow, ncol in cases:
                A, B = port.fill_tridiag(npdt, ncol, nrow, 1, 1, col0=2)
                want = port.tridiag("cr", A.copy(), B.copy())
                got, Aout = run_gpu("cr", A, B, ekat_b200.LEV_MAJOR, cuda, pad=(-ncol) % 4)
                name = lib().ekat_b200_last_kernel().decode()
                assert name.startswith("tridiag_cr_tma"), name
                assert bits_equal(got, want), (dt, variant, nrow, ncol, rel_diff(want, got))
                assert bits_equal(Aout, A)
        finally:
            lib().ekat_b200_set_tuning(0, 0, 0)


@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_cr_every_level_count(cuda, port, dt):
    """tridiag_cr_tma.cu at every level count 1..128 (every number of levels of the reduction, missing right
    neighbours, 1x1 / 2x2 bottom solve), ragged last tiles, and strongly dominant matrices whose reduced
    off-diagonals underflow towards zero within a few levels."""
    npdt, _ = DT[dt]
    from ekat_b200._lib import lib
    for nrow in range(1, 129):
        ncol = 9472 + 16 * (nrow % 3) + nrow % 7
        A, B = port.fill_tridiag(npdt, ncol, nrow, 1, 1, col0=nrow)
        if nrow % 5 == 0:  # off-diagonals 1e-3 of the diagonal
            A[:, 0] *= 1e-3
            A[:, 2] *= 1e-3
        want = port.tridiag("cr", A.copy(), B.copy())
        got, Aout = run_gpu("cr", A, B, ekat_b200.LEV_MAJOR, cuda, pad=(-ncol) % 4)
        assert lib().ekat_b200_last_kernel().decode().startswith("tridiag_cr_tma"), lib().ekat_b200_last_kernel()
        assert bits_equal(got, want), (dt, nrow, ncol, rel_diff(want, got))
        assert bits_equal(Aout, A)


@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_small_batch_kernel_two_sided_elimination(cuda, port, dt):
    """tridiag_small.cu (whole tile on chip, eliminated from both ends, divisions deferred to the substitution):
    forced through the tuning variant at every reference level count (1..129: no bottom half, partial top chunks,
    odd splits), ragged tiles, several matrices per launch, and at BASELINE config 1's own shape where it is the
    default -- against the oracle to the thomas tolerance, plus the reference's residual acceptance test."""
    from ekat_b200._lib import lib
    npdt, _ = DT[dt]
    try:
        lib().ekat_b200_set_tuning(0, 0, 13)
        for nrow in NROWS + [7, 9, 17, 24, 25, 48, 71, 72, 73, 96, 200]:
            for nrhs, nA, ncol in [(1, 1, 37), (3, 3, 70), (1, 1, 64)]:
                A, B = port.fill_tridiag(npdt, ncol, nrow, nA, nrhs, col0=13)
                want = port.tridiag("thomas", A.copy(), B.copy())
                got, Aout = run_gpu("thomas", A, B, ekat_b200.LEV_MAJOR, cuda, pad=(-ncol) % 4)
                assert lib().ekat_b200_last_kernel().decode().startswith("tridiag_small"), lib().ekat_b200_last_kernel()
                assert rel_diff(want, got) <= tol(npdt), (dt, nrow, nrhs, rel_diff(want, got))
                assert bits_equal(Aout, A)
                assert port.residual(A, B, got) <= 50 * eps(npdt)
    finally:
        lib().ekat_b200_set_tuning(0, 0, 0)
    A, B = port.fill_tridiag(npdt, 21600, 72, 1, 1, col0=0)  # config 1: the default route
    want = port.tridiag("thomas", A.copy(), B.copy())
    got, _ = run_gpu("thomas", A, B, ekat_b200.LEV_MAJOR, cuda)
    assert lib().ekat_b200_last_kernel().decode().startswith("tridiag_small"), lib().ekat_b200_last_kernel()
    assert rel_diff(want, got) <= tol(npdt)
    # dl(0) and du(n-1) are never used (ekat_tridiag.hpp:24-27): NaNs there must not reach the solution
    A2 = A.copy()
    A2[:, 0, 0, :] = np.nan
    A2[:, 2, -1, :] = np.nan
    got2, _ = run_gpu("thomas", A2, B, ekat_b200.LEV_MAJOR, cuda)
    assert bits_equal(got2, got)


@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_reference_layout_many_rhs(cuda, port, dt):
    """tridiag_cm_rhs.cu: the reference layout A[col][3][lev][nA], X[col][lev][nrhs] with several right-hand sides
    (BASELINE config 4 as a reference caller holds it; the reference perf driver's default 128 x 43,
    tests/algorithm/tridiag_tests_performance.cpp:33-37), one matrix per column and one per right-hand side, ragged
    last tiles, tile widths forced through the tuning knob, and the in-place side effects -- bfb bit for bit, thomas to
    the tolerance, dl / d as the reference leaves them."""
    from ekat_b200._lib import lib
    npdt, _ = DT[dt]
    cases = [(128, 10, 1, 601), (128, 10, 10, 77), (72, 4, 1, 1000), (72, 4, 4, 333), (128, 43, 1, 70), (16, 2, 1, 4999),
             (8, 32, 1, 130), (100, 6, 6, 41), (2, 2, 1, 50), (4, 1, 1, 300)]
    for nrow, nrhs, nA, ncol in cases:
        A, B = port.fill_tridiag(npdt, ncol, nrow, nA, nrhs, col0=5)
        for solver in ("bfb", "thomas"):
            want = port.tridiag(solver, A.copy(), B.copy())
            for clobber in (False, True):
                if nrhs == 1 and not clobber:
                    continue  # one right-hand side without side effects belongs to the TMA tensor kernels
                for cols in (0, 1, 3):
                    try:
                        lib().ekat_b200_set_tuning(cols, 0, 0)
                        got, Aout = run_gpu(solver, A, B, ekat_b200.COL_MAJOR, cuda, clobber=clobber)
                    finally:
                        lib().ekat_b200_set_tuning(0, 0, 0)
                    name = lib().ekat_b200_last_kernel().decode()
                    row_bytes = nrow * nA * np.dtype(npdt).itemsize, nrow * nrhs * np.dtype(npdt).itemsize
                    if all(b % 16 == 0 for b in row_bytes):  # bulk copies move multiples of 16 bytes; other rows take the tile kernel
                        assert name.startswith("tridiag_cm_rhs"), (name, nrow, nrhs, nA)
                    check(solver, got, want, npdt, (dt, nrow, nrhs, nA, ncol, clobber))
                    if not clobber:
                        assert bits_equal(Aout, A)
                    else:
                        # what the reference's in-place solvers leave: bfb multipliers in dl(1..) and pivots in d
                        # (:292-303), serial thomas pivots in d only (:216-227); du never changes
                        Aref = A.copy()
                        port.tridiag("bfb", Aref, B.copy())
                        assert bits_equal(Aout[:, 2], A[:, 2])
                        assert bits_equal(Aout[:, 1], Aref[:, 1])
                        assert bits_equal(Aout[:, 0, 1:], Aref[:, 0, 1:] if solver == "bfb" else A[:, 0, 1:])
                        assert bits_equal(Aout[:, 0, 0], A[:, 0, 0])  # dl(0) is never touched


@pytest.mark.parametrize("dt", ["f64", "f32"])
def test_clobber_level_major_stream(cuda, port, dt):
    """EKAT_B200_CLOBBER_DIAGS on level-major operands with one matrix per right-hand side takes the input-only ring
    kernel (tridiag_bfb_stream.cu), which stores the factorisation next to the solution: multipliers in dl(1..) and
    pivots in d bit for bit as the reference's bfb leaves them (ekat_tridiag.hpp:292-303), dl(0) and du untouched;
    thomas leaves the pivots in d (serial form, :216-227). Ragged top chunks, short and long columns, several
    matrices per column, ragged last tiles."""
    from ekat_b200._lib import lib
    npdt, _ = DT[dt]
    for nrow, nrhs, ncol in [(128, 1, 4099), (72, 1, 1000), (43, 1, 333), (129, 1, 700), (100, 3, 257), (8, 1, 64), (5, 2, 100),
                             (1, 1, 40), (2, 1, 40), (200, 1, 130)]:
        A, B = port.fill_tridiag(npdt, ncol, nrow, nrhs, nrhs, col0=3)
        Aref, Xref = A.copy(), B.copy()
        port.tridiag("bfb", Aref, Xref)
        for solver in ("bfb", "thomas"):
            got, Aout = run_gpu(solver, A, B, ekat_b200.LEV_MAJOR, cuda, clobber=True, pad=(-ncol) % 4)
            name = lib().ekat_b200_last_kernel().decode()
            assert name.startswith("tridiag_bfb_stream"), (name, nrow, nrhs)
            assert bits_equal(got, Xref), (solver, nrow, nrhs, ncol)  # thomas runs the bfb sweep here
            assert bits_equal(Aout[:, 2], A[:, 2]), "du untouched"
            assert bits_equal(Aout[:, 1], Aref[:, 1]), "pivots in d"
            assert bits_equal(Aout[:, 0, 0], A[:, 0, 0]), "dl(0) untouched"
            if solver == "bfb":
                assert bits_equal(Aout[:, 0, 1:], Aref[:, 0, 1:]), "multipliers in dl"
            else:
                assert bits_equal(Aout[:, 0], A[:, 0]), "serial thomas leaves dl alone"
