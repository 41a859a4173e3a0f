"""Small launches of every kernel family, for compute-sanitizer (memcheck / initcheck / synccheck / racecheck):

    compute-sanitizer --tool memcheck python tools/sanitize_cases.py [family ...]

families: shared bfb_stream spill tma cr tile levmajor lin_interp lin_interp_lm fused reduction util (default: all).
Shapes are small enough for the sanitizer's 10-100x slowdown but still give the persistent kernels several work
items per CTA, ragged last tiles and partial top chunks. Every result is also compared with the CPU oracle, so a
sanitizer-clean run is a parity run too (mirrors the reference's cuda-memcheck / valgrind test option,
cmake/EkatCreateUnitTest.cmake:314-325)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import ekat_b200  # noqa: E402
import oracle  # noqa: E402
from ekat_b200 import synth, tridiag as T  # noqa: E402
from ekat_b200._lib import lib  # noqa: E402
from ekat_b200.lin_interp import LinInterp, lin_interp_lm, tridiag_lin_interp  # noqa: E402
from helpers import a_to_lev, bits_equal, rel_diff, x_from_lev, x_to_lev  # noqa: E402

dev = torch.device("cuda:0")
orc = oracle.port()
LM, CMJ = ekat_b200.LEV_MAJOR, ekat_b200.COL_MAJOR
SMS = int(os.environ.get("SAN_SMS", "148"))  # persistent grids are sized by this many SMs


def solve(solver, ncol, nrow, nA, nrhs, layout, dt=np.float64, clobber=False, expect=None):
    A, B = orc.fill_tridiag(dt, ncol, nrow, nA, nrhs, col0=3)
    want = orc.tridiag(solver, A.copy(), B.copy())
    if layout == LM:
        pad = (-ncol) % 4
        dl, d, du = (torch.from_numpy(np.pad(a, ((0, 0), (0, 0), (0, pad)))).to(dev) for a in a_to_lev(A))
        x = torch.from_numpy(np.pad(x_to_lev(B), ((0, 0), (0, 0), (0, pad)))).to(dev)
        getattr(T, solver)(dl, d, du, x, layout=layout, clobber=clobber, ncol=ncol)
        torch.cuda.synchronize()
        got = x_from_lev(x.cpu().numpy()[:, :, :ncol])
    else:
        At, Xt = torch.from_numpy(A).to(dev), torch.from_numpy(B).to(dev)
        getattr(T, solver)(At[:, 0], At[:, 1], At[:, 2], Xt, layout=layout, clobber=clobber)
        torch.cuda.synchronize()
        got = Xt.cpu().numpy()
    k = lib().ekat_b200_last_kernel().decode()
    if expect:
        assert k.startswith(expect), (k, expect)
    if solver == "thomas":
        assert rel_diff(want, got) <= (1e-13 if dt == np.float64 else 1e-5), (k, rel_diff(want, got))
    else:
        assert bits_equal(got, want), (k, solver, ncol, nrow, nrhs)
    print(f"  ok {k:40s} {solver} {ncol} x {nrow} nA={nA} nrhs={nrhs} layout={layout} {np.dtype(dt).name}", flush=True)


def fam_shared():
    for dt, W in ((np.float64, 32), (np.float32, 64)):
        solve("thomas", SMS * W * 2 + 5, 128, 1, 1, LM, dt, expect="tridiag_tma_shared" if dt == np.float64 else None)
        solve("thomas", SMS * W * 2 + 5, 72, 1, 10, LM, dt, expect="tridiag_tma_shared")
        solve("bfb", SMS * W * 2 + 33, 43, 1, 4, LM, dt, expect="tridiag_tma_shared")
    solve("thomas", SMS * 32 * 2 + 7, 128, 1, 1, CMJ, expect="tridiag_tma_shared_cm")


def fam_bfb_stream():
    solve("bfb", SMS * 32 * 4 + 9, 128, 1, 1, LM, expect="tridiag_bfb_stream")
    solve("bfb", SMS * 32 * 2 + 9, 129, 1, 1, LM)
    solve("bfb", SMS * 32 * 2 + 3, 128, 1, 1, CMJ)
    solve("bfb", SMS * 64 * 2 + 9, 128, 1, 1, LM, np.float32)


def fam_spill():
    solve("bfb", SMS * 32 * 4 + 9, 72, 1, 1, LM)
    solve("bfb", SMS * 32 * 2 + 1, 43, 1, 1, CMJ)
    solve("bfb", SMS * 64 * 2 + 9, 72, 1, 1, LM, np.float32)


def fam_tma():
    solve("thomas", 4000, 128, 4, 4, LM)
    solve("bfb", 4000, 65, 3, 3, LM)
    solve("thomas", 9000, 128, 1, 1, LM, np.float32)
    solve("thomas", 4000, 72, 1, 1, CMJ, np.float32)


def fam_cr():
    solve("cr", 9472 + 5, 128, 1, 1, LM, expect="tridiag_cr")
    solve("cr", 9481, 43, 1, 1, LM, np.float32, expect="tridiag_cr")
    solve("cr", 9473, 3, 1, 1, LM)
    solve("cr", 64, 2048, 1, 1, LM)
    solve("cr", 300, 128, 1, 5, LM)


def fam_tile():
    solve("cr", 300, 128, 1, 5, CMJ)
    solve("cr", 70, 65, 3, 3, CMJ)
    solve("thomas", 500, 128, 1, 10, CMJ)
    solve("bfb", 500, 72, 4, 4, CMJ)
    solve("bfb", 37, 2048, 1, 1, CMJ, np.float32)
    solve("bfb", 333, 72, 1, 4, CMJ, clobber=True)


def fam_levmajor():
    solve("bfb", 1500, 128, 1, 12, LM, clobber=True)
    solve("thomas", 700, 128, 1, 43, LM, clobber=True)
    solve("bfb", 999, 72, 1, 1, LM, clobber=True)
    solve("bfb", 333, 33, 3, 3, LM, clobber=True)
    lib().ekat_b200_set_tuning(0, 0, 3)  # scalar cp.async path
    try:
        solve("bfb", 97, 33, 1, 3, LM)
    finally:
        lib().ekat_b200_set_tuning(0, 0, 0)


def li_case(km1, km2, N, ncol, dt):
    tdt = torch.float64 if dt == np.float64 else torch.float32
    x1, y1, x2 = synth.lin_interp_inputs(dt, ncol, km1, km2, N, col0=17)
    y1f = np.stack([y1, 2 * y1 + 1]).astype(dt)
    want, wim = orc.lin_interp(x1, x2, y1f, km1, km2, N)
    li = LinInterp(ncol, km1, km2, tdt, N, device=dev)
    tx1, tx2, ty1 = (torch.from_numpy(a).to(dev) for a in (x1, x2, y1f))
    for mode in ("fused", "split"):
        ty2 = torch.zeros(ty1.shape[:-1] + (li.pitch2,), dtype=tdt, device=dev)
        if mode == "fused":
            li.run(tx1, tx2, ty1, ty2, write_map=True)
        else:
            li.setup(tx1, tx2)
            li.lin_interp(tx1, tx2, ty1, ty2)
        torch.cuda.synchronize()
        assert np.array_equal(li.indx_map.cpu().numpy(), wim)
        assert bits_equal(ty2.cpu().numpy()[..., :km2], want[..., :km2])
    print(f"  ok {lib().ekat_b200_last_kernel().decode():40s} LinInterp {km1}->{km2} N={N} {ncol} cols {np.dtype(dt).name}", flush=True)


def fam_lin_interp():
    for dt in (np.float64, np.float32):
        li_case(72, 128, 16, SMS * 8 * 3 + 5, dt)
        li_case(128, 128, 1, SMS * 8 * 3 + 1, dt)
        li_case(33, 70, 4, 1300, dt)
        li_case(257, 300, 2, 119, dt)


def fam_lin_interp_lm():
    for dt, tdt in ((np.float64, torch.float64), (np.float32, torch.float32)):
        for km1, km2, nf, ncol in ((128, 128, 1, 6000), (72, 128, 4, 5001 + 3), (128, 64, 10, 4000)):
            ncol -= ncol % 4
            x1, y1, x2 = synth.lin_interp_inputs(dt, ncol, km1, km2, 1, col0=5)
            y1f = np.stack([(f + 1) * y1 for f in range(nf)]).astype(dt)
            want, _ = orc.lin_interp(x1, x2, y1f, km1, km2, 1)
            lev = lambda a: torch.from_numpy(np.ascontiguousarray(np.swapaxes(a, -1, -2))).to(dev)
            y2 = torch.zeros((nf, km2, ncol), dtype=tdt, device=dev)
            lin_interp_lm(lev(x1), lev(x2), lev(y1f), y2)
            torch.cuda.synchronize()
            assert bits_equal(np.ascontiguousarray(np.swapaxes(y2.cpu().numpy(), -1, -2)), want)
            print(f"  ok {lib().ekat_b200_last_kernel().decode():40s} {km1}->{km2} nf={nf} {ncol} {np.dtype(dt).name}", flush=True)


def fam_fused():
    for dt, tdt, TW in ((np.float64, torch.float64, 16), (np.float32, torch.float32, 32)):
        for nlev, km2, nrhs, ncol in ((128, 128, 10, SMS * TW * 3 + 8), (72, 128, 10, SMS * TW * 2 + 4), (43, 60, 6, SMS * TW * 2)):
            ncol -= ncol % 4
            A, B = orc.fill_tridiag(dt, ncol, nlev, 1, nrhs, col0=4)
            xs, _, xt = synth.lin_interp_inputs(dt, ncol, nlev, km2, 1, col0=4)
            lev = lambda a: torch.from_numpy(np.ascontiguousarray(a.T)).to(dev)
            Bd = torch.from_numpy(np.ascontiguousarray(B.transpose(2, 1, 0))).to(dev)
            Yd = torch.zeros((nrhs, km2, ncol), dtype=tdt, device=dev)
            Xd = torch.zeros((nrhs, nlev, ncol), dtype=tdt, device=dev)
            tridiag_lin_interp(lev(A[:, 0, :, 0]), lev(A[:, 1, :, 0]), lev(A[:, 2, :, 0]), Bd, lev(xs), lev(xt), Yd, X=Xd)
            torch.cuda.synchronize()
            k = lib().ekat_b200_last_kernel().decode()
            Xg = np.ascontiguousarray(Xd.cpu().numpy().transpose(0, 2, 1))  # [nrhs, ncol, nlev]
            want, _ = orc.lin_interp(xs, xt, Xg, nlev, km2, 1)               # remap of the solutions the kernel wrote
            assert bits_equal(np.ascontiguousarray(Yd.cpu().numpy().transpose(0, 2, 1)), want), k
            Xw = orc.tridiag("thomas", A.copy(), B.copy())
            assert rel_diff(Xw, Xg.transpose(1, 2, 0)) <= (1e-13 if dt == np.float64 else 1e-5)
            print(f"  ok {k:40s} fused {ncol} x {nlev}->{km2} x {nrhs} {np.dtype(dt).name}", flush=True)


def fam_reduction():
    from ekat_b200.reduction import adj_diff, view_reduction
    for dt, tdt in ((np.float64, torch.float64), (np.float32, torch.float32)):
        x = np.random.default_rng(3).random((5000, 80)).astype(dt)
        xd = torch.from_numpy(x).to(dev)
        for ser in (True, False):
            got = view_reduction(xd, 3, 72, serialize=ser, pack=16).cpu().numpy()
            assert bits_equal(got, orc.view_reduction(x, 3, 72, 0 if ser else 16))
        for fwd in (True, False):
            assert bits_equal(adj_diff(xd, fwd).cpu().numpy(), orc.adj_diff(x, fwd))
        xl = torch.from_numpy(np.ascontiguousarray(x.T)).to(dev)
        got = view_reduction(xl, 0, 80, serialize=True, layout=LM).cpu().numpy()
        assert bits_equal(got, orc.view_reduction(x, 0, 80, 0))
        print(f"  ok reduction / adj_diff {np.dtype(dt).name}", flush=True)


def fam_util():
    for tdt in (torch.float64, torch.float32):
        x = torch.rand((3001, 72), dtype=tdt, device=dev)
        y = T.col_to_lev(x)
        z = T.lev_to_col(y)
        torch.cuda.synchronize()
        assert torch.equal(x, z) and torch.equal(y, x.T.contiguous())
        dl, d, du, B = T.fill(tdt, 3001, 72, 1, 2, device=dev, pitch=3004)
        X = B.clone()
        T.thomas(dl, d, du, X, ncol=3001)
        r = float(T.residual(dl, d, du, B, X, ncol=3001).item())
        c = int(T.checksum(X, ncol=3001).item())
        assert r < 1e-4 and c != 0
    print("  ok util (transposes, fills, residual, checksum)", flush=True)


FAMS = {"shared": fam_shared, "bfb_stream": fam_bfb_stream, "spill": fam_spill, "tma": fam_tma, "cr": fam_cr, "tile": fam_tile,
        "levmajor": fam_levmajor, "lin_interp": fam_lin_interp, "lin_interp_lm": fam_lin_interp_lm, "fused": fam_fused,
        "reduction": fam_reduction, "util": fam_util}

if __name__ == "__main__":
    for name in (sys.argv[1:] or list(FAMS)):
        print(f"[{name}]", flush=True)
        FAMS[name]()
    print("all families ok")
