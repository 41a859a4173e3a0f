"""Randomised-schedule simulation of the plane hand-overs of k_solve_remap_roles (csrc/fused_tma.cu, BASELINE config 4 in
one kernel): NIW remap warps, NSW solver warps (TPS right-hand sides each per pass) and the factor warp walk the tiles of
a persistent CTA at their own pace; every plane of a tile (x1, the three factor planes, one plane per right-hand side) is
handed to the next tile the moment the current tile is done with it -- by the LAST remap warp out (an atomic counter), or by
the factor warp once every solver warp has arrived. Method as in sim_shared_protocol.py (tags on the planes, random copy
delays, a random runnable role per step). Fails on deadlock, on a read that finds another tile's data or an unfinished
factorisation, on a copy into a plane that still has readers, on a counter that is not back at zero, and if any
(tile, field) is not remapped by every warp exactly once.
    python tools/sim_fused_protocol.py [trials]
"""
import random
import sys

from sim_shared_protocol import Bar


class Sim:
    def __init__(self, nrhs, NIW, NSW, TPS, T, ntiles, two_boxes, rng, break_it=None):
        self.nrhs, self.NIW, self.NSW, self.TPS, self.T, self.ntiles, self.rng = nrhs, NIW, NSW, TPS, T, ntiles, rng
        self.two_boxes, self.break_it = two_boxes, break_it
        self.nsw = min(NSW, (nrhs + TPS - 1) // TPS)
        self.bx1, self.bA, self.bA2 = Bar(1), Bar(1), Bar(1)
        self.abar = Bar(self.nsw)
        self.lbar = [Bar(1) for _ in range(nrhs)]
        self.sbar = [Bar(1) for _ in range(nrhs)]
        self.fbar = [Bar(1) for _ in range(T)]
        self.cnt = [0] * (nrhs + 1)
        self.PX1 = [None]
        self.A = [None, None]          # upper / lower box of dl, d, du: ("A", tile) -> ("F", tile, steps done)
        self.PX = [None] * nrhs
        self.x1_readers = 0
        self.px_readers = [0] * nrhs
        self.solvers_in_tile = 0       # solver warps that have not arrived on abar for the tile whose factors are on chip
        self.inflight = []
        self.remapped = []

    def tma_load(self, where, idx, tag, bar, guard):
        guard()

        def done():
            guard()
            where[idx] = tag
            bar.complete_tx(1)
        self.inflight.append([self.rng.randint(1, 40), done])

    def load_A(self, tile):
        def guard():
            assert self.solvers_in_tile == 0, f"factor planes of tile {tile} requested with {self.solvers_in_tile} solver warps still in the previous tile"
        self.bA.arrive(tx=1)
        self.tma_load(self.A, 0, ("A", tile), self.bA, guard)
        if self.two_boxes:
            self.bA2.arrive(tx=1)
            self.tma_load(self.A, 1, ("A", tile), self.bA2, guard)
        else:
            self.A[1] = None

    def prologue(self):
        self.bx1.arrive(tx=1)
        self.tma_load(self.PX1, 0, ("x1", 0), self.bx1, lambda: None)
        self.load_A(0)
        for r in range(self.nrhs):
            self.lbar[r].arrive(tx=1)
            self.tma_load(self.PX, r, ("B", 0, r), self.lbar[r], lambda: None)

    def factor(self):
        for tile in range(self.ntiles):
            par = tile & 1
            if tile != 0:
                if self.break_it != "factor_without_abar":
                    yield lambda par=par: self.abar.passed(par ^ 1)
                self.load_A(tile)
            yield lambda par=par: self.bA.passed(par)
            if self.two_boxes:
                yield lambda par=par: self.bA2.passed(par)
            assert self.A[0] == ("A", tile) and (not self.two_boxes or self.A[1] == ("A", tile)), f"factor tile {tile}: planes hold {self.A}"
            self.solvers_in_tile = self.nsw
            for t in range(self.T):
                yield None
                self.A[0] = ("F", tile, t + 1)
                self.fbar[t].arrive()

    def solver(self, sw):
        if sw >= self.nsw:
            return
        for tile in range(self.ntiles):
            par = tile & 1
            for r0 in range(sw * self.TPS, self.nrhs, self.nsw * self.TPS):
                rs = [r for r in range(r0, r0 + self.TPS) if r < self.nrhs]
                for r in rs:  # (lanes of one warp: the warp goes on when every lane's right-hand side has landed)
                    yield lambda r=r, par=par: self.lbar[r].passed(par)
                    assert self.PX[r] == ("B", tile, r), f"solver {sw} tile {tile}: plane {r} holds {self.PX[r]}"
                for t in range(self.T):
                    yield lambda t=t, par=par: self.fbar[t].passed(par)
                    assert self.A[0][:2] == ("F", tile) and self.A[0][2] >= t + 1, f"solver {sw} tile {tile} step {t}: factors {self.A[0]}"
                yield None
                for r in rs:
                    self.PX[r] = ("S", tile, r)
                    self.px_readers[r] = self.NIW
                    self.sbar[r].arrive()
            self.solvers_in_tile -= 1
            self.abar.arrive()
            yield None

    def remap(self, w):
        last = (lambda c: c == 0) if self.break_it == "first_one_out" else (lambda c: c == self.NIW - 1)
        for tile in range(self.ntiles):
            par = tile & 1
            has_next = tile + 1 < self.ntiles
            yield lambda par=par: self.bx1.passed(par)
            assert self.PX1[0] == ("x1", tile), f"remap {w} tile {tile}: x1 plane holds {self.PX1[0]}"
            self.x1_readers += 1
            yield None  # the bracket search reads x1
            self.x1_readers -= 1
            c = self.cnt[self.nrhs]
            self.cnt[self.nrhs] += 1
            if last(c):
                if self.break_it != "first_one_out":
                    self.cnt[self.nrhs] = 0
                if has_next:
                    def guard():
                        assert self.x1_readers == 0, "x1 of the next tile requested while a warp is still searching"
                    self.bx1.arrive(tx=1)
                    self.tma_load(self.PX1, 0, ("x1", tile + 1), self.bx1, guard)
            if self.break_it == "first_one_out" and self.cnt[self.nrhs] == self.NIW:
                self.cnt[self.nrhs] = 0
            for f in range(self.nrhs):
                yield lambda f=f, par=par: self.sbar[f].passed(par)
                assert self.PX[f] == ("S", tile, f), f"remap {w} tile {tile}: solution plane {f} holds {self.PX[f]}"
                yield None  # remap + stores of field f
                self.remapped.append((tile, f, w))
                self.px_readers[f] -= 1
                c = self.cnt[f]
                self.cnt[f] += 1
                if last(c):
                    if self.break_it != "first_one_out":
                        self.cnt[f] = 0
                    if has_next:
                        def guard(f=f):
                            assert self.px_readers[f] == 0, f"right-hand side {f} of the next tile requested with {self.px_readers[f]} remap warps still reading the solution"
                        self.lbar[f].arrive(tx=1)
                        self.tma_load(self.PX, f, ("B", tile + 1, f), self.lbar[f], guard)
                if self.break_it == "first_one_out" and self.cnt[f] == self.NIW:
                    self.cnt[f] = 0

    def run(self):
        self.prologue()
        roles = [self.factor()] + [self.solver(sw) for sw in range(self.NSW)] + [self.remap(w) for w in range(self.NIW)]
        waiting = [None] * len(roles)
        alive = [True] * len(roles)
        steps = 0
        while any(alive):
            steps += 1
            assert steps < 2_000_000, "livelock"
            for op in self.inflight:
                op[0] -= 1
            due = [op for op in self.inflight if op[0] <= 0]
            self.inflight = [op for op in self.inflight if op[0] > 0]
            self.rng.shuffle(due)
            for op in due:
                op[1]()
            runnable = [i for i in range(len(roles)) if alive[i] and (waiting[i] is None or waiting[i]())]
            if not runnable:
                assert self.inflight, "DEADLOCK: " + ", ".join(f"role{i}" for i in range(len(roles)) if alive[i])
                continue
            i = self.rng.choice(runnable)
            try:
                waiting[i] = next(roles[i])
            except StopIteration:
                alive[i] = False
        assert self.cnt == [0] * (self.nrhs + 1), f"last-one-out counters not back at zero: {self.cnt}"
        want = sorted((t, f, w) for t in range(self.ntiles) for f in range(self.nrhs) for w in range(self.NIW))
        assert sorted(self.remapped) == want, "a (tile, field) was not remapped by every warp exactly once"


def main(trials=300, seed=1, break_it=None):
    rng = random.Random(seed)
    for _ in range(trials):
        nrhs = rng.choice([1, 2, 3, 5, 10, 13])
        NIW = rng.choice([2, 4, 10])
        NSW = rng.choice([1, 2, 5])
        TPS = rng.choice([1, 2])
        T = rng.choice([1, 2, 8, 16])
        ntiles = rng.choice([1, 2, 3, 4, 7])
        Sim(nrhs, NIW, NSW, TPS, T, ntiles, rng.random() < 0.5, rng, break_it=break_it).run()
    return trials


if __name__ == "__main__":
    n = main(int(sys.argv[1]) if len(sys.argv) > 1 else 300)
    print(f"protocol ok over {n} random configurations/schedules")
