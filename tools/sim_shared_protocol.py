"""Randomised-schedule simulation of the mbarrier / TMA protocol of k_tridiag_tma_shared (csrc/tridiag_tma_shared.cu).

No arithmetic: every shared-memory slot carries a tag saying what it holds, every role is a coroutine that follows
the kernel's control flow statement by statement (same barriers, same parities, same issue order), TMA copies
complete after random delays, and the scheduler picks a random runnable role at every step. The run fails on
  * deadlock (nobody runnable, nothing in flight),
  * a read that finds the wrong tag (a role ran ahead of its producer or a slot was overwritten too early),
  * an asynchronous write into a slot that still has readers or an unread tensor store pending.
There is no GPU in the build container, so this is how the protocol is checked before it is run on a B200.
    python tools/sim_shared_protocol.py [trials]
"""
import random
import sys


class Bar:
    def __init__(self, count):
        self.count, self.pending, self.tx, self.phase = count, count, 0, 0

    def _check(self):
        if self.pending == 0 and self.tx == 0:
            self.phase += 1
            self.pending = self.count

    def arrive(self, tx=0):
        assert self.pending > 0, "more arrivals than the barrier expects"
        self.tx += tx
        self.pending -= 1
        self._check()

    def complete_tx(self, n):
        self.tx -= n
        self._check()

    def passed(self, parity):
        return (self.phase & 1) != parity


class Sim:
    SD = 3

    def __init__(self, rh, nch, NS, nit, exact, rng, io=False):
        self.io = io
        self.rh, self.nch, self.NS, self.nit, self.exact, self.rng = rh, nch, NS, nit, exact, rng
        self.nslot = nch
        self.UPT = 2 * nch if exact else nch
        self.raw = [Bar(1) for _ in range(NS)]
        self.ready = [Bar(1) for _ in range(NS)]
        self.freeb = [Bar(rh) for _ in range(NS)]
        self.plc = [Bar(1) for _ in range(self.nslot)]
        self.cfree = [Bar(rh) for _ in range(self.nslot)]
        self.plx = [[Bar(1) for _ in range(self.nslot)] for _ in range(rh)]
        self.xdone = [[Bar(1) for _ in range(self.nslot)] for _ in range(rh)]
        self.PC = [None] * self.nslot
        self.PX = [[None] * self.nslot for _ in range(rh)]
        self.RING = [[None, None] for _ in range(NS)]
        self.readers = {}      # slot key -> outstanding generic readers expected before an async overwrite
        self.inflight = []     # [time_left, fn]
        self.store_groups = [[] for _ in range(rh)]  # per solver warp: list of [done] flags in commit order
        self.stored = set()

    def slot(self, k, it):
        return self.nch - 1 - k if (it & 1) else k

    def par(self, k, it):
        return it & 1

    # ---- async engine
    def tma_load(self, where, idx, tag, bar, nbytes, guard, after=None):
        guard()

        def done():
            guard()
            where[idx] = tag
            if after:
                after()
            bar.complete_tx(nbytes)
        self.inflight.append([self.rng.randint(1, 40), done])

    def tma_store(self, j, sl, tag):
        flag = [False]
        self.store_groups[j].append(flag)

        def done():
            assert self.PX[j][sl] == tag, f"store of {tag} read {self.PX[j][sl]}"
            self.stored.add((j,) + tag[1:])
            flag[0] = True
        self.inflight.append([self.rng.randint(1, 25), done])

    def stores_read(self, j, allow):
        g = self.store_groups[j]
        return all(f[0] for f in g[:len(g) - allow]) if len(g) > allow else True

    # ---- roles (generators yield a predicate to wait on, or None to just be rescheduled)
    def load_x(self, j, it2, k):
        sl = self.slot(k, it2)
        # the slot must not have an unread store pending, and the solver must be done with it
        for f, s2 in self._store_slots[j]:
            assert f[0] or s2 != sl, f"X load into slot {sl} with an unread store pending"
        assert self.PX[j][sl] is None or self.PX[j][sl][0] == "S", f"X load over live data {self.PX[j][sl]}"
        self.plx[j][sl].arrive(tx=1)
        self.tma_load(self.PX[j], sl, ("B", it2, k), self.plx[j][sl], 1, lambda: None)

    def io_warp(self, j):
        nch, nit, SD = self.nch, self.nit, self.SD
        jb0 = 1 if self.nfull_partial >= 0 else 0
        for k in range(nch):
            self.load_x(j, 0, k)
        yield None
        for it in range(nit):
            has1 = it + 1 < nit
            for jb in range(nch):
                k = nch - 1 - jb
                sl = self.slot(k, it)
                yield lambda sl=sl, it=it: self.xdone[j][sl].passed(it & 1)
                if k == self.nfull_partial:
                    if has1:
                        self.load_x(j, it + 1, jb)
                else:
                    flagpos = len(self.store_groups[j])
                    self.tma_store(j, sl, ("S", it, k))
                    self._store_slots[j].append((self.store_groups[j][flagpos], sl))
                    if jb - jb0 >= SD:
                        yield lambda: self.stores_read(j, SD)
                        if has1:
                            self.load_x(j, it + 1, jb - SD)
                yield None
            yield lambda: self.stores_read(j, 0)
            if has1:
                for jb in range(max(jb0, nch - SD), nch):
                    self.load_x(j, it + 1, jb)

    def solver(self, j):
        nch, NS, nit = self.nch, self.NS, self.nit
        PX = self.PX[j]

        def load_x(it2, k):
            self.load_x(j, it2, k)

        if not self.io:
            for k in range(nch):
                load_x(0, k)
        yield None
        for it in range(nit):
            ubase = it * self.UPT
            has1 = it + 1 < nit

            def refill(jb):
                if has1:
                    load_x(it + 1, jb)
            for k in range(nch):
                sl, u0 = self.slot(k, it), ubase + k
                s = u0 % NS
                yield lambda sl=sl, k=k: self.plx[j][sl].passed(self.par(k, it))
                yield lambda s=s, u0=u0: self.ready[s].passed((u0 // NS) & 1)
                assert self.RING[s][0] == ("m", it, k), f"solver {j} fwd it{it} k{k}: ring holds {self.RING[s][0]}"
                if not self.exact:
                    assert self.RING[s][1] == ("r", it, k), f"ring r holds {self.RING[s][1]}"
                assert PX[sl] == ("B", it, k), f"solver {j} fwd: X slot holds {PX[sl]} want B it{it} k{k}"
                self.ring_readers[s] -= 1
                self.freeb[s].arrive()
                yield None
                PX[sl] = ("Y", it, k)
            SD = self.SD
            jb0 = 1 if self.nfull_partial >= 0 else 0
            for jb in range(nch):
                k = nch - 1 - jb
                sl = self.slot(k, it)
                s = 0
                if self.exact:
                    u0 = ubase + nch + jb
                    s = u0 % NS
                    yield lambda s=s, u0=u0: self.ready[s].passed((u0 // NS) & 1)
                    yield lambda s=s, u0=u0: self.raw[s].passed((u0 // NS) & 1)
                    assert self.RING[s][0] == ("du", it, k), f"bwd du ring holds {self.RING[s][0]} want it{it} k{k}"
                    assert self.RING[s][1] == ("rr", it, k), f"bwd rr ring holds {self.RING[s][1]} want it{it} k{k}"
                assert PX[sl] == ("Y", it, k)
                assert self.PC[sl] == ("C'", it, k), f"solver {j} bwd it{it} k{k}: PC holds {self.PC[sl]}"
                self.pc_readers[sl] -= 1
                self.cfree[sl].arrive()
                if self.exact:
                    self.ring_readers[s] -= 1
                    self.freeb[s].arrive()
                yield None
                if k == self.nfull_partial:
                    PX[sl] = ("S", it, k)
                    self.stored.add((j, it, k))
                    if self.io:
                        self.xdone[j][sl].arrive()
                    else:
                        refill(jb)
                elif self.io:
                    PX[sl] = ("S", it, k)
                    self.xdone[j][sl].arrive()
                else:
                    PX[sl] = ("S", it, k)
                    flagpos = len(self.store_groups[j])
                    self.tma_store(j, sl, ("S", it, k))
                    self._store_slots[j].append((self.store_groups[j][flagpos], sl))
                    if jb - jb0 >= SD:
                        yield lambda: self.stores_read(j, SD)
                        refill(jb - SD)
                yield None
            if not self.io:
                yield lambda: self.stores_read(j, 0)
                for jb in range(max(jb0, nch - SD), nch):
                    refill(jb)

    def factor(self):
        nch, NS = self.nch, self.NS
        for it in range(self.nit):
            ubase = it * self.UPT
            for k in range(nch):
                sl, u0 = self.slot(k, it), ubase + k
                s = u0 % NS
                yield lambda s=s, u0=u0: self.raw[s].passed((u0 // NS) & 1)
                yield lambda sl=sl, k=k, it=it: self.plc[sl].passed(self.par(k, it))
                assert self.RING[s][0] == ("dl", it, k), f"factor it{it} k{k}: ring holds {self.RING[s][0]}"
                assert self.RING[s][1] == (("du" if self.exact else "d"), it, k)
                assert self.PC[sl] == ("C", it, k), f"factor it{it} k{k}: PC holds {self.PC[sl]}"
                yield None
                self.RING[s][0] = ("m", it, k)
                if not self.exact:
                    self.RING[s][1] = ("r", it, k)
                self.PC[sl] = ("C'", it, k)
                self.ring_readers[s] = self.rh
                self.pc_readers[sl] = self.rh
                self.ready[s].arrive()
                yield None
            if self.exact:
                for jb in range(nch):
                    u0 = ubase + nch + jb
                    s = u0 % NS
                    sl = self.slot(nch - 1 - jb, it)
                    assert self.PC[sl] == ("C'", it, nch - 1 - jb), f"factor bwd it{it} jb{jb}: PC holds {self.PC[sl]}"
                    yield lambda s=s, u0=u0: self.raw[s].passed((u0 // NS) & 1)
                    assert self.ring_readers[s] == self.rh or True
                    self.RING[s][1] = ("rr", it, nch - 1 - jb)
                    self.ready[s].arrive()
                    yield None

    def ring_producer(self):
        nch, NS = self.nch, self.NS
        for it in range(self.nit):
            ubase = it * self.UPT
            for k in range(self.UPT):
                u0 = ubase + k
                s = u0 % NS
                if u0 >= NS:
                    yield lambda s=s, u0=u0: self.freeb[s].passed((u0 // NS - 1) & 1)

                def guard(s=s):
                    assert self.ring_readers[s] == 0, f"ring stage {s} overwritten with {self.ring_readers[s]} readers left"
                if k < nch:
                    self.raw[s].arrive(tx=2)
                    self.tma_load(self.RING[s], 0, ("dl", it, k), self.raw[s], 1, guard)
                    self.tma_load(self.RING[s], 1, (("du" if self.exact else "d"), it, k), self.raw[s], 1, guard)
                else:
                    self.raw[s].arrive(tx=1)
                    kk = 2 * nch - 1 - k

                    def arm(s=s):
                        self.ring_readers[s] = self.rh
                    self.tma_load(self.RING[s], 0, ("du", it, kk), self.raw[s], 1, guard, after=arm)
                yield None

    def plane_producer(self):
        nch = self.nch

        def load_c(it2, k):
            sl = self.slot(k, it2)

            def guard(sl=sl):
                assert self.pc_readers[sl] == 0, f"PC slot {sl} overwritten with readers left"
            self.plc[sl].arrive(tx=1)
            self.tma_load(self.PC, sl, ("C", it2, k), self.plc[sl], 1, guard)
        for k in range(nch):
            load_c(0, k)
        yield None
        for it in range(self.nit - 1):
            for jb in range(nch):
                sl = self.slot(nch - 1 - jb, it)
                yield lambda sl=sl, it=it: self.cfree[sl].passed(it & 1)
                load_c(it + 1, jb)
            yield None

    def run(self, partial_tail):
        self.nfull_partial = self.nch - 1 if partial_tail else -1
        self.ring_readers = [0] * self.NS
        self.pc_readers = [0] * self.nslot
        self._store_slots = {j: [] for j in range(self.rh)}
        roles = [self.solver(j) for j in range(self.rh)] + [self.factor(), self.ring_producer(), self.plane_producer()]
        if self.io:
            roles += [self.io_warp(j) for j in range(self.rh)]
        waiting = [None] * len(roles)
        alive = [True] * len(roles)
        steps = 0
        while any(alive):
            steps += 1
            assert steps < 5_000_000, "livelock"
            # advance the async engine
            for op in self.inflight:
                op[0] -= 1
            due = [op for op in self.inflight if op[0] <= 0]
            self.inflight = [op for op in self.inflight if op[0] > 0]
            self.rng.shuffle(due)
            for op in due:
                op[1]()
            runnable = [i for i in range(len(roles)) if alive[i] and (waiting[i] is None or waiting[i]())]
            if not runnable:
                assert self.inflight, "DEADLOCK: " + ", ".join(
                    f"role{i}" for i in range(len(roles)) if alive[i])
                continue
            i = self.rng.choice(runnable)
            try:
                waiting[i] = next(roles[i])
            except StopIteration:
                alive[i] = False
        # drain
        while self.inflight:
            for op in self.inflight:
                op[0] -= 1
            due = [op for op in self.inflight if op[0] <= 0]
            self.inflight = [op for op in self.inflight if op[0] > 0]
            for op in due:
                op[1]()
        want = {(j, it, k) for j in range(self.rh) for it in range(self.nit) for k in range(self.nch)}
        assert self.stored == want, f"missing stores: {sorted(want - self.stored)[:5]}"


def main(trials=300, seed=1):
    rng = random.Random(seed)
    for t in range(trials):
        rh = rng.choice([1, 2, 5])
        nch = rng.choice([1, 2, 3, 9, 16])
        NS = rng.choice([2, 3, 4, 5, 8])
        nit = rng.choice([1, 2, 3, 4, 7])
        exact = rng.random() < 0.5
        partial = rng.random() < 0.5
        Sim(rh, nch, NS, nit, exact, rng, io=rng.random() < 0.6).run(partial)
    return trials


if __name__ == "__main__":
    n = main(int(sys.argv[1]) if len(sys.argv) > 1 else 300)
    print(f"protocol ok over {n} random configurations/schedules")
