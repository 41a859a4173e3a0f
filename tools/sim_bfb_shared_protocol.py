"""Randomised-schedule simulation of the mbarrier / TMA protocol of k_tridiag_bfb_shared (csrc/tridiag_bfb_shared.cu):
rh forward warps, rh backward warps, the factor warp and the three producer threads, with the kernel's barriers, parities,
ring positions and issue order, statement by statement (see sim_shared_protocol.py for the method: slots carry tags, TMA
copies complete after random delays, a random runnable role runs at every step). The run fails on a deadlock, on a read
that finds the wrong tag, on an asynchronous write into a slot that still has readers or an unread tensor store pending,
and if any chunk of any right-hand side is not stored exactly once.
compute-sanitizer is closed on this GPU pool; this is how the hand-offs are checked besides repeat-run bit-identity.
    python tools/sim_bfb_shared_protocol.py [trials]
"""
import random
import sys

from sim_shared_protocol import Bar


class Ring:
    """Ring position of a role, as the kernel's rs / rph / ring_advance."""

    def __init__(self, n):
        self.n, self.s, self.ph = n, 0, 0

    def advance(self):
        self.s += 1
        if self.s == self.n:
            self.s, self.ph = 0, self.ph ^ 1


class Sim:
    SD = 1  # kSD of the kernel

    def __init__(self, rh, nch, NS, NU, nit, rng, partial, break_it=None):
        self.rh, self.nch, self.NS, self.NU, self.nit, self.rng = rh, nch, NS, NU, nit, rng
        self.nfull_partial = nch - 1 if partial else -1
        self.break_it = break_it  # deliberately broken variants, to show that the model notices (tests)
        ns = nch
        self.raw = [Bar(1) for _ in range(NS)]
        self.ready = [Bar(1) for _ in range(NS)]
        self.freeb = [Bar(rh) for _ in range(NS)]
        self.uraw = [Bar(1) for _ in range(NU)]
        self.ufree = [Bar(rh) for _ in range(NU)]
        self.plc = [Bar(1) for _ in range(ns)]
        self.cfree = [Bar(rh) for _ in range(ns)]
        self.fdone = [Bar(1) for _ in range(rh)]
        self.plx = [[Bar(1) for _ in range(ns)] for _ in range(rh)]
        self.PD = [None] * ns
        self.PX = [[None] * ns for _ in range(rh)]
        self.RING = [[None, None] for _ in range(NS)]
        self.URING = [None] * NU
        self.ring_readers = [0] * NS
        self.u_readers = [0] * NU
        self.pd_readers = [0] * ns
        self.inflight = []
        self.store_groups = [[] for _ in range(rh)]
        self.store_slots = [[] for _ in range(rh)]
        self.stored = []

    def slot(self, k, it):
        return self.nch - 1 - k if (it & 1) else k

    # ---- async engine
    def tma_load(self, where, idx, tag, bar, guard, after=None):
        guard()

        def done():
            guard()
            where[idx] = tag
            if after:
                after()
            bar.complete_tx(1)
        self.inflight.append([self.rng.randint(1, 40), done])

    def tma_store(self, j, sl, tag):
        flag = [False]
        self.store_groups[j].append(flag)
        self.store_slots[j].append((flag, sl))

        def done():
            assert self.PX[j][sl] == tag, f"store of {tag} read {self.PX[j][sl]}"
            self.stored.append((j,) + tag[1:])
            flag[0] = True
        self.inflight.append([self.rng.randint(1, 25), done])

    def stores_read(self, j, allow):
        g = self.store_groups[j]
        return all(f[0] for f in g[:len(g) - allow]) if len(g) > allow else True

    def load_x(self, j, it2, k):
        sl = self.slot(k, it2)
        for f, s2 in self.store_slots[j]:
            assert f[0] or s2 != sl, f"X load into slot {sl} of rhs {j} with an unread store pending"
        assert self.PX[j][sl] is None or self.PX[j][sl][0] == "S", f"X load over live data {self.PX[j][sl]}"
        self.plx[j][sl].arrive(tx=1)
        self.tma_load(self.PX[j], sl, ("B", it2, k), self.plx[j][sl], lambda: None)

    # ---- roles
    def forward(self, j):
        r = Ring(self.NS)
        for it in range(self.nit):
            for k in range(self.nch):
                sl, s, ph = self.slot(k, it), r.s, r.ph
                yield lambda sl=sl, it=it: self.plx[j][sl].passed(it & 1)
                yield lambda s=s, ph=ph: self.ready[s].passed(ph)
                r.advance()
                assert self.RING[s][0] == ("m", it, k), f"fwd {j} it{it} k{k}: ring holds {self.RING[s][0]}"
                assert self.PX[j][sl] == ("B", it, k), f"fwd {j} it{it} k{k}: X slot holds {self.PX[j][sl]}"
                self.ring_readers[s] -= 1
                self.freeb[s].arrive()
                yield None
                self.PX[j][sl] = ("Y", it, k)
            self.fdone[j].arrive()
            yield None

    def backward(self, j):
        r = Ring(self.NU)
        nch, SD = self.nch, self.SD
        jb0 = 1 if self.nfull_partial >= 0 else 0
        for k in range(nch):
            self.load_x(j, 0, k)
        yield None
        for it in range(self.nit):
            has1 = it + 1 < self.nit

            def refill(jb):
                if has1:
                    self.load_x(j, it + 1, jb)
            yield lambda it=it: self.fdone[j].passed(it & 1)
            for jb in range(nch):
                k = nch - 1 - jb
                sl, s, ph = self.slot(k, it), r.s, r.ph
                yield lambda s=s, ph=ph: self.uraw[s].passed(ph)
                r.advance()
                assert self.URING[s] == ("du", it, k), f"bwd {j} it{it} k{k}: du ring holds {self.URING[s]}"
                assert self.PD[sl] == ("C'", it, k), f"bwd {j} it{it} k{k}: pivot slot holds {self.PD[sl]}"
                assert self.PX[j][sl] == ("Y", it, k), f"bwd {j} it{it} k{k}: X slot holds {self.PX[j][sl]}"
                self.pd_readers[sl] -= 1
                self.u_readers[s] -= 1
                self.cfree[sl].arrive()
                self.ufree[s].arrive()
                yield None
                self.PX[j][sl] = ("S", it, k)
                if k == self.nfull_partial:
                    self.stored.append((j, it, k))  # plain stores
                    refill(jb)
                else:
                    self.tma_store(j, sl, ("S", it, k))
                    if jb - jb0 >= SD:
                        if self.break_it != "refill_without_wait":
                            yield lambda: self.stores_read(j, SD)
                        refill(jb - SD)
                yield None
            yield lambda: self.stores_read(j, 0)
            for jb in range(max(jb0, nch - SD), nch):
                refill(jb)

    def factor(self):
        r = Ring(self.NS)
        for it in range(self.nit):
            for k in range(self.nch):
                sl, s, ph = self.slot(k, it), r.s, r.ph
                yield lambda s=s, ph=ph: self.raw[s].passed(ph)
                r.advance()
                yield lambda sl=sl, it=it: self.plc[sl].passed(it & 1)
                assert self.RING[s] == [("dl", it, k), ("du", it, k)], f"factor it{it} k{k}: ring holds {self.RING[s]}"
                assert self.PD[sl] == ("C", it, k), f"factor it{it} k{k}: pivot slot holds {self.PD[sl]}"
                yield None
                self.RING[s][0] = ("m", it, k)
                self.PD[sl] = ("C'", it, k)
                self.ring_readers[s] = self.rh
                self.pd_readers[sl] = self.rh
                self.ready[s].arrive()
                yield None

    def ring_producer(self):
        r = Ring(self.NS)
        wrapped = False
        for it in range(self.nit):
            for k in range(self.nch):
                s, ph = r.s, r.ph
                if wrapped and self.break_it != "ring_without_wait":
                    yield lambda s=s, ph=ph: self.freeb[s].passed(ph ^ 1)
                r.advance()
                if r.s == 0:
                    wrapped = True

                def guard(s=s):
                    assert self.ring_readers[s] == 0, f"ring stage {s} overwritten with {self.ring_readers[s]} readers left"
                self.raw[s].arrive(tx=2)
                self.tma_load(self.RING[s], 0, ("dl", it, k), self.raw[s], guard)
                self.tma_load(self.RING[s], 1, ("du", it, k), self.raw[s], guard)
                yield None

    def du_producer(self):
        r = Ring(self.NU)
        wrapped = False
        for it in range(self.nit):
            for jb in range(self.nch):
                s, ph = r.s, r.ph
                if wrapped:
                    yield lambda s=s, ph=ph: self.ufree[s].passed(ph ^ 1)
                r.advance()
                if r.s == 0:
                    wrapped = True

                def guard(s=s):
                    assert self.u_readers[s] == 0, f"du ring stage {s} overwritten with {self.u_readers[s]} readers left"

                def arm(s=s):
                    self.u_readers[s] = self.rh
                self.uraw[s].arrive(tx=1)
                self.tma_load(self.URING, s, ("du", it, self.nch - 1 - jb), self.uraw[s], guard, after=arm)
                yield None

    def plane_producer(self):
        nch = self.nch

        def load_c(it2, k):
            sl = self.slot(k, it2)

            def guard(sl=sl):
                assert self.pd_readers[sl] == 0, f"pivot slot {sl} overwritten with readers left"
            self.plc[sl].arrive(tx=1)
            self.tma_load(self.PD, sl, ("C", it2, k), self.plc[sl], guard)
        for k in range(nch):
            load_c(0, k)
        yield None
        for it in range(self.nit - 1):
            for jb in range(nch):
                sl = self.slot(nch - 1 - jb, it)
                if self.break_it != "plane_without_wait":
                    yield lambda sl=sl, it=it: self.cfree[sl].passed(it & 1)
                load_c(it + 1, jb)
                yield None

    def run(self):
        roles = ([self.forward(j) for j in range(self.rh)] + [self.backward(j) for j in range(self.rh)] +
                 [self.factor(), self.ring_producer(), self.du_producer(), self.plane_producer()])
        waiting = [None] * len(roles)
        alive = [True] * len(roles)
        steps = 0
        while any(alive):
            steps += 1
            assert steps < 5_000_000, "livelock"
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
        while self.inflight:
            for op in self.inflight:
                op[0] -= 1
            due = [op for op in self.inflight if op[0] <= 0]
            self.inflight = [op for op in self.inflight if op[0] > 0]
            for op in due:
                op[1]()
        want = sorted((j, it, k) for j in range(self.rh) for it in range(self.nit) for k in range(self.nch))
        assert sorted(self.stored) == want, "a chunk was stored twice or not at all"


def main(trials=300, seed=1, break_it=None):
    rng = random.Random(seed)
    for _ in range(trials):
        rh = rng.choice([1, 2, 3, 5, 6])
        nch = rng.choice([1, 2, 3, 9, 16])
        NS = rng.choice([4, 5, 7, 8])
        NU = rng.choice([3, 4])
        nit = rng.choice([1, 2, 3, 4, 7])
        Sim(rh, nch, NS, NU, nit, rng, partial=rng.random() < 0.5, break_it=break_it).run()
    return trials


if __name__ == "__main__":
    n = main(int(sys.argv[1]) if len(sys.argv) > 1 else 300)
    print(f"protocol ok over {n} random configurations/schedules")
