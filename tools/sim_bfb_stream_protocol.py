"""Randomised-schedule simulation of the hand-offs of k_tridiag_bfb_stream (csrc/tridiag_bfb_stream.cu), the form with two
sweeping warps (PP) on level-major operands: the io thread (ring of input stages, pivot plane, tensor stores of the results
out of two output buffers, the read-back of x' through L2 once the forward stores have COMPLETED), the two sweeping warps
taking the chunks of a tile in turn, the pair of named barriers under which the chain state crosses between them, and the
mbarriers with the kernel's phase formulas. Method as in sim_shared_protocol.py: every buffer carries a tag saying what it
holds, copies complete after random delays, a random runnable role runs at every step; the run fails on deadlock, on a read
that finds the wrong tag, on a write into a buffer an unread tensor store still needs, on a named barrier completed
without its waiter, and if any chunk is not stored exactly once.
    python tools/sim_bfb_stream_protocol.py [trials]
"""
import random
import sys

from sim_shared_protocol import Bar


class NamedBar:
    """bar.arrive / bar.sync id, 64: completes when both warps have arrived; one of them must be the waiter."""

    def __init__(self):
        self.count, self.gen, self.has_sync = 0, 0, False

    def _arrive(self, is_sync):
        self.count += 1
        self.has_sync = self.has_sync or is_sync
        if self.count == 2:
            assert self.has_sync, "named barrier completed by two bar.arrive: a hand-over was never waited for"
            self.count, self.has_sync = 0, False
            self.gen += 1

    def arrive(self):
        self._arrive(False)

    def sync_enter(self):
        g = self.gen
        self._arrive(True)
        return lambda: self.gen > g


class Sim:
    def __init__(self, nch, NS, pd, rng, partial, break_it=None):
        self.nch, self.NS, self.pd, self.rng, self.break_it = nch, NS, pd, rng, break_it
        self.nfull = nch - 1 if partial else nch  # index of a ragged top chunk (nch: none)
        self.NU = 2 * nch
        self.rbar = [Bar(1) for _ in range(NS)]
        self.ebar = [Bar(1) for _ in range(NS)]
        self.dbar = [Bar(1) for _ in range(2)]
        self.obar = [Bar(1) for _ in range(2)]
        self.pbar = [Bar(1) for _ in range(nch)]
        self.nb = {1: NamedBar(), 2: NamedBar()}
        self.RG = [[None, None, None] for _ in range(NS)]
        self.stage_held = [False] * NS   # operands not yet in a sweeping warp's registers
        self.P0 = [None] * nch
        self.OB = [None, None]
        self.HS = None
        self.inflight = []
        self.groups = []                 # tensor-store commit groups of the io thread: lists of {"read":, "done":}
        self.open_group = []
        self.xprime_done = [False] * nch
        self.stored = []
        self.ob_store_pending = [None, None]  # store that still has to read OB[b]

    # ---- async engine
    def later(self, lo, hi, fn):
        self.inflight.append([self.rng.randint(lo, hi), fn])

    def tma_load(self, where, idx, tag, bar, guard):
        guard()

        def done():
            guard()
            where[idx] = tag
            bar.complete_tx(1)
        self.later(1, 40, done)

    def tma_store_ob(self, b, tag, on_done):
        st = {"read": False, "done": False}
        self.open_group.append(st)
        self.ob_store_pending[b] = st

        def read():
            assert self.OB[b] == tag, f"store of {tag} read {self.OB[b]} from OB[{b}]"
            st["read"] = True

            def done():
                st["done"] = True
                on_done()
            self.later(1, 30, done)
        self.later(1, 25, read)

    def commit(self):
        self.groups.append(self.open_group)
        self.open_group = []

    def reads_done(self, allow):
        g = self.groups
        return all(s["read"] for grp in g[:max(0, len(g) - allow)] for s in grp)

    def all_done(self):
        return all(s["done"] for grp in self.groups for s in grp)

    # ---- io thread
    def io(self):
        nch, NS, NU, pd, nfull = self.nch, self.NS, self.NU, self.pd, self.nfull
        state = {"next_u": 0, "last_e": -1, "fwd_stores": 0, "pending": -1, "flushed": nch <= 2}
        eph = [0] * NS
        dph = [0, 0]

        def issue_use(u):
            s = u % NS

            assert not self.stage_held[s], f"ring stage {s} reissued before its operands were taken"
            self.stage_held[s] = True  # from the issue of a use until a sweeping warp has its operands in registers

            def guard():
                pass
            if u < nch:
                self.rbar[s].arrive(tx=3)
                self.tma_load(self.RG[s], 0, ("dl", u), self.rbar[s], guard)
                self.tma_load(self.RG[s], 1, ("du", u), self.rbar[s], guard)
                self.tma_load(self.RG[s], 2, ("x", u), self.rbar[s], guard)
            else:
                jb = u - nch
                k = nch - 1 - jb
                self.rbar[s].arrive(tx=2 if jb >= 2 else 1)
                self.tma_load(self.RG[s], 1, ("du", k), self.rbar[s], guard)
                if jb >= 2:
                    assert self.xprime_done[k], f"x' of chunk {k} read back before its store completed"
                    self.tma_load(self.RG[s], 2, ("x'", k), self.rbar[s], guard)

        def issue_plane(k):
            self.pbar[k].arrive(tx=1)
            self.tma_load(self.P0, k, ("d", k), self.pbar[k], lambda: None)

        def pump():
            # generator: may have to block on the completion of the forward stores
            while state["next_u"] < NU and state["next_u"] - NS <= state["last_e"]:
                if state["next_u"] >= nch + 2 and not state["flushed"]:
                    if state["fwd_stores"] < nch - 2:
                        break
                    if self.break_it != "readback_without_flush":
                        yield self.all_done
                    state["flushed"] = True
                issue_use(state["next_u"])
                state["next_u"] += 1

        def handle_d(t):
            k = t if t < nch else 2 * nch - 1 - t
            b = k & 1
            yield lambda b=b: self.dbar[b].passed(dph[b])
            dph[b] ^= 1
            st = (k <= nch - 3) if t < nch else (k != nfull)
            if st:
                if t < nch:
                    self.tma_store_ob(b, ("x'", k), lambda k=k: self.xprime_done.__setitem__(k, True))
                    state["fwd_stores"] += 1
                else:
                    self.tma_store_ob(b, ("x", k), lambda k=k: self.stored.append(k))
                self.commit()
            if state["pending"] >= 0:
                if self.break_it != "ob_free_without_wait":
                    yield (lambda: self.reads_done(1)) if st else (lambda: self.reads_done(0))
                self.obar[state["pending"]].arrive()
                state["pending"] = -1
            if st:
                state["pending"] = b
            else:
                self.obar[b].arrive()

        for k in range(min(pd, nch)):
            issue_plane(k)
        yield from pump()
        for t in range(NU):
            s = t % NS
            yield lambda s=s: self.ebar[s].passed(eph[s])
            eph[s] ^= 1
            state["last_e"] = t
            yield from pump()
            if t + pd < nch:
                issue_plane(t + pd)
            if t >= 1:
                yield from handle_d(t - 1)
                yield from pump()
            yield None
        yield from handle_d(NU - 1)
        yield lambda: self.reads_done(0)

    # ---- sweeping warp cw
    def sweeper(self, cw):
        nch, NS, nfull = self.nch, self.NS, self.nfull

        def hand_wait():
            return self.nb[1 + cw].sync_enter()

        def hand_over():
            self.nb[2 - cw].arrive()

        def write_ob(b, tag):
            st = self.ob_store_pending[b]
            assert st is None or st["read"], f"OB[{b}] overwritten with {tag} while a tensor store still has to read it"
            self.OB[b] = tag

        # forward sweep
        for k in range(cw, nch, 2):
            s, b = k % NS, k & 1
            yield lambda s=s, k=k: self.rbar[s].passed((k // NS) & 1)
            yield lambda k=k: self.pbar[k].passed(0)
            assert self.RG[s] == [("dl", k), ("du", k), ("x", k)], f"fwd chunk {k}: stage holds {self.RG[s]}"
            assert self.P0[k] == ("d", k), f"fwd chunk {k}: plane holds {self.P0[k]}"
            self.stage_held[s] = False
            self.ebar[s].arrive()
            yield None
            if k > 0:
                yield hand_wait()
                assert self.HS == ("fwd", k - 1), f"fwd chunk {k} took chain state {self.HS}"
            yield None
            if k < nch - 1:
                self.HS = ("fwd", k)
                hand_over()
            self.P0[k] = ("d'", k)
            m = k >> 1
            if m >= 1:
                yield lambda b=b, m=m: self.obar[b].passed((m - 1) & 1)
            write_ob(b, ("x'", k))
            if k == nch - 1:
                hand_over()
            self.dbar[b].arrive()
            yield None
        # back substitution
        for jb in range((nch + cw) & 1, nch, 2):
            k = nch - 1 - jb
            u0 = nch + jb
            s, b = u0 % NS, k & 1
            kmax = (nch - 1) - (1 if ((nch - 1) & 1) != b else 0)
            m = ((nch - b + 1) >> 1) + ((kmax - k) >> 1)
            yield lambda s=s, u0=u0: self.rbar[s].passed((u0 // NS) & 1)
            if jb == 0:
                yield hand_wait()
            if jb == 1:
                yield hand_wait()
                assert self.HS == ("x", 0), f"bwd step 1 took {self.HS}"
            # loads
            assert self.RG[s][1] == ("du", k), f"bwd chunk {k}: stage holds du {self.RG[s][1]}"
            if jb <= 1:
                assert self.OB[b] == ("x'", k), f"bwd chunk {k}: OB[{b}] holds {self.OB[b]}"
            else:
                assert self.RG[s][2] == ("x'", k), f"bwd chunk {k}: stage holds x {self.RG[s][2]}"
            assert self.P0[k] == ("d'", k), f"bwd chunk {k}: plane holds {self.P0[k]}"
            self.stage_held[s] = False
            self.ebar[s].arrive()
            yield None
            if jb > 1:
                yield hand_wait()
                assert self.HS == ("x", jb - 1), f"bwd step {jb} took {self.HS}"
            yield None
            if k == nfull:  # ragged top chunk (jb == 0): plain stores; wait_ob comes before them in the kernel
                if m >= 1:
                    yield lambda b=b, m=m: self.obar[b].passed((m - 1) & 1)
                self.stored.append(k)
                if jb < nch - 1:
                    self.HS = ("x", jb)
                    hand_over()
                self.dbar[b].arrive()
            else:
                if jb < nch - 1:
                    self.HS = ("x", jb)
                    hand_over()
                if m >= 1:
                    yield lambda b=b, m=m: self.obar[b].passed((m - 1) & 1)
                write_ob(b, ("x", k))
                self.dbar[b].arrive()
            yield None

    def run(self):
        roles = [self.sweeper(0), self.sweeper(1), self.io()]
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
        while self.inflight:
            for op in self.inflight:
                op[0] -= 1
            due = [op for op in self.inflight if op[0] <= 0]
            self.inflight = [op for op in self.inflight if op[0] > 0]
            for op in due:
                op[1]()
        assert sorted(self.stored) == list(range(self.nch)), f"chunks stored: {sorted(self.stored)}"


def main(trials=300, seed=1, break_it=None):
    rng = random.Random(seed)
    for _ in range(trials):
        nch = rng.choice([1, 2, 3, 4, 5, 9, 12, 16, 17])
        NS = rng.choice([2, 3, 4, 5, 6])
        pd = rng.choice([1, 2, 4, 8])
        Sim(nch, NS, pd, rng, partial=rng.random() < 0.5, break_it=break_it).run()
    return trials


if __name__ == "__main__":
    n = main(int(sys.argv[1]) if len(sys.argv) > 1 else 300)
    print(f"protocol ok over {n} random configurations/schedules")
