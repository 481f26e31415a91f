"""Randomised model check of the ping-pong attention forward's synchronisation protocol (csrc/attention_tcgen05.cu,
attn_fwd_pp_kernel), no GPU needed.

The kernel's four roles (TMA loader, MMA issuer, softmax group A, softmax group B) are transcribed as coroutines over
a small model of the hardware primitives they use:

* mbarrier with an arrival count and a phase bit; ``wait(parity)`` passes once the phase of that parity has completed;
* TMA loads complete (arrive on their barrier) asynchronously;
* tcgen05 MMAs and ``tcgen05.commit`` execute in issue order on one tensor pipe, at arbitrary later times.

A random scheduler interleaves the roles, the pipe and the TMA completions.  Every run must terminate (no deadlock, no
lost wake-up from a phase-parity mistake) and must respect the data hazards: no buffer (Q, K/V stage, S, P, O) is
overwritten while a reader that was issued earlier has not consumed it, and nothing is read before it was produced.
"""
import random

import pytest

KV_STAGES = 3


class Bar:
    def __init__(self, count):
        self.count, self.pending, self.phase = count, count, 0

    def arrive(self):
        self.pending -= 1
        assert self.pending >= 0, "more arrivals than the barrier was initialised for"
        if self.pending == 0:
            self.pending, self.phase = self.count, self.phase ^ 1

    def done(self, parity):
        return self.phase != parity


class World:
    """Shared state + hazard bookkeeping.  Buffers hold a tag describing their contents; readers check the tag."""

    def __init__(self, nq, n_items_for_cta):
        self.bar = {}
        for i in range(2):
            for name, cnt in (("q_full", 1), ("q_empty", 1), ("s_full", 1), ("p_full", 4), ("o_done", 1)):
                self.bar[(name, i)] = Bar(cnt)
        for i in range(KV_STAGES):
            for name in ("k_full", "v_full", "kv_empty"):
                self.bar[(name, i)] = Bar(1)
        self.pipe = []           # tensor pipe FIFO: ("mma", fn) or ("commit", bar)
        self.tma = []            # outstanding TMA loads: (bar, fn)
        self.buf = {}            # buffer name -> content tag
        self.readers = {}        # buffer name -> number of issued-but-not-executed MMA readers
        self.errors = []

    # hazard helpers
    def write(self, name, tag):
        if self.readers.get(name, 0):
            self.errors.append(f"{name} overwritten with {tag} while {self.readers[name]} issued readers are pending")
        self.buf[name] = tag

    def expect(self, name, tag, who):
        if self.buf.get(name) != tag:
            self.errors.append(f"{who} read {name} = {self.buf.get(name)} but expected {tag}")


def items_of(nq, HB, cta, grid):
    n_pairs = (nq + 1) // 2
    n_items = n_pairs * HB
    return [(w, n_pairs - 1 - w // HB) for w in range(cta, n_items, grid)]


def loader(W, nq, items):
    g = 0
    for it, (w, pair) in enumerate(items):
        has_b = 2 * pair + 1 < nq
        n_kv = 2 * pair + 2 if has_b else 2 * pair + 1
        qb = it & 1
        yield lambda: W.bar[("q_empty", qb)].done(((it >> 1) & 1) ^ 1)

        def q_land(w=w, qb=qb, has_b=has_b):
            W.write(("Q", qb, 0), (w, "A"))
            if has_b:
                W.write(("Q", qb, 1), (w, "B"))
        W.tma.append((W.bar[("q_full", qb)], q_land))
        for j in range(n_kv):
            st = g % KV_STAGES
            yield lambda st=st, g=g: W.bar[("kv_empty", st)].done(((g // KV_STAGES) & 1) ^ 1)
            W.tma.append((W.bar[("k_full", st)], lambda st=st, w=w, j=j: W.write(("K", st), (w, j))))
            W.tma.append((W.bar[("v_full", st)], lambda st=st, w=w, j=j: W.write(("V", st), (w, j))))
            g += 1


def mma(W, nq, items):
    g = 0
    cnt = [0, 0]
    for it, (w, pair) in enumerate(items):
        has_b = 2 * pair + 1 < nq
        n_kv = 2 * pair + 2 if has_b else 2 * pair + 1
        n_of = [2 * pair + 1, 2 * pair + 2 if has_b else 0]
        qb = it & 1
        g0 = g

        def issue_qk(x, j):
            st = (g0 + j) % KV_STAGES
            yield lambda: W.bar[("k_full", st)].done(((g0 + j) // KV_STAGES) & 1)
            for name in (("Q", qb, x), ("K", st)):
                W.readers[name] = W.readers.get(name, 0) + 1

            def run(x=x, j=j, st=st, w=w, qb=qb):
                W.expect(("Q", qb, x), (w, "AB"[x]), f"QK_{'AB'[x]}({j})")
                W.expect(("K", st), (w, j), f"QK_{'AB'[x]}({j})")
                for name in (("Q", qb, x), ("K", st)):
                    W.readers[name] -= 1
                W.write(("S", x), (w, j))
            W.pipe.append(("mma", run))
            W.pipe.append(("commit", W.bar[("s_full", x)]))

        yield lambda: W.bar[("q_full", qb)].done((it >> 1) & 1)
        yield from issue_qk(0, 0)
        if has_b:
            yield from issue_qk(1, 0)
        if n_kv == 1:
            W.pipe.append(("commit", W.bar[("q_empty", qb)]))
        for j in range(n_kv):
            st = (g0 + j) % KV_STAGES
            v_ready = False
            for x in range(2):
                if j >= n_of[x]:
                    continue
                yield lambda x=x: W.bar[("p_full", x)].done(cnt[x] & 1)
                if not v_ready:
                    yield lambda: W.bar[("v_full", st)].done(((g0 + j) // KV_STAGES) & 1)
                    v_ready = True
                for name in (("P", x), ("V", st)):
                    W.readers[name] = W.readers.get(name, 0) + 1

                def run(x=x, j=j, st=st, w=w):
                    W.expect(("P", x), (w, j), f"PV_{'AB'[x]}({j})")
                    W.expect(("V", st), (w, j), f"PV_{'AB'[x]}({j})")
                    for name in (("P", x), ("V", st)):
                        W.readers[name] -= 1
                    if j == 0:
                        W.write(("O", x), (w, 0))        # accumulate = 0: overwrites the previous item's O
                    else:
                        W.expect(("O", x), (w, j - 1), f"PV_{'AB'[x]}({j}) accumulate")
                        W.buf[("O", x)] = (w, j)
                W.pipe.append(("mma", run))
                W.pipe.append(("commit", W.bar[("o_done", x)]))
                cnt[x] += 1
                if j + 1 < n_of[x]:
                    yield from issue_qk(x, j + 1)
                    if j + 2 == n_kv and (x == 1 or not has_b):
                        W.pipe.append(("commit", W.bar[("q_empty", qb)]))
            W.pipe.append(("commit", W.bar[("kv_empty", st)]))
        g += n_kv


def softmax(W, nq, items, x, warp):
    """One of the four warps of group x (each arrives once on p_full[x]); warp 0 does the hazard bookkeeping."""
    cnt = 0
    for (w, pair) in items:
        qt = 2 * pair + x
        if qt >= nq:
            continue
        for j in range(qt + 1):
            yield lambda: W.bar[("s_full", x)].done(cnt & 1)
            if warp == 0:
                W.expect(("S", x), (w, j), f"softmax_{'AB'[x]}({j}) pass 1")
            if j > 0:
                yield lambda: W.bar[("o_done", x)].done((cnt - 1) & 1)
                if warp == 0:
                    W.expect(("O", x), (w, j - 1), f"softmax_{'AB'[x]}({j}) rescale")
            yield lambda: True      # (pass 2 takes time: let the others run)
            if warp == 0:
                W.expect(("S", x), (w, j), f"softmax_{'AB'[x]}({j}) pass 2")
                W.write(("P", x), (w, j))
            W.bar[("p_full", x)].arrive()
            cnt += 1
        yield lambda: W.bar[("o_done", x)].done((cnt - 1) & 1)
        if warp == 0:
            W.expect(("O", x), (w, qt), f"epilogue_{'AB'[x]}")
            W.buf[("O", x)] = (w, "read")       # the next item's PV(0) may now overwrite it


def run_once(nq, HB, grid, cta, seed):
    rng = random.Random(seed)
    items = items_of(nq, HB, cta, grid)
    W = World(nq, len(items))
    roles = [loader(W, nq, items), mma(W, nq, items)]
    for x in range(2):
        for warp in range(4):
            roles.append(softmax(W, nq, items, x, warp))
    waiting = [None] * len(roles)
    alive = [True] * len(roles)
    for i, r in enumerate(roles):
        try:
            waiting[i] = next(r)
        except StopIteration:
            alive[i] = False
    steps = 0
    while any(alive) or W.pipe or W.tma:
        steps += 1
        assert steps < 2_000_000, "runaway simulation"
        choices = []
        for i in range(len(roles)):
            if alive[i] and waiting[i]():
                choices.append(("role", i))
        if W.pipe:
            choices.append(("pipe", 0))
        for k in range(len(W.tma)):
            choices.append(("tma", k))
        if not choices:
            blocked = [i for i in range(len(roles)) if alive[i]]
            raise AssertionError(f"deadlock: roles {blocked} blocked (0 = loader, 1 = MMA, 2-5 = A, 6-9 = B); "
                                 f"nq={nq} items={items}")
        kind, i = rng.choice(choices)
        if kind == "role":
            try:
                waiting[i] = next(roles[i])
            except StopIteration:
                alive[i] = False
        elif kind == "pipe":
            op, arg = W.pipe.pop(0)
            if op == "mma":
                arg()
            else:
                arg.arrive()
        else:
            bar, fn = W.tma.pop(i)
            fn()
            bar.arrive()
        assert not W.errors, W.errors[:3]
    return steps


@pytest.mark.parametrize("nq", [1, 2, 3, 4, 5, 8])
def test_pingpong_protocol_has_no_deadlock_or_hazard(nq):
    # HB work items per pair and a small grid so that a CTA walks several items of different sizes back to back
    for HB, grid in ((3, 2), (4, 3), (1, 1)):
        for cta in range(grid):
            for seed in range(6):
                run_once(nq, HB, grid, cta, seed)
