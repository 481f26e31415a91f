"""Randomised model check of the attention-backward look-ahead protocol (csrc/attention_tcgen05.cu,
attn_bwd_persistent_kernel<64, V2 = true>, the kLook path), no GPU needed.

Roles transcribed from the kernel: TMA loader, MMA issuer (S / dP of step g+1 issued as soon as the math warps have read
S / dP of step g), eight math warps (P / dS, dQ drain, item epilogue).  Primitives: mbarriers with arrival counts and
phase parity, asynchronous TMA completions, one in-order tensor pipe for MMAs and commits.  A random scheduler
interleaves everything; each run must finish (no deadlock / lost wake-up) and respect the hazards the barriers are
there for: K/V buffers, Q/dO stages, S/dP TMEM columns, P/dS smem tiles, the dQ columns and the dK/dV accumulators.
"""
import random

import pytest

STAGES, KVBUFS = 2, 2


class Bar:
    def __init__(self, count):
        self.count, self.pending, self.phase = count, count, 0

    def arrive(self):
        self.pending -= 1
        assert self.pending >= 0
        if self.pending == 0:
            self.pending, self.phase = self.count, self.phase ^ 1

    def done(self, parity):
        return self.phase != parity


class World:
    def __init__(self):
        b = self.bar = {}
        for i in range(2):
            b[("kv_full", i)], b[("kv_empty", i)] = Bar(1), Bar(1)
            b[("qdo_full", i)], b[("qdo_empty", i)] = Bar(1), Bar(1)
        b["sdp_full"], b["pds_full"], b["dq_full"], b["s_free"], b["sdp_read"] = Bar(1), Bar(8), Bar(1), Bar(8), Bar(8)
        self.pipe, self.tma, self.buf, self.readers, self.errors = [], [], {}, {}, []

    def write(self, name, tag):
        if self.readers.get(name, 0):
            self.errors.append(f"{name} overwritten with {tag} while {self.readers[name]} issued readers are pending")
        self.buf[name] = tag

    def expect(self, name, tag, who):
        if self.buf.get(name) != tag:
            self.errors.append(f"{who} read {name} = {self.buf.get(name)} but expected {tag}")

    def reading(self, names, delta):
        for n in names:
            self.readers[n] = self.readers.get(n, 0) + delta


def items_of(nq, HB, cta, grid):
    """(w, n_it) for the snake order of the V2 kernel; item w = kv tile w // HB, n_it = nq - kv tile."""
    out, k, w = [], 0, cta
    while w < nq * HB:
        out.append((w, nq - w // HB))
        k += 1
        w = k * grid + ((grid - 1 - cta) if (k & 1) else cta)
    return out


def loader(W, items):
    gs = 0
    for item, (w, n_it) in enumerate(items):
        kb = item % KVBUFS
        yield lambda: W.bar[("kv_empty", kb)].done(((item // KVBUFS) & 1) ^ 1)
        W.tma.append((W.bar[("kv_full", kb)], lambda kb=kb, w=w: W.write(("KV", kb), w)))
        for it in range(n_it):
            st = gs % STAGES
            yield lambda st=st, gs=gs: W.bar[("qdo_empty", st)].done(((gs // STAGES) & 1) ^ 1)
            W.tma.append((W.bar[("qdo_full", st)], lambda st=st, w=w, it=it: W.write(("QdO", st), (w, it))))
            gs += 1


def mma(W, items):
    steps = [(item, w, it, n_it) for item, (w, n_it) in enumerate(items) for it in range(n_it)]

    def issue_sdp(g, item, w, it):
        st = g % STAGES
        yield lambda: W.bar[("qdo_full", st)].done((g // STAGES) & 1)
        names = (("QdO", st), ("KV", item % KVBUFS))
        W.reading(names, +1)

        def run():
            W.expect(("QdO", st), (w, it), f"SdP({g})")
            W.expect(("KV", item % KVBUFS), w, f"SdP({g})")
            W.reading(names, -1)
            W.write("SdP", g)
        W.pipe.append(("mma", run))
        W.pipe.append(("commit", W.bar["sdp_full"]))

    if steps:
        yield lambda: W.bar[("kv_full", 0)].done(0)
        yield from issue_sdp(0, *steps[0][:3])
    for gs, (item, w, it, n_it) in enumerate(steps):
        st, kb = gs % STAGES, item % KVBUFS
        has_next = gs + 1 < len(steps)
        if it + 1 == n_it and has_next:
            nitem = item + 1
            yield lambda nitem=nitem: W.bar[("kv_full", nitem % KVBUFS)].done((nitem // KVBUFS) & 1)
        if has_next:
            yield lambda gs=gs: W.bar["sdp_read"].done(gs & 1)
            yield from issue_sdp(gs + 1, *steps[gs + 1][:3])
        yield lambda gs=gs: W.bar["pds_full"].done(gs & 1)
        if gs > 0:
            yield lambda gs=gs: W.bar["s_free"].done((gs - 1) & 1)
        W.reading(("PdS", ("KV", kb)), +1)

        def run_dq(gs=gs, kb=kb, w=w):
            W.expect("PdS", gs, f"dQ({gs})")
            W.expect(("KV", kb), w, f"dQ({gs})")
            W.reading(("PdS", ("KV", kb)), -1)
            W.write("dQ", gs)
        W.pipe.append(("mma", run_dq))
        W.pipe.append(("commit", W.bar["dq_full"]))
        W.reading(("PdS", ("QdO", st)), +1)

        def run_dvk(gs=gs, st=st, w=w, it=it):
            W.expect("PdS", gs, f"dV/dK({gs})")
            W.expect(("QdO", st), (w, it), f"dV/dK({gs})")
            W.reading(("PdS", ("QdO", st)), -1)
            if it == 0:
                W.write("dVK", (w, 0))            # accumulate = 0: overwrites the previous item's accumulators
            else:
                W.expect("dVK", (w, it - 1), f"dV/dK({gs}) accumulate")
                W.buf["dVK"] = (w, it)
        W.pipe.append(("mma", run_dvk))
        W.pipe.append(("commit", W.bar[("qdo_empty", st)]))
        if it + 1 == n_it:
            W.pipe.append(("commit", W.bar[("kv_empty", kb)]))


def math(W, items, warp):
    gs = 0
    for item, (w, n_it) in enumerate(items):
        for it in range(n_it):
            yield lambda gs=gs: W.bar["sdp_full"].done(gs & 1)
            # first chunk: TMEM read, then (gs > 0) wait until dV / dK(gs-1) have consumed P / dS before storing
            if warp == 0:
                W.expect("SdP", gs, f"math({gs}) chunk 0")
            if gs > 0:
                yield lambda gs=gs: W.bar[("qdo_empty", (gs - 1) % STAGES)].done(((gs - 1) // STAGES) & 1)
            yield lambda: True
            # second chunk: TMEM read, release S / dP
            if warp == 0:
                W.expect("SdP", gs, f"math({gs}) chunk 1")
            W.bar["sdp_read"].arrive()
            yield lambda: True
            if warp == 0:
                W.write("PdS", gs)
            W.bar["pds_full"].arrive()
            yield lambda gs=gs: W.bar["dq_full"].done(gs & 1)
            if warp == 0:
                W.expect("dQ", gs, f"dQ drain({gs})")
            W.bar["s_free"].arrive()
            gs += 1
        kb = item % KVBUFS
        yield lambda kb=kb, item=item: W.bar[("kv_empty", kb)].done((item // KVBUFS) & 1)
        if warp == 0:
            W.expect("dVK", (w, n_it - 1), f"epilogue(item {item})")
            W.buf["dVK"] = (w, "read")


def run_once(nq, HB, grid, cta, seed):
    rng = random.Random(seed)
    items = items_of(nq, HB, cta, grid)
    W = World()
    roles = [loader(W, items), mma(W, items)] + [math(W, items, i) for i in range(8)]
    waiting, alive = [None] * len(roles), [True] * len(roles)
    for i, r in enumerate(roles):
        try:
            waiting[i] = next(r)
        except StopIteration:
            alive[i] = False
    steps = 0
    while any(alive) or W.pipe or W.tma:
        steps += 1
        assert steps < 2_000_000
        choices = [("role", i) for i in range(len(roles)) if alive[i] and waiting[i]()]
        if W.pipe:
            choices.append(("pipe", 0))
        choices += [("tma", k) for k in range(len(W.tma))]
        if not choices:
            raise AssertionError(f"deadlock: roles {[i for i in range(len(roles)) if alive[i]]} blocked "
                                 f"(0 loader, 1 MMA, 2-9 math); nq={nq} items={items}")
        kind, i = rng.choice(choices)
        if kind == "role":
            try:
                waiting[i] = next(roles[i])
            except StopIteration:
                alive[i] = False
        elif kind == "pipe":
            op, arg = W.pipe.pop(0)
            arg() if op == "mma" else arg.arrive()
        else:
            bar, fn = W.tma.pop(i)
            fn()
            bar.arrive()
        assert not W.errors, W.errors[:3]


@pytest.mark.parametrize("nq", [1, 2, 3, 5, 8])
def test_lookahead_protocol_has_no_deadlock_or_hazard(nq):
    for HB, grid in ((3, 2), (4, 3), (1, 1)):
        for cta in range(grid):
            for seed in range(6):
                run_once(nq, HB, grid, cta, seed)
