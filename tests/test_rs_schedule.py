"""Host-side check of the fused GEMM -> reduce-scatter schedule (csrc/gemm_tcgen05.cu: rs_block_of and the owner
signalling), no GPU needed: a block that is never computed, computed twice, or flagged to a rank that does not wait for
it would hang the kernel on hardware, so the schedule is verified here for every (M, n) the model can produce."""
import ctypes

import pytest

from distributed_pytorch_from_scratch_b200.ops import _lib

BM = 128


def _schedule(M, n):
    tiles = (M + BM - 1) // BM
    arr = lambda: (ctypes.c_int * tiles)()  # noqa: E731
    order, lo, hi = arr(), arr(), arr()
    rc = _lib.lib().dpfs_debug_rs_schedule(M, n, ctypes.addressof(order), ctypes.addressof(lo), ctypes.addressof(hi))
    return rc, list(order), list(lo), list(hi)


@pytest.mark.parametrize("n", [2, 4, 8])
@pytest.mark.parametrize("rows_per_rank", [128, 129, 200, 256, 1000, 4000, 4001, 8000, 16000])
def test_schedule_is_a_fair_bijection(n, rows_per_rank):
    M = rows_per_rank * n
    rc, order, lo, hi = _schedule(M, n)
    assert rc == 0
    tiles = (M + BM - 1) // BM
    assert sorted(order) == list(range(tiles)), "every row-block is computed exactly once"
    for b in range(tiles):
        assert 0 <= lo[b] <= hi[b] < n and hi[b] - lo[b] <= 1, "a block touches at most two owners"
        assert lo[b] == (b * BM) // rows_per_rank and hi[b] == (min(b * BM + BM, M) - 1) // rows_per_rank
    # what each owner's comm warps wait for == what gets flagged to it
    for r in range(n):
        row_lo, row_hi = r * rows_per_rank, (r + 1) * rows_per_rank
        waited = set(range(row_lo // BM, (row_hi - 1) // BM + 1))
        flagged = {b for b in range(tiles) if lo[b] == r or hi[b] == r}
        assert waited == flagged, (r, sorted(waited ^ flagged))
    # fairness: after any prefix of the compute order the owners' counts of finished primary blocks differ by <= 1
    done = [0] * n
    for b in order:
        done[lo[b]] += 1
        assert max(done) - min(done) <= 1 or min(done) == min(
            sum(1 for x in range(tiles) if lo[x] == r) for r in range(n)), done
    # primary blocks of an owner are computed in ascending order (the comm warps wait for them in that order)
    for r in range(n):
        mine = [b for b in order if lo[b] == r]
        assert mine == sorted(mine)


def test_rejects_unsupported_shapes():
    assert _schedule(1000, 8)[0] == -53      # 125 rows per rank < one tile
    assert _schedule(1001, 2)[0] == -53      # not divisible


# ---- GEMM + hand-off in one kernel (csrc/gemm_handoff.cu): same tile order, comm work items of 16 / 8 rows ----------
def _gh_schedule(M, N, n, rank):
    tiles = (M + BM - 1) // BM
    cap = M // n // 8 + 2
    order = (ctypes.c_int * tiles)()
    arrs = [(ctypes.c_int * cap)() for _ in range(4)]
    rc = _lib.lib().dpfs_debug_gh_schedule(M, N, n, rank, ctypes.addressof(order), *[ctypes.addressof(a) for a in arrs],
                                           cap)
    return rc, list(order), [list(a)[:max(rc, 0)] for a in arrs]


@pytest.mark.parametrize("n", [2, 4, 8])
@pytest.mark.parametrize("rows_per_rank", [128, 130, 200, 1000, 4000, 4001])
@pytest.mark.parametrize("N", [256, 512, 1024])
def test_gemm_handoff_schedule(n, rows_per_rank, N):
    """Every rank computes every row-block once, in the fused reduce-scatter's order; a rank's comm items tile its rows
    exactly, and every item only waits for row-blocks whose producers flag this rank (a wait without a matching flag
    would spin until the watchdog traps)."""
    M = rows_per_rank * n
    tiles = (M + BM - 1) // BM
    rs_order = _schedule(M, n)[1]
    rows_per_item = 16 if N <= 512 else 8
    for rank in range(n):
        rc, order, (ilo, ihi, blo, bhi) = _gh_schedule(M, N, n, rank)
        assert rc == (rows_per_rank + rows_per_item - 1) // rows_per_item
        assert order == rs_order and sorted(order) == list(range(tiles))
        assert ilo[0] == 0 and ihi[-1] == rows_per_rank and all(a == b for a, b in zip(ihi[:-1], ilo[1:]))
        g0 = rank * rows_per_rank
        for lo, hi, b0, b1 in zip(ilo, ihi, blo, bhi):
            assert 0 < hi - lo <= rows_per_item and 0 <= b1 - b0 <= 1
            assert b0 == (g0 + lo) // BM and b1 == (g0 + hi - 1) // BM
            for b in range(b0, b1 + 1):     # the producers of block b flag the owners of its first and last row
                owners = {(b * BM) // rows_per_rank, (min(b * BM + BM, M) - 1) // rows_per_rank}
                assert rank in owners, (rank, b, owners)
