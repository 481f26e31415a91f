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
