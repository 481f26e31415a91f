"""Work distribution of the persistent attention kernels (csrc/attention_tcgen05.cu), checked on the host: every item is
processed by exactly one CTA under both orders, and the snake order of the opt-in kernels shortens the critical path
for the causal workload (items sorted heavy-first) -- most of all when there are only a couple of rounds (TP = 8)."""
import ctypes

import pytest

from distributed_pytorch_from_scratch_b200.ops import _lib


def _items(n_items, grid, bid, snake):
    buf = (ctypes.c_int * 64)()
    n = _lib.lib().dpfs_debug_attn_items(n_items, grid, bid, int(snake), ctypes.addressof(buf), 64)
    assert n <= 64
    return list(buf[:n])


def _forward_costs(nq, HB):
    # forward item w: q tile nq-1 - w // HB -> (q tile + 1) kv steps
    return [nq - (w // HB) for w in range(nq * HB)]


@pytest.mark.parametrize("nq,HB,grid", [(8, 256, 148), (8, 32, 148), (8, 64, 148), (16, 128, 148), (3, 5, 4), (1, 7, 3)])
def test_every_item_once_and_snake_is_never_worse(nq, HB, grid):
    n_items = nq * HB
    g = min(grid, n_items)
    cost = _forward_costs(nq, HB)
    span = {}
    for snake in (False, True):
        seen, loads = [], []
        for bid in range(g):
            mine = _items(n_items, g, bid, snake)
            assert mine[0] == bid if mine else True
            seen += mine
            loads.append(sum(cost[w] for w in mine))
        assert sorted(seen) == list(range(n_items)), "each work item belongs to exactly one CTA"
        span[snake] = max(loads)
    assert span[True] <= span[False]
    ideal = -(-sum(cost) // g)
    assert span[True] <= ideal + max(cost), (span, ideal)


def test_tp8_flagship_makespan():
    # b32 x t1000, one head per rank: 256 items on 148 CTAs -> round-robin 12 steps on the busiest CTA, snake 8
    cost = _forward_costs(8, 32)
    def makespan(snake):
        return max(sum(cost[w] for w in _items(256, 148, bid, snake)) for bid in range(148))
    assert makespan(False) == 12 and makespan(True) == 8
