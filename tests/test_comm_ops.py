"""Split / Gather / Copy / Reduce forward + backward on gloo (reference models/comm_ops.py:7-83 semantics)."""
import pytest
import torch

from conftest import free_port
from dist_utils import run_distributed


def _worker(rank, world):
    from models.comm_ops import Copy, Gather, Reduce, Split
    torch.manual_seed(0)
    full = torch.randn(3, 5, 8 * world)
    c = full.shape[-1] // world
    # Split: local chunk forward, all-gather backward
    x = full.clone().requires_grad_(True)
    y = Split.apply(x)
    assert torch.equal(y, full[..., rank * c:(rank + 1) * c])
    (y * (rank + 1)).sum().backward()
    expect = torch.cat([torch.full((3, 5, c), float(r + 1)) for r in range(world)], dim=-1)
    assert torch.equal(x.grad, expect)
    # Gather: concat forward, local chunk backward
    loc = (full[..., rank * c:(rank + 1) * c]).clone().requires_grad_(True)
    g = Gather.apply(loc)
    assert torch.equal(g, full)
    w = torch.arange(full.numel(), dtype=torch.float32).view_as(full)
    (g * w).sum().backward()
    assert torch.equal(loc.grad, w[..., rank * c:(rank + 1) * c])
    # Reduce: sum forward (input untouched), identity backward
    a = torch.full((4, 4), float(rank + 1), requires_grad=True)
    a0 = a.detach().clone()
    r = Reduce.apply(a)
    assert torch.equal(a.detach(), a0), "Reduce must not mutate its input"
    assert torch.allclose(r, torch.full((4, 4), float(sum(range(1, world + 1)))))
    r.sum().backward()
    assert torch.equal(a.grad, torch.ones(4, 4))
    # Copy: identity forward, all-reduce backward
    b = torch.ones(2, 3, requires_grad=True)
    cpy = Copy.apply(b)
    assert torch.equal(cpy, b)
    (cpy * (rank + 1)).sum().backward()
    assert torch.allclose(b.grad, torch.full((2, 3), float(sum(range(1, world + 1)))))


@pytest.mark.parametrize("world", [2, 4])
def test_comm_ops_gloo(world):
    run_distributed(_worker, world, free_port())


def test_identity_at_tp1():
    import process_manager as pm
    from models.comm_ops import Copy, Gather, Reduce, Split
    pm.init_pgm(1)
    x = torch.randn(2, 6, requires_grad=True)
    for op in (Copy, Gather, Reduce, Split):
        y = op.apply(x)
        assert torch.equal(y, x)
    assert repr(pm.pgm) == "TP1" and pm.pgm.tp_rank == 0 and pm.pgm.grid.tolist() == [0]
