"""Column / Row parallel linear and the vocab-parallel embedding vs vanilla nn.Linear / nn.Embedding on gloo:
the reference's own assertions (tests/test_column_parallel_linear.py, test_row_parallel_linear.py,
test_parallel_vocab_embedding.py) with its tolerances and real failure propagation, on grids scaled down to CPU speed
(18 layer shapes x 3 batch shapes, 300-step loss histories).  The reference's full grids and 1000-step histories run on
GPUs over NCCL in tests/test_gpu_layer_grids.py."""
import math

import pytest
import torch
import torch.nn as nn

from conftest import free_port
from dist_utils import run_distributed


def _rng_states():
    return torch.get_rng_state()


def _col_worker(rank, world):
    import process_manager as pm
    from models.layers import ColumnParallelLinear
    torch.manual_seed(42)
    state = _rng_states()
    import itertools
    for idim, odim, add_bias in itertools.product((64, 128, 256), (64, 128, 512), (True, False)):
        layer = ColumnParallelLinear(idim, odim, add_bias=add_bias, gather_output=True)
        torch.set_rng_state(state)
        layer.reset_parameters()
        ref = nn.Linear(idim, odim, bias=add_bias)
        torch.set_rng_state(state)
        nn.init.kaiming_uniform_(ref.weight, a=math.sqrt(5))
        if add_bias:
            nn.init.zeros_(ref.bias)
        op = layer.odim_partition
        sl = slice(rank * op, (rank + 1) * op)
        assert layer.weight.shape == (odim // world, idim)
        assert torch.allclose(ref.weight.data[sl], layer.weight.data)
        for bs, seq in [(1, 32), (8, 64), (16, 128)]:
            x = torch.rand(bs, seq, idim, requires_grad=True)
            x2 = x.detach().clone().requires_grad_(True)
            y = layer(x)
            y.mean().backward()
            yr = ref(x2)
            yr.mean().backward()
            assert y.shape == yr.shape
            assert torch.allclose(y, yr, atol=1e-4)
            assert torch.allclose(x.grad, x2.grad, atol=1e-7)
            assert torch.allclose(layer.weight.grad, ref.weight.grad[sl], atol=1e-7)
            if add_bias:
                assert torch.allclose(layer.bias.grad, ref.bias.grad[sl], atol=1e-7)
            layer.zero_grad()
            ref.zero_grad()
    # multi-step SGD: weights and loss histories stay identical
    layer = ColumnParallelLinear(64, 128, gather_output=True)
    torch.set_rng_state(state)
    layer.reset_parameters()
    ref = nn.Linear(64, 128)
    torch.set_rng_state(state)
    nn.init.kaiming_uniform_(ref.weight, a=math.sqrt(5))
    nn.init.zeros_(ref.bias)
    hist = []
    for model in (layer, ref):
        torch.set_rng_state(state)
        opt = torch.optim.SGD(model.parameters(), lr=1e-2)
        losses = []
        for _ in range(300):
            bs = torch.randint(1, 8, (1,)).item()
            x = torch.rand(bs, 16, 64)
            loss = model(x).mean()
            opt.zero_grad()
            loss.backward()
            opt.step()
            losses.append(loss.item())
        hist.append(torch.tensor(losses))
    assert torch.allclose(hist[0], hist[1], atol=1e-6)
    op = layer.odim_partition
    assert torch.allclose(layer.weight.data, ref.weight.data[rank * op:(rank + 1) * op], atol=1e-6)


def _row_worker(rank, world):
    from models.layers import RowParallelLinear
    torch.manual_seed(42)
    state = _rng_states()
    import itertools
    for idim, odim, add_bias in itertools.product((64, 128, 256), (32, 128, 512), (True, False)):
        layer = RowParallelLinear(idim, odim, add_bias=add_bias, split_input=True)
        torch.set_rng_state(state)
        layer.reset_parameters()
        ref = nn.Linear(idim, odim, bias=add_bias)
        torch.set_rng_state(state)
        nn.init.kaiming_uniform_(ref.weight, a=math.sqrt(5))
        if add_bias:
            nn.init.zeros_(ref.bias)
        ip = layer.idim_partition
        sl = slice(rank * ip, (rank + 1) * ip)
        assert layer.weight.shape == (odim, idim // world)
        assert torch.allclose(ref.weight.data[:, sl], layer.weight.data)
        if add_bias:
            assert layer.bias.shape == (odim,)
        x = torch.rand(4, 16, idim, requires_grad=True)
        x2 = x.detach().clone().requires_grad_(True)
        y = layer(x)
        y.mean().backward()
        yr = ref(x2)
        yr.mean().backward()
        assert torch.allclose(y, yr, atol=1e-4)
        assert torch.allclose(x.grad, x2.grad, atol=1e-6)
        assert torch.allclose(layer.weight.grad, ref.weight.grad.chunk(world, dim=1)[rank], atol=1e-6)
        if add_bias:
            assert torch.allclose(layer.bias.grad, ref.bias.grad, atol=1e-6)


class _ParallelToy(nn.Module):
    def __init__(self, V, h, o):
        super().__init__()
        from models.layers import ColumnParallelLinear, ParallelVocabularyEmbedding
        self.emb = ParallelVocabularyEmbedding(V, h)
        self.lin = ColumnParallelLinear(h, o, gather_output=True)

    def reset_parameters(self):
        self.emb.reset_parameters()
        self.lin.reset_parameters()

    def forward(self, x):
        return self.lin(self.emb(x))


class _VanillaToy(nn.Module):
    def __init__(self, V, h, o):
        super().__init__()
        self.emb = nn.Embedding(V, h)
        self.lin = nn.Linear(h, o)

    def reset_parameters(self):
        nn.init.normal_(self.emb.weight)
        nn.init.kaiming_uniform_(self.lin.weight, a=math.sqrt(5))
        nn.init.zeros_(self.lin.bias)

    def forward(self, x):
        return self.lin(self.emb(x))


def _emb_worker(rank, world):
    from models.layers import ParallelVocabEmbedding, ParallelVocabularyEmbedding
    assert ParallelVocabEmbedding is ParallelVocabularyEmbedding
    torch.manual_seed(42)
    state = _rng_states()
    for V, h in [(64, 16), (1280, 64)]:
        emb = ParallelVocabularyEmbedding(V, h)
        torch.set_rng_state(state)
        emb.reset_parameters()
        ref = nn.Embedding(V, h)
        torch.set_rng_state(state)
        nn.init.normal_(ref.weight)
        assert emb.weight.shape == (V // world, h)
        assert torch.allclose(emb.weight.data, ref.weight.data[emb.vocab_st_idx:emb.vocab_ed_idx], atol=1e-6)
        x = torch.randint(0, V, (8, 16))
        x0 = x.clone()
        out = emb(x)
        assert torch.equal(x, x0), "the embedding must not mutate input_ids (reference defect D3)"
        assert torch.allclose(out, ref(x), atol=1e-6)
    # toy model, Adam steps: loss histories match
    hist, models = [], []
    for cls in (_ParallelToy, _VanillaToy):
        m = cls(128, 32, 64)
        torch.set_rng_state(state)
        m.reset_parameters()
        torch.set_rng_state(state)
        opt = torch.optim.Adam(m.parameters(), lr=1e-3)
        losses = []
        for _ in range(300):
            x = torch.randint(0, 128, (4, 12))
            loss = m(x).pow(2).mean()
            opt.zero_grad()
            loss.backward()
            opt.step()
            losses.append(loss.item())
        hist.append(torch.tensor(losses))
        models.append(m)
    assert torch.allclose(hist[0], hist[1], atol=1e-5)
    p, v = models
    assert torch.allclose(p.emb.weight.data, v.emb.weight.data[p.emb.vocab_st_idx:p.emb.vocab_ed_idx], atol=1e-5)


def _nondiv_vocab_worker(rank, world):
    from models.layers import ParallelVocabularyEmbedding
    torch.manual_seed(1)
    state = torch.get_rng_state()
    V = 10 * world + 3
    emb = ParallelVocabularyEmbedding(V, 8)
    torch.set_rng_state(state)
    emb.reset_parameters()       # the reference raises a shape mismatch here on the last rank (defect D4)
    ref = nn.Embedding(V, 8)
    torch.set_rng_state(state)
    nn.init.normal_(ref.weight)
    x = torch.arange(V).view(1, -1)
    assert torch.allclose(emb(x), ref(x), atol=1e-6)


@pytest.mark.parametrize("world", [2, 4])
def test_column_parallel_linear(world):
    run_distributed(_col_worker, world, free_port())


@pytest.mark.parametrize("world", [2, 4])
def test_row_parallel_linear(world):
    run_distributed(_row_worker, world, free_port())


@pytest.mark.parametrize("world", [2, 4])
def test_parallel_vocab_embedding(world):
    run_distributed(_emb_worker, world, free_port())


def test_non_divisible_vocab():
    run_distributed(_nondiv_vocab_worker, 2, free_port())


def test_rmsnorm_matches_reference_formula():
    import process_manager as pm
    from models.layers import LayerNorm, RMSNorm
    pm.init_pgm(1)
    n = RMSNorm(32)
    assert [k for k, _ in n.named_parameters()] == ["scale"] and n.eps == 1e-5
    x = torch.randn(4, 7, 32)
    ref = n.scale * (x.float() * (x.float().pow(2).mean(-1, keepdim=True) + 1e-5).rsqrt())
    assert torch.allclose(n(x), ref)
    ln = LayerNorm(32)
    assert torch.allclose(ln(x), torch.nn.functional.layer_norm(x, (32,)), atol=1e-6)
