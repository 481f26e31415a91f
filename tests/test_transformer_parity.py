"""Tensor-parallel Transformer vs the single-device VallinaTransformer oracle on gloo (BASELINE.json config 1):
weight shards equal the oracle's slices (init-parity contract), loss / logits / sharded grads match.
Mirrors the reference's tests/test_transformers.py — which cannot even import (defect D1) — with tight tolerances."""
import pytest
import torch
import torch.nn.functional as F

from conftest import free_port
from dist_utils import run_distributed

CFG = dict(attn_dim=128, ffn_dim=256, num_heads=4, num_layers=2, vocab_size=512, maxlen=64)


def _compare_weights(para, van, rank, world):
    emb = para.embedding
    assert torch.allclose(emb.weight, van.embedding.weight[emb.vocab_st_idx:emb.vocab_ed_idx], atol=1e-6)

    def col(p, v):
        op = p.odim_partition
        sl = slice(op * rank, op * (rank + 1))
        assert torch.allclose(p.weight, v.weight[sl], atol=1e-6)
        assert torch.allclose(p.bias, v.bias[sl], atol=1e-6)

    def row(p, v):
        ip = p.idim_partition
        assert torch.allclose(p.weight, v.weight[:, ip * rank:ip * (rank + 1)], atol=1e-6)
        assert torch.allclose(p.bias, v.bias, atol=1e-6)

    col(para.lm_head, van.lm_head)
    for pl, vl in zip(para.layers, van.layers):
        for name in ("wq", "wk", "wv"):
            col(getattr(pl.attn, name), getattr(vl.attn, name))
        row(pl.attn.wo, vl.attn.wo)
        col(pl.ffn.gate_proj, vl.ffn.gate_proj)
        col(pl.ffn.up_proj, vl.ffn.up_proj)
        row(pl.ffn.down_proj, vl.ffn.down_proj)


def _worker(rank, world, cfg=None):
    from models.model import Transformer, VallinaTransformer
    CFG = cfg or globals()["CFG"]
    torch.manual_seed(7)
    state = torch.get_rng_state()
    para = Transformer(*CFG.values())            # positional, like the reference tests do
    torch.set_rng_state(state)
    para.reset_parameters()
    van = VallinaTransformer(**CFG)
    torch.set_rng_state(state)
    van.reset_parameters()
    _compare_weights(para, van, rank, world)
    assert len(para.state_dict()) == 4 + 16 * CFG["num_layers"]

    torch.manual_seed(100)
    for step in range(3):
        b, t = int(torch.randint(1, 5, (1,))), int(torch.randint(8, 48, (1,)))
        ids = torch.randint(0, CFG["vocab_size"], (b, t))
        ids0 = ids.clone()
        tgt = torch.randint(0, CFG["vocab_size"], (b, t))
        tgt[:, -2:] = -1
        pos = torch.arange(t).unsqueeze(0).repeat(b, 1)
        logits = para(ids, pos)
        assert torch.equal(ids, ids0)
        ref_logits = van(ids, pos)
        assert logits.shape == (b, t, CFG["vocab_size"])
        assert torch.allclose(logits, ref_logits, atol=2e-5), (logits - ref_logits).abs().max()
        loss = F.cross_entropy(logits.float().view(-1, logits.size(-1)), tgt.view(-1), ignore_index=-1)
        ref_loss = F.cross_entropy(ref_logits.view(-1, ref_logits.size(-1)), tgt.view(-1), ignore_index=-1)
        assert abs(loss.item() - ref_loss.item()) < 1e-5
        loss2 = para.loss(ids, pos, tgt)           # public loss API == forward + cross-entropy
        assert abs(loss2.item() - ref_loss.item()) < 1e-5
        para.zero_grad()
        van.zero_grad()
        loss.backward()
        ref_loss.backward()
        op = para.layers[0].attn.wq.odim_partition
        g, gr = para.layers[0].attn.wq.weight.grad, van.layers[0].attn.wq.weight.grad[op * rank:op * (rank + 1)]
        assert torch.allclose(g, gr, atol=1e-6)
        ip = para.layers[-1].ffn.down_proj.idim_partition
        g, gr = para.layers[-1].ffn.down_proj.weight.grad, van.layers[-1].ffn.down_proj.weight.grad[:, ip * rank:ip * (rank + 1)]
        assert torch.allclose(g, gr, atol=1e-6)
        # replicated parameters get identical grads on every rank
        assert torch.allclose(para.norm.scale.grad, van.norm.scale.grad, atol=1e-6)
        assert torch.allclose(para.layers[0].attn.wo.bias.grad, van.layers[0].attn.wo.bias.grad, atol=1e-6)
        e = para.embedding
        assert torch.allclose(e.weight.grad, van.embedding.weight.grad[e.vocab_st_idx:e.vocab_ed_idx], atol=1e-6)


@pytest.mark.parametrize("world", [2, 4])
def test_transformer_tp_matches_oracle(world):
    run_distributed(_worker, world, free_port())


def _grid_worker(rank, world):
    """A corner of the reference's model grid (tests/test_transformers.py:73-116: widths, head sizes, depths, vocab
    sizes, maxlens) scaled to CPU speed; the full 144-point grid runs on GPUs in tests/test_gpu_layer_grids.py."""
    import itertools
    for attn_dim, head_dim, layers, vocab in itertools.product((64, 128), (16, 32), (1, 3), (256, 1024)):
        if (attn_dim, head_dim, layers) in ((64, 32, 3), (128, 16, 3)):
            continue        # keep the CPU suite short: 12 of the 16 corners
        _worker(rank, world, dict(attn_dim=attn_dim, ffn_dim=2 * attn_dim, num_heads=attn_dim // head_dim,
                                  num_layers=layers, vocab_size=vocab, maxlen=48 + 16 * layers))


def test_transformer_tp_grid_matches_oracle():
    run_distributed(_grid_worker, 2, free_port())


def test_tp1_matches_oracle_and_api():
    import process_manager as pm
    from constants import IGNORE_INDEX, ModelArgumments
    from models.model import Transformer, VallinaTransformer, apply_rotary_pos_emb, get_cos_sin, rotate_half
    pm.init_pgm(1)
    assert IGNORE_INDEX == -1
    a = ModelArgumments()
    assert (a.attn_dim, a.ffn_dim, a.num_heads, a.num_layers, a.vocab_size, a.maxlen) == (512, 2048, 8, 12, 1024, 1000)
    torch.manual_seed(3)
    state = torch.get_rng_state()
    m = Transformer(**CFG)
    torch.set_rng_state(state)
    m.reset_parameters()
    v = VallinaTransformer(**CFG)
    torch.set_rng_state(state)
    v.reset_parameters()
    ids = torch.randint(0, CFG["vocab_size"], (2, 16))
    pos = torch.arange(16).unsqueeze(0).repeat(2, 1)
    assert torch.allclose(m(ids, pos), v(ids, pos), atol=1e-5)
    names = set(m.state_dict().keys())
    for k in ("embedding.weight", "layers.0.attn.wq.weight", "layers.0.attn.wo.bias", "layers.1.ffn.gate_proj.bias",
              "layers.1.ffn.down_proj.weight", "layers.0.norm1.scale", "layers.1.norm2.scale", "norm.scale",
              "lm_head.weight", "lm_head.bias"):
        assert k in names
    assert not any("rope" in k or "cos" in k for k in names), "RoPE tables must not be in the state_dict"
    m.retain_grad()
    # RoPE helpers keep the reference conventions
    x = torch.arange(8.).view(1, 1, 1, 8)
    assert torch.equal(rotate_half(x), torch.tensor([-4., -5, -6, -7, 0, 1, 2, 3]).view(1, 1, 1, 8))
    cos, sin = get_cos_sin(10, 8, 10000., device=torch.device("cpu"), dtype=torch.float32)
    assert cos.shape == (10, 8) and torch.allclose(cos[:, :4], cos[:, 4:])
    q, k = apply_rotary_pos_emb(x, x, cos[:1].unsqueeze(0), sin[:1].unsqueeze(0))
    assert torch.allclose(q, x) and torch.allclose(k, x)      # position 0 is the identity rotation


def test_default_model_parameter_count():
    import process_manager as pm
    from dataclasses import asdict
    from constants import ModelArgumments
    from models.model import Transformer
    pm.init_pgm(1)
    m = Transformer(**asdict(ModelArgumments()))
    n = sum(p.numel() for p in m.parameters())
    assert abs(n / 1e6 - 51.4739) < 1e-3, n          # 51.4739 M at TP=1
    assert len(list(m.parameters())) == 196
