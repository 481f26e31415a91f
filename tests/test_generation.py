"""KV-cached greedy decoding (models/generation.py) against the reference-style full re-forward per token
(reference test.py:143-160) on gloo, TP=1 and TP=2: same tokens, same stop condition; the cached last-position
hidden state reproduces the module path's logits; the vocab-parallel arg-max equals arg-max over gathered logits."""
import pytest
import torch
import torch.nn.functional as F

from conftest import free_port
from dist_utils import run_distributed

CFG = dict(attn_dim=64, ffn_dim=128, num_heads=4, num_layers=2, vocab_size=96, maxlen=48)
EOS = 1


def _reference_decode(model, tokens, eos_id, max_decode_len):
    """The reference's loop: whole prefix through the model for every new token."""
    while True:
        pos = torch.arange(tokens.size(-1)).unsqueeze(0)
        logits = model(tokens.clone(), pos)[0, -1]
        pred = int(logits.float().argmax(dim=-1).item())
        tokens = F.pad(tokens, (0, 1), value=pred)
        if pred == eos_id or tokens.size(-1) > max_decode_len:
            hit = pred == eos_id
            return (tokens[:, :-1] if hit else tokens), hit


def _worker(rank, world):
    from distributed_pytorch_from_scratch_b200.models import generation as G
    from models.model import Transformer
    torch.manual_seed(11)
    model = Transformer(**CFG)
    model.reset_parameters()
    model.eval()
    with torch.no_grad():
        # spread the logits so greedy choices are not decided by rounding noise
        model.lm_head.weight.mul_(4.0)
    torch.manual_seed(5)
    for trial in range(4):
        t0 = int(torch.randint(1, 9, (1,)))
        prompt = torch.randint(3, CFG["vocab_size"], (1, t0))
        max_len = 24 if trial < 3 else 200          # the last trial runs into the RoPE-table bound
        with torch.no_grad():
            if max_len < CFG["maxlen"]:
                want, want_eos = _reference_decode(model, prompt.clone(), EOS, max_len)
            got, got_eos = G.greedy_generate(model, prompt.clone(), EOS, max_len, dtype=torch.float32)
        if max_len < CFG["maxlen"]:
            assert got_eos == want_eos
            assert torch.equal(got, want), (got, want)
        else:
            assert got.size(1) <= CFG["maxlen"] + 1

    # prefill + single steps reproduce the full forward's last-position logits
    ids = torch.randint(3, CFG["vocab_size"], (2, 10))
    pos = torch.arange(10).unsqueeze(0).repeat(2, 1)
    with torch.no_grad():
        full = model(ids.clone(), pos)                                  # (2, 10, V) gathered logits
        cache = G.KVCache.allocate(model, 2, 16, torch.float32, ids.device)
        h = G.forward_cached(model, ids[:, :7], cache, torch.float32)
        for j in range(7, 10):
            assert cache.length == j
            nxt = G.vocab_parallel_argmax(model, h, cache, torch.float32)
            assert torch.equal(nxt, full[:, j - 1].argmax(-1))
            h = G.forward_cached(model, ids[:, j:j + 1], cache, torch.float32)
        local = F.linear(h, model.lm_head.weight, model.lm_head.bias)
        vl = model.lm_head.odim_partition
        assert torch.allclose(local, full[:, -1, rank * vl:(rank + 1) * vl], atol=2e-5)
        with pytest.raises(ValueError):
            G.forward_cached(model, ids[:, :7], cache, torch.float32)   # 10 + 7 > capacity 16

        # batched generation: every row equals its own single-sequence decode (up to its EOS)
        out = G.greedy_generate_batch(model, ids[:, :6], EOS, 8, dtype=torch.float32)
        for r in range(2):
            single, hit = G.greedy_generate(model, ids[r:r + 1, :6].clone(), EOS, 6 + 8 - 1, dtype=torch.float32)
            n = single.size(1)
            assert torch.equal(out[r, :n], single[0])
            if hit:
                assert bool((out[r, n:] == EOS).all())


@pytest.mark.parametrize("world", [1, 2])
def test_kv_cache_decode_matches_full_reforward(world):
    run_distributed(_worker, world, free_port())
