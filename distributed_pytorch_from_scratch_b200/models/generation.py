"""Incremental (KV-cached) greedy decoding for the tensor-parallel Transformer.

The reference decodes by re-running the whole prefix through the model for every generated token
(reference test.py:143-150: ``model(input_ids, position_ids)[0, -1]`` inside a ``while True``), i.e. O(t^2) layer
passes per sequence and a (b, H/n, t, t) score tensor per layer per token.  Serving does not want that, so this module
keeps the per-layer K / V of the prefix (already rotated by RoPE, sharded by heads exactly like training:
``models/model.py:55-60`` of the reference) and feeds one new position per step:

* prefill:  one causal pass over the prompt, K / V written to the cache;
* step:     q of the new token against the cached K / V (no mask needed: everything in the cache is in the past);
* the next token is the arg-max over the *vocab-parallel* logits: every rank reduces its own V/n slice to
  (max, arg-max) and only those two numbers per sequence are all-gathered -- the (b, V) logits are never assembled
  (the reference gathers the full logits on every rank, models/layers.py:99).

Collectives per generated token: 2 all-reduces of (b, 1, d) per layer + one (b, 2) all-gather, on the TP group.
The math is the model's own (same weights, same op order as ``Transformer.forward``), so greedy tokens match the
full re-forward; ``tests/test_generation.py`` checks that at TP=1 and TP=2.

Everything here is plain PyTorch on whatever device / dtype the caller picks: single-token GEMVs are bandwidth-bound
library calls, not a place for the tcgen05 kernels (those serve the prefill-shaped training path).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import List, Optional, Tuple

import torch
import torch.nn.functional as F

from ..parallel import comm_ops
from ..parallel import process_manager as pm
from .model import rotate_half


@dataclass
class KVCache:
    """Per-layer K / V of the positions processed so far: ``k[l], v[l]`` are (b, H/n, capacity, hd)."""
    k: List[torch.Tensor]
    v: List[torch.Tensor]
    length: int = 0
    weights: dict = field(default_factory=dict)   # compute-dtype copies of the parameters, made once per cache

    @property
    def capacity(self) -> int:
        return self.k[0].size(2)

    @staticmethod
    def allocate(model, batch: int, capacity: int, dtype: torch.dtype, device) -> "KVCache":
        n = pm.pgm.tp_size
        nh, hd = model.num_heads // n, model.head_dim
        capacity = min(capacity, model.maxlen)
        mk = lambda: torch.zeros(batch, nh, capacity, hd, dtype=dtype, device=device)  # noqa: E731
        return KVCache([mk() for _ in model.layers], [mk() for _ in model.layers])


def _w(cache: KVCache, p: Optional[torch.Tensor], dtype: torch.dtype) -> Optional[torch.Tensor]:
    """Parameter in the compute dtype (cast once per cache, not once per token)."""
    if p is None or p.dtype == dtype:
        return p
    key = id(p)
    hit = cache.weights.get(key)
    if hit is None or hit[0] != p._version:
        hit = (p._version, p.detach().to(dtype))
        cache.weights[key] = hit
    return hit[1]


def _rmsnorm(x: torch.Tensor, norm) -> torch.Tensor:
    # the module's own formula (fp32 statistics, cast back, then the fp32 scale) -- models/layers.py RMSNorm.forward
    xf = x.float()
    return (norm.scale * (xf * (xf.pow(2).mean(-1, keepdim=True) + norm.eps).rsqrt()).type_as(x)).to(x.dtype)


def _rope(t: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor) -> torch.Tensor:
    return t * cos + rotate_half(t) * sin


def _embed(model, ids: torch.Tensor) -> torch.Tensor:
    emb = model.embedding
    if pm.pgm.tp_size == 1:
        return F.embedding(ids, emb.weight)
    mask = (ids >= emb.vocab_st_idx) & (ids < emb.vocab_ed_idx)
    local = torch.where(mask, ids - emb.vocab_st_idx, torch.zeros_like(ids))
    out = F.embedding(local, emb.weight) * mask.unsqueeze(-1).to(emb.weight.dtype)
    return comm_ops.all_reduce_sum_(out)


@torch.no_grad()
def forward_cached(model, ids: torch.Tensor, cache: KVCache, dtype: torch.dtype) -> torch.Tensor:
    """Run ``ids`` (b, t_new) as positions [cache.length, cache.length + t_new) and append their K / V.
    Returns the hidden state of the LAST new position after the final norm: (b, d)."""
    b, t_new = ids.shape
    p0, p1 = cache.length, cache.length + t_new
    if p1 > cache.capacity:
        raise ValueError(f"KV cache overflow: {p1} positions > capacity {cache.capacity} (model.maxlen bounds RoPE)")
    hd = model.head_dim
    cos = model.rope_cos_f32[p0:p1].to(dtype)[None, None]       # (1, 1, t_new, hd)
    sin = model.rope_sin_f32[p0:p1].to(dtype)[None, None]
    x = _embed(model, ids).to(dtype)
    causal = None
    if t_new > 1:   # prefill: new position i sees cached positions and new positions <= i
        causal = torch.ones(t_new, p1, dtype=torch.bool, device=ids.device).tril(diagonal=p0)
    for li, layer in enumerate(model.layers):
        a, f = layer.attn, layer.ffn
        nh = a.num_local_heads
        h = _rmsnorm(x, layer.norm1)
        q, k, v = (F.linear(h, _w(cache, m.weight, dtype), _w(cache, m.bias, dtype)) for m in (a.wq, a.wk, a.wv))
        q, k, v = (z.view(b, t_new, nh, hd).transpose(1, 2) for z in (q, k, v))        # (b, nh, t_new, hd)
        q, k = _rope(q, cos, sin), _rope(k, cos, sin)
        cache.k[li][:, :, p0:p1] = k
        cache.v[li][:, :, p0:p1] = v
        keys, vals = cache.k[li][:, :, :p1], cache.v[li][:, :, :p1]
        score = torch.matmul(q, keys.transpose(-1, -2)) / math.sqrt(hd)                 # (b, nh, t_new, p1)
        if causal is not None:
            score = score.masked_fill(~causal, -10000.)                                 # the reference's mask value
        o = torch.matmul(torch.softmax(score, dim=-1), vals)
        o = o.transpose(1, 2).reshape(b, t_new, nh * hd)
        y = comm_ops.all_reduce_sum_(F.linear(o, _w(cache, a.wo.weight, dtype)))
        x = x + (y + _w(cache, a.wo.bias, dtype))
        h = _rmsnorm(x, layer.norm2)
        g = F.linear(h, _w(cache, f.gate_proj.weight, dtype), _w(cache, f.gate_proj.bias, dtype))
        u = F.linear(h, _w(cache, f.up_proj.weight, dtype), _w(cache, f.up_proj.bias, dtype))
        y = comm_ops.all_reduce_sum_(F.linear(F.silu(g) * u, _w(cache, f.down_proj.weight, dtype)))
        x = x + (y + _w(cache, f.down_proj.bias, dtype))
    cache.length = p1
    return _rmsnorm(x[:, -1], model.norm)


@torch.no_grad()
def vocab_parallel_argmax(model, h_last: torch.Tensor, cache: KVCache, dtype: torch.dtype) -> torch.Tensor:
    """arg-max over the full vocabulary of ``lm_head(h_last)`` without gathering the logits: (b,) int64."""
    head = model.lm_head
    logits = F.linear(h_last, _w(cache, head.weight, dtype), _w(cache, head.bias, dtype)).float()   # (b, V/n)
    val, idx = logits.max(dim=-1)
    n, rank = pm.pgm.tp_size, pm.pgm.tp_rank
    if n == 1:
        return idx
    idx = idx + rank * head.odim_partition
    both = comm_ops.all_gather_rows(torch.stack((val, idx.to(val.dtype)), dim=-1))       # (n, b, 2)
    # ties resolve to the lowest rank = lowest vocabulary index, like arg-max over the gathered logits
    best = both[..., 0].argmax(dim=0)                                                     # (b,)
    return both[best, torch.arange(both.size(1), device=both.device), 1].to(torch.long)


@torch.no_grad()
def greedy_generate(model, tokens: torch.Tensor, eos_id: int, max_decode_len: int,
                    dtype: torch.dtype = torch.float32) -> Tuple[torch.Tensor, bool]:
    """KV-cached equivalent of the reference's decode loop (test.py:143-160) for ONE sequence ``tokens`` (1, t):
    append arg-max tokens until EOS or until the sequence is longer than ``max_decode_len``.
    Returns (tokens without the EOS, hit_eos)."""
    assert tokens.dim() == 2 and tokens.size(0) == 1
    device = tokens.device
    capacity = min(model.maxlen, max(tokens.size(1), max_decode_len) + 1)
    cache = KVCache.allocate(model, 1, capacity, dtype, device)
    h_last = forward_cached(model, tokens, cache, dtype)
    while True:
        nxt = vocab_parallel_argmax(model, h_last, cache, dtype)                         # (1,)
        pred = int(nxt.item())
        tokens = torch.cat((tokens, nxt.view(1, 1)), dim=1)
        if pred == eos_id or tokens.size(-1) > max_decode_len:
            hit_eos = pred == eos_id
            return (tokens[:, :-1] if hit_eos else tokens), hit_eos
        if cache.length >= cache.capacity:      # RoPE table exhausted (the reference would index out of range here)
            return tokens, False
        h_last = forward_cached(model, nxt.view(1, 1), cache, dtype)


@torch.no_grad()
def greedy_generate_batch(model, tokens: torch.Tensor, eos_id: int, max_new_tokens: int,
                          dtype: torch.dtype = torch.float32) -> torch.Tensor:
    """Batched variant for equal-length prompts (b, t): every sequence advances together, finished ones keep
    emitting EOS.  Returns (b, t + generated)."""
    device = tokens.device
    b, t = tokens.shape
    cache = KVCache.allocate(model, b, min(model.maxlen, t + max_new_tokens), dtype, device)
    done = torch.zeros(b, dtype=torch.bool, device=device)
    h_last = forward_cached(model, tokens, cache, dtype)
    for _ in range(max_new_tokens):
        nxt = vocab_parallel_argmax(model, h_last, cache, dtype)
        nxt = torch.where(done, torch.full_like(nxt, eos_id), nxt)
        tokens = torch.cat((tokens, nxt.unsqueeze(1)), dim=1)
        done |= nxt == eos_id
        if bool(done.all()) or cache.length >= cache.capacity:
            break
        h_last = forward_cached(model, nxt.unsqueeze(1), cache, dtype)
    return tokens
