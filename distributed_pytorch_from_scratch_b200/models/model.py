"""LLaMA-style tensor-parallel decoder (RMSNorm, RoPE, MHA, SwiGLU, untied vocab-parallel head) and its
single-device oracle.  Same constructor, submodule names, state_dict keys and init order as the reference
(models/model.py:17-158).

Two execution paths share the same parameters:
  * module path  — layer by layer through the parallel layers (PyTorch ops on CPU / fp32: the parity oracle);
  * fused path   — on CUDA in bf16 mode the whole block runs as hand-scheduled sm_100a kernels
                   (models/fused_transformer.py): fused QKV / gate-up GEMMs on tcgen05, RoPE, flash attention,
                   residual+bias GEMM epilogues, fused norms, vocab-parallel cross-entropy.
"""
from __future__ import annotations

import math
import os
from typing import Optional, Tuple

import torch
import torch.nn as nn
from torch.nn import functional as F

from ..ops import dispatch
from ..parallel import process_manager as pm
from .layers import ColumnParallelLinear, ParallelVocabularyEmbedding, RMSNorm, RowParallelLinear


# ------------------------------------------------------------------------------------------------ RoPE helpers
def rotate_half(x: torch.Tensor) -> torch.Tensor:
    """cat(-x[d/2:], x[:d/2]) — the HF / NeoX "half" convention (reference models/model.py:17-21)."""
    x1 = x[..., :x.shape[-1] // 2]
    x2 = x[..., x.shape[-1] // 2:]
    return torch.cat((-x2, x1), dim=-1)


def apply_rotary_pos_emb(q: torch.Tensor, k: torch.Tensor, cos: torch.Tensor,
                         sin: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """q, k: (b, n, t, d); cos, sin: (b, t, d).  ref: models/model.py:25-31."""
    assert cos.ndim == 3 and sin.ndim == 3, \
        f"Expected cos and sin to be 3D tensors (batch_size, seq_length, head_dim), got {cos.ndim} and {sin.ndim}"
    cos = cos.unsqueeze(1)
    sin = sin.unsqueeze(1)
    q_embed = (q * cos) + (rotate_half(q) * sin)
    k_embed = (k * cos) + (rotate_half(k) * sin)
    return q_embed, k_embed


def _env_dtype(default=None):
    v = os.getenv("DTYPE")
    if v is None:
        return default
    return torch.bfloat16 if v == "bfloat16" else torch.float32


def _env_device(default=None):
    v = os.getenv("DEVICE")
    if v is None:
        return default
    return torch.device("cuda") if v == "cuda" else torch.device("cpu")


def get_cos_sin(seq_length: int, head_dim: int, base: float, device=None,
                dtype=None) -> Tuple[torch.Tensor, torch.Tensor]:
    """(seq_length, head_dim) cos / sin tables; theta is computed on the CPU like the reference
    (models/model.py:35-46).  ``device``/``dtype`` default to the DEVICE / DTYPE env vars the reference uses, then to
    the process-group device and fp32."""
    assert head_dim % 2 == 0
    theta = 1.0 / (base ** (torch.arange(0, head_dim, 2, dtype=torch.int64).float().to("cpu") / head_dim))
    if dtype is None:
        dtype = _env_dtype(torch.float32)
    if device is None:
        device = _env_device(None)
        if device is None:
            device = pm.pgm.device if pm.pgm is not None else torch.device("cpu")
        if device.type == "cuda" and not torch.cuda.is_available():
            device = torch.device("cpu")
    position = torch.arange(seq_length).to(device).unsqueeze(1).float()
    theta = theta.to(device)
    cos = torch.cos(position.float() * theta.float()).to(dtype).repeat(1, 2)
    sin = torch.sin(position.float() * theta.float()).to(dtype).repeat(1, 2)
    return cos, sin


# ------------------------------------------------------------------------------------------------ blocks
def _causal_attention(q, k, v, head_dim):
    """The reference's attention math (models/model.py:73-77): additive -10000 mask, softmax, PV."""
    t = q.size(-2)
    attn_score = torch.matmul(q, k.transpose(-1, -2)) / math.sqrt(head_dim)
    mask = torch.triu(torch.ones(t, t, device=q.device, dtype=torch.bool), diagonal=1)
    attn_score = attn_score.masked_fill(mask, -10000.)
    attn = torch.softmax(attn_score, dim=-1)
    return torch.matmul(attn, v)


class Attention(nn.Module):
    def __init__(self, attn_dim: int, num_heads: int):
        super().__init__()
        self.attn_dim = attn_dim
        self.num_heads = num_heads
        self.head_dim = attn_dim // num_heads
        n = pm.pgm.tp_size
        assert num_heads % n == 0 and attn_dim % num_heads == 0
        self.num_local_heads = num_heads // n
        self.wq = ColumnParallelLinear(attn_dim, attn_dim, gather_output=False)
        self.wk = ColumnParallelLinear(attn_dim, attn_dim, gather_output=False)
        self.wv = ColumnParallelLinear(attn_dim, attn_dim, gather_output=False)
        self.wo = RowParallelLinear(attn_dim, attn_dim, split_input=False)

    def reset_parameters(self):
        self.wq.reset_parameters()
        self.wk.reset_parameters()
        self.wv.reset_parameters()
        self.wo.reset_parameters()

    def forward(self, x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor) -> torch.Tensor:
        b, t, _ = x.size()
        nh, hd = self.num_local_heads, self.head_dim
        q, k, v = self.wq(x), self.wk(x), self.wv(x)
        q, k, v = (z.view(b, t, nh, hd).transpose(1, 2) for z in (q, k, v))          # b t (n d) -> b n t d
        q, k = apply_rotary_pos_emb(q, k, cos, sin)
        o = _causal_attention(q, k, v.to(q.dtype), hd)
        o = o.transpose(1, 2).reshape(b, t, nh * hd)                                   # b n t d -> b t (n d)
        return self.wo(o)


class FFN(nn.Module):
    def __init__(self, idim: int, hdim: int):
        super().__init__()
        self.idim, self.hdim = idim, hdim
        self.gate_proj = ColumnParallelLinear(idim, hdim, gather_output=False)
        self.up_proj = ColumnParallelLinear(idim, hdim, gather_output=False)
        self.down_proj = RowParallelLinear(hdim, idim, split_input=False)

    def reset_parameters(self):
        self.gate_proj.reset_parameters()
        self.up_proj.reset_parameters()
        self.down_proj.reset_parameters()

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.down_proj(F.silu(self.gate_proj(x)) * self.up_proj(x))


class DecoderLayer(nn.Module):
    """Pre-norm residual block.  ref: models/model.py:98-121.  The RoPE tables are shared by all layers of a
    Transformer (reference defect D8 kept one copy per layer as plain attributes)."""

    def __init__(self, attn_dim: int, ffn_dim: int, num_heads: int, maxlen: int = 2048, rope_theta: float = 10000.,
                 rope_owner: Optional[nn.Module] = None):
        super().__init__()
        assert attn_dim % num_heads == 0
        self.head_dim = attn_dim // num_heads
        self.attn = Attention(attn_dim, num_heads)
        self.ffn = FFN(attn_dim, ffn_dim)
        self.norm1 = RMSNorm(attn_dim)
        self.norm2 = RMSNorm(attn_dim)
        self._rope_owner = [rope_owner] if rope_owner is not None else None   # list: keep it out of nn.Module
        if rope_owner is None:
            cos, sin = get_cos_sin(maxlen, self.head_dim, rope_theta)
            self.register_buffer("_cos", cos, persistent=False)
            self.register_buffer("_sin", sin, persistent=False)

    @property
    def cos(self):
        return self._rope_owner[0].rope_cos if self._rope_owner is not None else self._cos

    @property
    def sin(self):
        return self._rope_owner[0].rope_sin if self._rope_owner is not None else self._sin

    def reset_parameters(self):
        self.attn.reset_parameters()
        self.ffn.reset_parameters()

    def forward(self, x: torch.Tensor, position_ids: torch.Tensor) -> torch.Tensor:
        with torch.no_grad():
            cos, sin = self.cos[position_ids], self.sin[position_ids]      # (b, t, head_dim)
        x = x + self.attn(self.norm1(x), cos, sin)
        x = x + self.ffn(self.norm2(x))
        return x


class Transformer(nn.Module):
    """ref: models/model.py:124-158.  ``forward(input_ids, position_ids) -> (b, t, V)`` logits on every rank;
    ``loss(input_ids, position_ids, target_ids)`` is the vocab-parallel fused cross-entropy entry that never
    materialises the gathered logits (used by train.py)."""

    def __init__(self, attn_dim: int, ffn_dim: int, num_heads: int, num_layers: int, vocab_size: int,
                 maxlen: int = 2048, rope_theta: float = 10000., sequence_parallel: bool = False):
        super().__init__()
        self.attn_dim, self.ffn_dim, self.num_heads, self.num_layers = attn_dim, ffn_dim, num_heads, num_layers
        self.vocab_size, self.maxlen, self.rope_theta = vocab_size, maxlen, rope_theta
        self.head_dim = attn_dim // num_heads
        self.sequence_parallel = sequence_parallel
        self.compute_dtype = None        # None: follow DTYPE env / autocast (reference behaviour)
        cos, sin = get_cos_sin(maxlen, self.head_dim, rope_theta, dtype=torch.float32)
        self.register_buffer("rope_cos_f32", cos, persistent=False)
        self.register_buffer("rope_sin_f32", sin, persistent=False)
        self.embedding = ParallelVocabularyEmbedding(vocab_size, attn_dim)
        self.layers = nn.ModuleList([
            DecoderLayer(attn_dim, ffn_dim, num_heads, maxlen, rope_theta, rope_owner=self)
            for _ in range(num_layers)
        ])
        self.norm = RMSNorm(attn_dim)
        self.lm_head = ColumnParallelLinear(attn_dim, vocab_size, gather_output=True)
        self._fused = None               # models.fused_transformer.FusedRuntime, built lazily on CUDA

    # -- RoPE tables seen by the module path (dtype follows the compute dtype like the reference's env switch)
    @property
    def rope_cos(self):
        return self.rope_cos_f32.to(self._dtype())

    @property
    def rope_sin(self):
        return self.rope_sin_f32.to(self._dtype())

    def _dtype(self) -> torch.dtype:
        if self.compute_dtype is not None:
            return self.compute_dtype
        return _env_dtype(torch.float32)

    def reset_parameters(self):
        self.embedding.reset_parameters()
        for layer in self.layers:
            layer.reset_parameters()
        self.lm_head.reset_parameters()

    def retain_grad(self):
        # test helper (reference models/model.py:145-149)
        for _, param in self.named_parameters():
            if param.requires_grad:
                param.retain_grad()

    # ------------------------------------------------------------------------------------------ fused path
    def _fused_ok(self, input_ids: torch.Tensor) -> bool:
        if not input_ids.is_cuda or not dispatch.native_available():
            return False
        if self.embedding.weight.dtype != torch.float32 or not self.embedding.weight.is_cuda:
            return False
        bf16 = self._dtype() == torch.bfloat16 or (
            torch.is_autocast_enabled("cuda") and torch.get_autocast_dtype("cuda") == torch.bfloat16)
        if not bf16:
            return False
        from .fused_transformer import fused_supported
        return fused_supported(self)

    def fused_runtime(self):
        from .fused_transformer import FusedRuntime
        if self._fused is None or not self._fused.valid():
            self._fused = FusedRuntime(self)
        return self._fused

    # ------------------------------------------------------------------------------------------ forward / loss
    def forward(self, input_ids: torch.Tensor, position_ids: torch.Tensor) -> torch.Tensor:
        if self._fused_ok(input_ids):
            return self.fused_runtime().forward_logits(input_ids, position_ids)
        x = self.embedding(input_ids)
        x = x.to(self._dtype())     # explicitly convert token embedding to the compute dtype if necessary
        for layer in self.layers:
            x = layer(x, position_ids)
        logits = self.lm_head(self.norm(x))
        return logits

    def loss(self, input_ids: torch.Tensor, position_ids: torch.Tensor, target_ids: torch.Tensor,
             ignore_index: int = -1) -> torch.Tensor:
        """Mean cross-entropy over non-ignored targets (reference train.py:101-104)."""
        if self._fused_ok(input_ids):
            return self.fused_runtime().forward_loss(input_ids, position_ids, target_ids, ignore_index)
        logits = self.forward(input_ids, position_ids)
        return F.cross_entropy(logits.float().view(-1, logits.size(-1)), target_ids.view(-1),
                               ignore_index=ignore_index, reduction="mean")


# ------------------------------------------------------------------------------------------------ oracle
class _VanillaAttention(nn.Module):
    def __init__(self, attn_dim: int, num_heads: int):
        super().__init__()
        self.num_heads, self.head_dim = num_heads, attn_dim // num_heads
        self.wq = nn.Linear(attn_dim, attn_dim)
        self.wk = nn.Linear(attn_dim, attn_dim)
        self.wv = nn.Linear(attn_dim, attn_dim)
        self.wo = nn.Linear(attn_dim, attn_dim)

    def forward(self, x, cos, sin):
        b, t, _ = x.size()
        q, k, v = (lin(x).view(b, t, self.num_heads, self.head_dim).transpose(1, 2)
                   for lin in (self.wq, self.wk, self.wv))
        q, k = apply_rotary_pos_emb(q, k, cos, sin)
        o = _causal_attention(q, k, v.to(q.dtype), self.head_dim)
        return self.wo(o.transpose(1, 2).reshape(b, t, -1))


class _VanillaFFN(nn.Module):
    def __init__(self, idim: int, hdim: int):
        super().__init__()
        self.gate_proj = nn.Linear(idim, hdim)
        self.up_proj = nn.Linear(idim, hdim)
        self.down_proj = nn.Linear(hdim, idim)

    def forward(self, x):
        return self.down_proj(F.silu(self.gate_proj(x)) * self.up_proj(x))


class _VanillaLayer(nn.Module):
    def __init__(self, attn_dim, ffn_dim, num_heads):
        super().__init__()
        self.attn = _VanillaAttention(attn_dim, num_heads)
        self.ffn = _VanillaFFN(attn_dim, ffn_dim)
        self.norm1 = RMSNorm(attn_dim)
        self.norm2 = RMSNorm(attn_dim)

    def forward(self, x, cos, sin):
        x = x + self.attn(self.norm1(x), cos, sin)
        x = x + self.ffn(self.norm2(x))
        return x


class VallinaTransformer(nn.Module):
    """(sic) Single-device oracle with plain nn.Linear / nn.Embedding and the same math; the reference's own test
    imports this class but never ships it (defect D1, tests/test_transformers.py:14).  ``reset_parameters`` draws
    from the RNG in the same order as ``Transformer.reset_parameters`` so that, from the same RNG state, every TP
    shard equals the corresponding slice of these weights."""

    def __init__(self, attn_dim: int, ffn_dim: int, num_heads: int, num_layers: int, vocab_size: int,
                 maxlen: int = 2048, rope_theta: float = 10000.):
        super().__init__()
        self.head_dim = attn_dim // num_heads
        self.embedding = nn.Embedding(vocab_size, attn_dim)
        self.layers = nn.ModuleList([_VanillaLayer(attn_dim, ffn_dim, num_heads) for _ in range(num_layers)])
        self.norm = RMSNorm(attn_dim)
        self.lm_head = nn.Linear(attn_dim, vocab_size)
        cos, sin = get_cos_sin(maxlen, self.head_dim, rope_theta, dtype=torch.float32)
        self.register_buffer("rope_cos", cos, persistent=False)
        self.register_buffer("rope_sin", sin, persistent=False)

    @torch.no_grad()
    def reset_parameters(self):
        nn.init.normal_(self.embedding.weight, mean=0., std=1.)
        for layer in self.layers:
            for lin in (layer.attn.wq, layer.attn.wk, layer.attn.wv, layer.attn.wo, layer.ffn.gate_proj,
                        layer.ffn.up_proj, layer.ffn.down_proj):
                nn.init.kaiming_uniform_(lin.weight, a=math.sqrt(5))
                nn.init.zeros_(lin.bias)
        nn.init.kaiming_uniform_(self.lm_head.weight, a=math.sqrt(5))
        nn.init.zeros_(self.lm_head.bias)

    def retain_grad(self):
        for _, param in self.named_parameters():
            if param.requires_grad:
                param.retain_grad()

    def forward(self, input_ids: torch.Tensor, position_ids: torch.Tensor) -> torch.Tensor:
        x = self.embedding(input_ids)
        with torch.no_grad():
            cos = self.rope_cos[position_ids].to(x.dtype)
            sin = self.rope_sin[position_ids].to(x.dtype)
        for layer in self.layers:
            x = layer(x, cos, sin)
        return self.lm_head(self.norm(x))


VanillaTransformer = VallinaTransformer
