"""Causal flash attention on tcgen05 (csrc/attention_tcgen05.cu).

Operates on the packed QKV activation ``[b*t, 3*Dl]`` produced by the fused QKV GEMM (q | k | v segments, ``H``
local heads of ``hd``), with RoPE already applied, and returns ``[b*t, Dl]`` ready for the ``wo`` GEMM: the
``b t (n d) <-> b n t d`` rearranges of the reference (models/model.py:71,77) are folded into TMA tensor maps.
"""
from __future__ import annotations

import math

import torch

from . import _lib


def attn_fwd(qkv: torch.Tensor, B: int, T: int, H: int, HD: int, scale: float):
    assert qkv.dtype == torch.bfloat16 and qkv.is_contiguous() and qkv.shape == (B * T, 3 * H * HD)
    out = torch.empty((B * T, H * HD), dtype=torch.bfloat16, device=qkv.device)
    lse = torch.empty((B, H, T), dtype=torch.float32, device=qkv.device)
    rc = _lib.lib().dpfs_attn_fwd(qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), B, T, H, HD, scale, _lib.stream())
    _lib.check(rc, "dpfs_attn_fwd")
    return out, lse


def attn_bwd(qkv, out, d_out, lse, B: int, T: int, H: int, HD: int, scale: float):
    assert d_out.dtype == torch.bfloat16 and d_out.is_contiguous()
    dqkv = torch.empty_like(qkv)
    delta = torch.empty((B, H, T), dtype=torch.float32, device=qkv.device)
    dq_acc = torch.empty((B * T, H * HD), dtype=torch.float32, device=qkv.device)
    rc = _lib.lib().dpfs_attn_bwd(qkv.data_ptr(), out.data_ptr(), d_out.data_ptr(), lse.data_ptr(), delta.data_ptr(),
                                  dq_acc.data_ptr(), dqkv.data_ptr(), B, T, H, HD, scale, _lib.stream())
    _lib.check(rc, "dpfs_attn_bwd")
    return dqkv


class FlashAttnQKVPacked(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, B, T, H, HD, scale):
        out, lse = attn_fwd(qkv, B, T, H, HD, scale)
        ctx.save_for_backward(qkv, out, lse)
        ctx.dims = (B, T, H, HD, scale)
        return out

    @staticmethod
    def backward(ctx, d_out):
        qkv, out, lse = ctx.saved_tensors
        B, T, H, HD, scale = ctx.dims
        dqkv = attn_bwd(qkv, out, d_out.contiguous(), lse, B, T, H, HD, scale)
        return dqkv, None, None, None, None, None


def flash_attn_qkvpacked(qkv: torch.Tensor, B: int, T: int, H: int, HD: int, scale: float | None = None):
    if scale is None:
        scale = 1.0 / math.sqrt(HD)
    return FlashAttnQKVPacked.apply(qkv, B, T, H, HD, scale)


def attention_reference(qkv: torch.Tensor, B: int, T: int, H: int, HD: int, scale: float | None = None):
    """Plain PyTorch fp32 oracle with the reference's exact masking semantics (models/model.py:73-77)."""
    if scale is None:
        scale = 1.0 / math.sqrt(HD)
    Dl = H * HD
    q, k, v = (qkv[:, i * Dl:(i + 1) * Dl].reshape(B, T, H, HD).permute(0, 2, 1, 3) for i in range(3))
    s = torch.matmul(q, k.transpose(-1, -2)) * scale
    mask = torch.triu(torch.ones(T, T, device=qkv.device), diagonal=1).bool()
    s = s.masked_fill(mask, -10000.)
    p = torch.softmax(s, dim=-1)
    o = torch.matmul(p, v)
    return o.permute(0, 2, 1, 3).reshape(B * T, Dl)
