"""Autograd wrapper around the tcgen05 GEMM for a single linear layer (standalone use of the parallel layers).

    y = x W^T (+ b)         forward : gemm_nt(x, W)                       K-major / K-major
    dx = dy W               dgrad   : gemm_nt(dy, W, b_mn_major=True)     no transpose of W in memory
    dW = dy^T x             wgrad   : gemm_nt(dy, x, a_mn_major, b_mn_major), fp32 output, split-K over tokens
    db = colsum(dy)

ref: F.linear + autograd in models/layers.py:49,93.
"""
from __future__ import annotations

import torch

from . import fused
from .dispatch import shadow_bf16
from .gemm import gemm_nt


def _as_bf16_2d(x: torch.Tensor) -> torch.Tensor:
    x2 = x.reshape(-1, x.shape[-1])
    if x2.dtype != torch.bfloat16:
        x2 = x2.to(torch.bfloat16)
    return x2.contiguous()


class NativeLinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias):
        x2 = _as_bf16_2d(x)
        w = shadow_bf16(weight)
        b = None
        if bias is not None:
            b = bias.detach() if bias.dtype == torch.float32 else bias.detach().float()
        y = gemm_nt(x2, w, bias=b, out_dtype=torch.bfloat16)
        ctx.save_for_backward(x2, weight, bias)
        ctx.x_shape = x.shape
        ctx.x_dtype = x.dtype
        return y.view(*x.shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, weight, bias = ctx.saved_tensors
        dy2 = _as_bf16_2d(dy)
        w = shadow_bf16(weight)
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = gemm_nt(dy2, w, b_mn_major=True, out_dtype=torch.bfloat16)
            dx = dx.view(ctx.x_shape).to(ctx.x_dtype)
        if ctx.needs_input_grad[1]:
            dw = gemm_nt(dy2, x2, a_mn_major=True, b_mn_major=True, out_dtype=torch.float32)
            if weight.dtype != torch.float32:
                dw = dw.to(weight.dtype)
        if bias is not None and ctx.needs_input_grad[2]:
            db = fused.colsum(dy2)
            if bias.dtype != torch.float32:
                db = db.to(bias.dtype)
        return dx, dw, db


def native_linear(x, weight, bias=None):
    return NativeLinearFn.apply(x, weight, bias)


class NativeRMSNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, scale, eps):
        xc = x.contiguous()
        if xc.dtype not in (torch.float32, torch.bfloat16):
            xc = xc.float()
        y, rstd = fused.rmsnorm_fwd(xc, scale.detach().float(), eps, out_dtype=torch.bfloat16)
        ctx.save_for_backward(xc, scale, rstd)
        return y

    @staticmethod
    def backward(ctx, dy):
        xc, scale, rstd = ctx.saved_tensors
        dyc = dy.contiguous()
        if dyc.dtype not in (torch.float32, torch.bfloat16):
            dyc = dyc.float()
        dx, _, dscale = fused.rmsnorm_bwd(dyc, xc, rstd, scale.detach().float())
        return dx.view(xc.shape).to(xc.dtype), dscale.to(scale.dtype), None


def native_rmsnorm(x, scale, eps):
    return NativeRMSNormFn.apply(x, scale, eps)
