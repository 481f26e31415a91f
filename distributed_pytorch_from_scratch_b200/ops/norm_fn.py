"""Autograd wrappers for the fused LayerNorm and activation kernels (standalone module use)."""
from __future__ import annotations

import torch

from . import fused


class NativeLayerNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, eps):
        xc = x.contiguous()
        if xc.dtype not in (torch.float32, torch.bfloat16):
            xc = xc.float()
        y, mean, rstd = fused.layernorm_fwd(xc, gamma.detach().float(),
                                            None if beta is None else beta.detach().float(), eps)
        ctx.save_for_backward(xc, gamma, mean, rstd)
        ctx.has_beta = beta is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        xc, gamma, mean, rstd = ctx.saved_tensors
        dyc = dy.contiguous()
        if dyc.dtype not in (torch.float32, torch.bfloat16):
            dyc = dyc.float()
        dx, dgamma, dbeta = fused.layernorm_bwd(dyc, xc, mean, rstd, gamma.detach().float(), ctx.has_beta)
        return dx.to(xc.dtype), dgamma.to(gamma.dtype), dbeta, None


def native_layernorm(x, gamma, beta, eps):
    return NativeLayerNormFn.apply(x, gamma, beta, eps)


class SwiGLUFn(torch.autograd.Function):
    """act = silu(gu[:, :F]) * gu[:, F:] on a packed bf16 [M, 2F] tensor."""

    @staticmethod
    def forward(ctx, gu):
        ctx.save_for_backward(gu)
        return fused.swiglu_fwd(gu)

    @staticmethod
    def backward(ctx, dact):
        (gu,) = ctx.saved_tensors
        return fused.swiglu_bwd(dact.contiguous(), gu)


class GELUFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        ctx.save_for_backward(x)
        return fused.gelu_fwd(x)

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        return fused.gelu_bwd(dy.contiguous(), x)


def swiglu(gu):
    return SwiGLUFn.apply(gu)


def gelu(x):
    return GELUFn.apply(x)
