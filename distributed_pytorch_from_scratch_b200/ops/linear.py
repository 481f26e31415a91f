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


# ------------------------------------------------------------------------------------------------ tensor-parallel forms
# The same three GEMMs with the layer's collective INSIDE the GEMM kernel (csrc/gemm_tcgen05.cu, FUSE_AR): finished
# tiles are summed across the ranks over NVLink by comm warps of the same CTAs while the tensor cores compute the next
# tiles.  Used by the public layer classes when the symmetric-memory communicator is active (pgm.symm).
def _symm():
    from ..parallel import process_manager as pm
    return pm.pgm.symm if pm.pgm is not None else None


class RowParallelLinearFn(torch.autograd.Function):
    """y = all_reduce(x_r W_r^T): forward GEMM + all-reduce in one kernel; backward is local (identity through the
    reduce).  The replicated bias is added by the caller after the reduce, like the reference
    (models/layers.py:49-54, models/comm_ops.py:39)."""

    @staticmethod
    def forward(ctx, x, weight):
        x2 = _as_bf16_2d(x)
        w = shadow_bf16(weight)
        symm = _symm()
        y = symm.gemm_allreduce(x2, w).clone()      # the result is a view of the symmetric buffer: take a copy
        ctx.save_for_backward(x2, weight)
        ctx.x_shape, ctx.x_dtype = x.shape, x.dtype
        return y.view(*x.shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, weight = ctx.saved_tensors
        dy2 = _as_bf16_2d(dy)
        w = shadow_bf16(weight)
        dx = dw = None
        if ctx.needs_input_grad[0]:
            dx = gemm_nt(dy2, w, b_mn_major=True, out_dtype=torch.bfloat16).view(ctx.x_shape).to(ctx.x_dtype)
        if ctx.needs_input_grad[1]:
            dw = gemm_nt(dy2, x2, a_mn_major=True, b_mn_major=True, out_dtype=torch.float32)
            if weight.dtype != torch.float32:
                dw = dw.to(weight.dtype)
        return dx, dw


class ColumnParallelLinearFn(torch.autograd.Function):
    """y_r = x W_r^T + b_r; backward: dx = all_reduce(dy_r W_r) as ONE kernel (dgrad GEMM + all-reduce), replacing the
    separate Copy.backward all-reduce of the reference (models/comm_ops.py:59 <- models/layers.py:91)."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        x2 = _as_bf16_2d(x)
        w = shadow_bf16(weight)
        b = None
        if bias is not None:
            b = bias.detach() if bias.dtype == torch.float32 else bias.detach().float()
        y = gemm_nt(x2, w, bias=b, out_dtype=torch.bfloat16)
        ctx.save_for_backward(x2, weight, bias)
        ctx.x_shape, ctx.x_dtype = x.shape, x.dtype
        return y.view(*x.shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, weight, bias = ctx.saved_tensors
        dy2 = _as_bf16_2d(dy)
        w = shadow_bf16(weight)
        symm = _symm()
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            if symm is not None and symm.can_fuse_gemm_ar(dy2, w, b_mn_major=True):
                dx = symm.gemm_allreduce(dy2, w, b_mn_major=True).clone()
            else:
                from ..parallel.comm_ops import all_reduce_sum_
                dx = all_reduce_sum_(gemm_nt(dy2, w, b_mn_major=True, out_dtype=torch.bfloat16))
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


def can_fuse_row_parallel(x, weight) -> bool:
    """True when RowParallelLinear.forward can run as the one-kernel GEMM + all-reduce."""
    symm = _symm()
    K = x.shape[-1]
    return (symm is not None and K % 8 == 0 and weight.shape[1] == K
            and symm.can_fuse_shape(x.numel() // K, weight.shape[0]))


class VocabParallelEmbeddingFn(torch.autograd.Function):
    """Vocab-parallel lookup + its collective (reference models/layers.py:134-141).  With NVLS multicast the owner of
    each token scatters the bf16 row into every rank's output (one kernel, nothing is summed); otherwise masked lookup
    kernel + all-reduce.  Backward: scatter-add of the owned rows into the local shard."""

    @staticmethod
    def forward(ctx, ids, weight, st, ed):
        from ..parallel.comm_ops import all_reduce_sum_
        symm = _symm()
        flat = ids.reshape(-1).contiguous()
        D = weight.shape[1]
        M = flat.numel()
        if symm is not None and symm.push_enabled and M * D * 2 <= symm.push_zone_bytes and (M * D * 2) % 16 == 0:
            out = fused.embedding_fwd_scatter(flat, weight.detach(), st, ed, symm).float()
        else:
            out = all_reduce_sum_(fused.embedding_fwd(flat, weight.detach(), st, ed, out_dtype=torch.float32))
        ctx.save_for_backward(flat)
        ctx.range, ctx.vshape = (st, ed), weight.shape
        return out.view(*ids.shape, D)

    @staticmethod
    def backward(ctx, dout):
        (flat,) = ctx.saved_tensors
        st, ed = ctx.range
        d2 = dout.reshape(-1, dout.shape[-1]).contiguous()
        if d2.dtype not in (torch.float32, torch.bfloat16):
            d2 = d2.float()
        dw = fused.embedding_bwd(flat, d2, ctx.vshape[0], st, ed)
        return None, dw, None, None


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
