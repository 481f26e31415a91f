"""Communication autograd ops of Megatron-style 1-D tensor parallelism (reference models/comm_ops.py:7-83).

    Copy   : identity forward, all-reduce backward        ("copy to the TP region")
    Reduce : all-reduce forward, identity backward        ("reduce from the TP region")
    Split  : keep the local last-dim chunk forward, all-gather backward
    Gather : all-gather along the last dim forward, keep the local chunk backward

Copy/Reduce and Split/Gather are conjugate pairs.  All ops are the identity at TP=1.  Unlike the reference the
collectives never mutate their autograd inputs (its in-place all_reduce without mark_dirty is reference defect D9).

Data plane: on CUDA with the symmetric-memory communicator enabled (``pgm.symm``) the collectives run as our own
NVLink kernels (parallel/symm.py, csrc/comm.cu); otherwise torch.distributed (NCCL or gloo) is used, which is also
the correctness oracle for the fused paths.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from . import process_manager as pm


def _tp():
    return pm.pgm.tp_size, pm.pgm.tp_rank, pm.pgm.tp_group


def all_reduce_sum(t: torch.Tensor) -> torch.Tensor:
    """Out-of-place SUM all-reduce over the TP group."""
    n, _, group = _tp()
    if n == 1:
        return t
    symm = pm.pgm.symm
    if symm is not None and t.is_cuda and symm.supports(t):
        return symm.all_reduce(t)
    out = t.clone(memory_format=torch.contiguous_format)
    dist.all_reduce(out, op=dist.ReduceOp.SUM, group=group)
    return out


def _all_gather_stacked(t: torch.Tensor, n: int, group) -> torch.Tensor:
    """(...) -> (n, ...) in rank order; one flat buffer, no per-rank list (the reference builds a zeros list + cat)."""
    t = t.contiguous()
    symm = pm.pgm.symm
    if symm is not None and t.is_cuda and symm.can_gather(t):
        return symm.all_gather_rows(t)          # our own NVLink all-gather kernel (multimem.st / peer stores)
    flat = t.reshape(-1)
    out = torch.empty(n * flat.numel(), dtype=t.dtype, device=t.device)
    dist.all_gather_into_tensor(out, flat, group=group)
    return out.view((n,) + tuple(t.shape))


def all_gather_last_dim(t: torch.Tensor) -> torch.Tensor:
    n, _, group = _tp()
    if n == 1:
        return t
    gathered = _all_gather_stacked(t, n, group)
    # (n, ..., c) -> (..., n*c)
    return torch.cat(list(gathered.unbind(0)), dim=-1)


def local_chunk_last_dim(t: torch.Tensor) -> torch.Tensor:
    n, r, _ = _tp()
    if n == 1:
        return t
    dim = t.size(-1)
    assert dim % n == 0, f"last dim {dim} is not divisible by tp_size {n}"
    c = dim // n
    return t[..., r * c:(r + 1) * c].contiguous()


class Split(torch.autograd.Function):
    """(..., idim) -> (..., idim/n) forward; all-gather of the grads along the last dim backward."""

    @staticmethod
    def forward(ctx, input: torch.Tensor) -> torch.Tensor:
        return local_chunk_last_dim(input)

    @staticmethod
    def backward(ctx, grad_output: torch.Tensor) -> torch.Tensor:
        return all_gather_last_dim(grad_output)


class Reduce(torch.autograd.Function):
    """Sum over the TP group forward; identity backward."""

    @staticmethod
    def forward(ctx, input: torch.Tensor) -> torch.Tensor:
        return all_reduce_sum(input)

    @staticmethod
    def backward(ctx, grad_output: torch.Tensor) -> torch.Tensor:
        return grad_output


class Copy(torch.autograd.Function):
    """Identity forward; sum of the grads over the TP group backward."""

    @staticmethod
    def forward(ctx, input: torch.Tensor) -> torch.Tensor:
        return input.view_as(input) if pm.pgm.tp_size > 1 else input

    @staticmethod
    def backward(ctx, grad_output: torch.Tensor) -> torch.Tensor:
        return all_reduce_sum(grad_output)


class Gather(torch.autograd.Function):
    """(..., idim/n) -> (..., idim) forward (rank-ordered concat); local chunk of the grads backward."""

    @staticmethod
    def forward(ctx, input: torch.Tensor) -> torch.Tensor:
        return all_gather_last_dim(input)

    @staticmethod
    def backward(ctx, grad_output: torch.Tensor) -> torch.Tensor:
        return local_chunk_last_dim(grad_output)


def all_reduce_sum_(t: torch.Tensor) -> torch.Tensor:
    """In-place SUM all-reduce of a tensor we own (never an autograd input); returns ``t``."""
    n, _, group = _tp()
    if n == 1:
        return t
    symm = pm.pgm.symm
    if symm is not None and t.is_cuda and symm.supports(t):
        return symm.all_reduce(t, out=t)
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t


def all_gather_rows(t: torch.Tensor) -> torch.Tensor:
    """Gather a tensor from every TP rank along a new leading dim: (...) -> (n, ...)."""
    n, _, group = _tp()
    if n == 1:
        return t.unsqueeze(0)
    return _all_gather_stacked(t, n, group)
