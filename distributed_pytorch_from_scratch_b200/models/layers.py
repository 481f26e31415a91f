"""Tensor-parallel layers: RowParallelLinear, ColumnParallelLinear, ParallelVocabularyEmbedding, RMSNorm, LayerNorm.

API, parameter names/shapes, initialisation order and checkpoint layout follow the reference
(models/layers.py:14-155); the math behind them is the sm_100a kernel library when running on CUDA in bf16 mode
(``ops.dispatch.use_native``) and plain PyTorch otherwise (CPU / fp32 oracle mode).

Real shapes (the reference's inline comments are wrong, defect D6):
    RowParallelLinear.weight    (odim, idim / n)   bias (odim,)   replicated, added after the reduce
    ColumnParallelLinear.weight (odim / n, idim)   bias (odim / n,)
    ParallelVocabularyEmbedding.weight (V / n, hdim)
"""
from __future__ import annotations

import math
from typing import Tuple

import torch
import torch.distributed as dist
import torch.nn as nn
from torch.nn import functional as F

from ..ops import dispatch
from ..parallel import process_manager as pm
from ..parallel.comm_ops import Copy, Gather, Reduce, Split


def _full_init_then_shard(param: torch.Tensor, full_shape, init_fn, shard_dim: int, part: int):
    """Init-parity contract (reference models/layers.py:33-42): draw the FULL weight with the current RNG on the
    parameter's device, broadcast rank 0's copy, keep this rank's slice — so TP=k and TP=1 start identical."""
    n, r = pm.pgm.tp_size, pm.pgm.tp_rank
    full = torch.empty(*full_shape, dtype=param.dtype, device=param.device, requires_grad=False)
    init_fn(full)
    if n > 1:
        if dist.is_initialized():
            dist.broadcast(full, src=0)
        full = torch.split(full, part, dim=shard_dim)[r]
    param.copy_(full.contiguous())


class RowParallelLinear(nn.Module):
    """y = sum_ranks(x_r W_r^T) + b with the input dim sharded.  ref: models/layers.py:14-55."""

    def __init__(self, idim: int, odim: int, add_bias: bool = True, split_input: bool = True):
        super().__init__()
        self.idim, self.odim = idim, odim
        self.split_input = split_input
        n = pm.pgm.tp_size
        assert idim % n == 0, f"idim {idim} must be divisible by tp_size {n}"
        self.idim_partition = idim // n
        self.weight = nn.Parameter(torch.empty(self.odim, self.idim_partition))
        self.add_bias = add_bias
        if add_bias:
            self.bias = nn.Parameter(torch.empty(self.odim))
        else:
            self.register_parameter("bias", None)

    @torch.no_grad()
    def reset_parameters(self):
        _full_init_then_shard(self.weight, (self.odim, self.idim),
                              lambda w: nn.init.kaiming_uniform_(w, a=math.sqrt(5)), -1, self.idim_partition)
        if self.add_bias:
            nn.init.zeros_(self.bias)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self.split_input:
            x = Split.apply(x)
        native = dispatch.use_native(x) and self.idim_partition % 8 == 0 and self.odim % 8 == 0
        if native and pm.pgm.tp_size > 1:
            from ..ops.linear import RowParallelLinearFn, can_fuse_row_parallel
            if can_fuse_row_parallel(x, self.weight):
                # ONE kernel: tcgen05 GEMM whose finished tiles are summed over NVLink while the next ones compute
                y = RowParallelLinearFn.apply(x, self.weight)
                return y + self.bias if self.add_bias else y
        if native:
            from ..ops.linear import native_linear
            y = native_linear(x, self.weight, None)
        else:
            y = F.linear(x, self.weight)
        y = Reduce.apply(y)
        if self.add_bias:
            y = y + self.bias
        return y

    def extra_repr(self):
        return f"idim={self.idim}, odim={self.odim}, idim_partition={self.idim_partition}, bias={self.add_bias}"


class ColumnParallelLinear(nn.Module):
    """y_r = x W_r^T + b_r with the output dim sharded (optionally gathered).  ref: models/layers.py:58-100."""

    def __init__(self, idim: int, odim: int, add_bias: bool = True, gather_output: bool = True):
        super().__init__()
        self.idim, self.odim = idim, odim
        self.gather_output = gather_output
        n = pm.pgm.tp_size
        assert odim % n == 0, f"odim {odim} must be divisible by tp_size {n}"
        self.odim_partition = odim // n
        self.weight = nn.Parameter(torch.empty(self.odim_partition, self.idim))
        self.add_bias = add_bias
        if add_bias:
            self.bias = nn.Parameter(torch.empty(self.odim_partition))
        else:
            self.register_parameter("bias", None)

    @torch.no_grad()
    def reset_parameters(self):
        _full_init_then_shard(self.weight, (self.odim, self.idim),
                              lambda w: nn.init.kaiming_uniform_(w, a=math.sqrt(5)), 0, self.odim_partition)
        if self.add_bias:
            nn.init.zeros_(self.bias)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        native = dispatch.use_native(x) and self.idim % 8 == 0 and self.odim_partition % 8 == 0
        if native and pm.pgm.tp_size > 1 and pm.pgm.symm is not None:
            # backward = dgrad GEMM + all-reduce in ONE kernel (takes the place of Copy.backward)
            from ..ops.linear import ColumnParallelLinearFn
            y = ColumnParallelLinearFn.apply(x, self.weight, self.bias if self.add_bias else None)
            return Gather.apply(y) if self.gather_output else y
        x = Copy.apply(x)
        if native:
            from ..ops.linear import native_linear
            y = native_linear(x, self.weight, self.bias if self.add_bias else None)  # bias fused in the epilogue
        else:
            y = F.linear(x, self.weight)
            if self.add_bias:
                y = y + self.bias
        if self.gather_output:
            y = Gather.apply(y)
        return y

    def extra_repr(self):
        return f"idim={self.idim}, odim={self.odim}, odim_partition={self.odim_partition}, bias={self.add_bias}"


class ParallelVocabularyEmbedding(nn.Module):
    """Vocabulary-sharded embedding: each rank looks up the ids it owns, others contribute zeros, then all-reduce.
    ref: models/layers.py:103-141.  The caller's ``input_ids`` are never modified (reference defect D3)."""

    def __init__(self, vocab_size: int, hdim: int):
        super().__init__()
        self.vocab_size = vocab_size
        self.hdim = hdim
        self.vocab_st_idx, self.vocab_ed_idx = self._get_vocab_range(vocab_size)
        self.weight = nn.Parameter(torch.empty(self.vocab_ed_idx - self.vocab_st_idx, self.hdim))

    @torch.no_grad()
    def reset_parameters(self):
        n, r = pm.pgm.tp_size, pm.pgm.tp_rank
        full = torch.empty(self.vocab_size, self.hdim, device=self.weight.device, dtype=self.weight.dtype)
        nn.init.normal_(full, mean=0., std=1.)
        if n > 1:
            if dist.is_initialized():
                dist.broadcast(full, src=0)
            full = full[self.vocab_st_idx:self.vocab_ed_idx]  # also correct for a non-divisible vocab (defect D4)
        self.weight.copy_(full.contiguous())

    def _get_vocab_range(self, vocab_size: int) -> Tuple[int, int]:
        n, r = pm.pgm.tp_size, pm.pgm.tp_rank
        assert n < vocab_size
        per = vocab_size // n
        st = per * r
        ed = st + per
        if r == n - 1 and ed != vocab_size:
            print(f"[WARNING]: In ParallelVocabularyEmbedding, the vocab size {vocab_size} is not divisible by "
                  f"tp_size {n}. The last partition will have {vocab_size - st} embeddings.")
            ed = vocab_size
        return st, ed

    def forward(self, x: torch.Tensor):
        assert x.ndim == 2, f"Input should be 2D tensor (B, L), but got {x.ndim}D tensor."
        if pm.pgm.tp_size == 1:
            return F.embedding(x, self.weight)
        if (x.is_cuda and dispatch.use_native(self.weight) and self.weight.dtype == torch.float32
                and self.hdim % 8 == 0 and self.vocab_size % pm.pgm.tp_size == 0):
            # lookup kernel (+ owner scatter through the NVSwitch when multicast is available) and native backward
            from ..ops.linear import VocabParallelEmbeddingFn
            return VocabParallelEmbeddingFn.apply(x, self.weight, self.vocab_st_idx, self.vocab_ed_idx)
        mask = torch.logical_and(self.vocab_st_idx <= x, x < self.vocab_ed_idx)
        local = torch.where(mask, x - self.vocab_st_idx, torch.zeros_like(x))
        out = F.embedding(local, self.weight)
        out = out * mask.unsqueeze(-1).to(out.dtype)
        return Reduce.apply(out)

    def extra_repr(self):
        return f"vocab_size={self.vocab_size}, hdim={self.hdim}, range=[{self.vocab_st_idx}, {self.vocab_ed_idx})"


ParallelVocabEmbedding = ParallelVocabularyEmbedding  # BASELINE.json spells it this way


class RMSNorm(nn.Module):
    """scale * x * rsqrt(mean(x^2) + eps), statistics in fp32; the parameter is called ``scale``.
    ref: models/layers.py:145-155."""

    def __init__(self, hdim: int, eps: float = 1e-5):
        super().__init__()
        self.eps = eps
        self.scale = nn.Parameter(torch.ones(hdim))

    def _norm(self, x: torch.Tensor) -> torch.Tensor:
        return x * (x.pow(2).mean(-1, keepdim=True) + self.eps).rsqrt()

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if dispatch.use_native(x) and x.shape[-1] % 4 == 0 and x.shape[-1] <= 2048:
            from ..ops.linear import native_rmsnorm
            return native_rmsnorm(x, self.scale, self.eps)
        return self.scale * self._norm(x.float()).type_as(x)


class LayerNorm(nn.Module):
    """Standard LayerNorm (weight/bias).  Not used by the reference model; provided for GPT-2 style blocks."""

    def __init__(self, hdim: int, eps: float = 1e-5, bias: bool = True):
        super().__init__()
        self.eps = eps
        self.weight = nn.Parameter(torch.ones(hdim))
        self.bias = nn.Parameter(torch.zeros(hdim)) if bias else None

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if dispatch.use_native(x) and x.shape[-1] % 4 == 0 and x.shape[-1] <= 2048:
            from ..ops.norm_fn import native_layernorm
            return native_layernorm(x, self.weight, self.bias, self.eps)
        return F.layer_norm(x.float(), (x.shape[-1],), self.weight, self.bias, self.eps).type_as(x)
