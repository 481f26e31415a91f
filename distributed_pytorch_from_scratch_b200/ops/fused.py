"""Thin Python wrappers over the memory-bound sm_100a kernels (csrc/elementwise.cu, csrc/train_ops.cu).

Every function launches on the current CUDA stream and returns torch tensors; none of them synchronises.
Scratch buffers are cached per device (they are tiny) so steady-state steps do not allocate.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import _lib

_SCRATCH = {}
_RETIRED = []   # outgrown scratch buffers stay allocated: a CUDA graph captured earlier has their addresses baked in
BF16 = torch.bfloat16
F32 = torch.float32


def _scratch(device, key: str, nfloats: int) -> torch.Tensor:
    k = (device, key)
    buf = _SCRATCH.get(k)
    if buf is None or buf.numel() < nfloats:
        if buf is not None:
            # Never hand an outgrown buffer back to the caching allocator: TrainStep keeps one captured graph per batch
            # shape, and a graph captured for a smaller shape still writes its partial sums through the old pointer on
            # every replay.  The buffers are a few hundred KB; keeping them costs nothing.
            _RETIRED.append(buf)
        buf = torch.empty(max(nfloats, 1024), dtype=F32, device=device)
        _SCRATCH[k] = buf
    return buf


def _is_bf16(t: torch.Tensor) -> int:
    if t.dtype == BF16:
        return 1
    assert t.dtype == F32, f"expected fp32 or bf16, got {t.dtype}"
    return 0


# ------------------------------------------------------------------------------------------------ RMSNorm
def rmsnorm_fwd(x: torch.Tensor, scale: torch.Tensor, eps: float, out_dtype=BF16, want_rstd: bool = True):
    """x: [..., D] fp32 or bf16 (contiguous); scale fp32 [D]. Returns (y[out_dtype], rstd fp32 [rows] | None)."""
    D = x.shape[-1]
    x2 = x.reshape(-1, D)
    assert x2.is_contiguous() and scale.dtype == F32
    M = x2.shape[0]
    y = torch.empty(x2.shape, dtype=out_dtype, device=x.device)
    rstd = torch.empty(M, dtype=F32, device=x.device) if want_rstd else None
    rc = _lib.lib().dpfs_rmsnorm_fwd(x2.data_ptr(), _is_bf16(x2), scale.data_ptr(), y.data_ptr(), _is_bf16(y),
                                     _lib.ptr(rstd), M, D, eps, _lib.stream())
    _lib.check(rc, "dpfs_rmsnorm_fwd")
    return y.view(x.shape), rstd


def add_rmsnorm_fwd(x: torch.Tensor, y: torch.Tensor, bias: Optional[torch.Tensor], scale: Optional[torch.Tensor],
                    eps: float, want_norm: bool = True):
    """x_out = x + y + bias (fp32); h = rmsnorm(x_out) * scale (bf16).  x fp32 [M, D], y bf16 [M, D] (the all-reduced
    row-parallel GEMM output).  Returns (x_out, h | None, rstd | None)."""
    M, D = x.shape
    assert x.dtype == F32 and y.dtype == BF16 and x.is_contiguous() and y.is_contiguous() and y.shape == x.shape
    x_out = torch.empty_like(x)
    h = torch.empty((M, D), dtype=BF16, device=x.device) if want_norm else None
    rstd = torch.empty(M, dtype=F32, device=x.device) if want_norm else None
    rc = _lib.lib().dpfs_add_rmsnorm_fwd(x.data_ptr(), y.data_ptr(), _lib.ptr(bias),
                                         _lib.ptr(scale) if want_norm else 0, x_out.data_ptr(), _lib.ptr(h),
                                         _lib.ptr(rstd), M, D, eps, _lib.stream())
    _lib.check(rc, "dpfs_add_rmsnorm_fwd")
    return x_out, h, rstd


def rmsnorm_bwd(dy: torch.Tensor, x: torch.Tensor, rstd: torch.Tensor, scale: torch.Tensor,
                dres: Optional[torch.Tensor] = None, want_fp32: bool = True, want_bf16: bool = False,
                dscale_out: Optional[torch.Tensor] = None, accumulate: bool = False,
                dbias_out: Optional[torch.Tensor] = None):
    """Returns (dx fp32 | None, dx bf16 | None, dscale fp32 [D]).  dx = dres + rmsnorm'(dy).

    dbias_out (fp32 [D], overwritten): column sums of dx, i.e. the bias gradient of the row-parallel linear whose
    output gradient dx is (ref models/layers.py:51-54) -- produced by the same pass instead of a colsum of its own."""
    D = x.shape[-1]
    x2, dy2 = x.reshape(-1, D), dy.reshape(-1, D)
    assert x2.is_contiguous() and dy2.is_contiguous()
    M = x2.shape[0]
    dx = torch.empty((M, D), dtype=F32, device=x.device) if want_fp32 else None
    dxb = torch.empty((M, D), dtype=BF16, device=x.device) if want_bf16 else None
    dscale = dscale_out if dscale_out is not None else torch.empty(D, dtype=F32, device=x.device)
    lib = _lib.lib()
    scratch = _scratch(x.device, "norm", lib.dpfs_norm_bwd_blocks(M) * D * (2 if dbias_out is not None else 1))
    if dres is not None:
        dres = dres.reshape(-1, D)
        assert dres.dtype == F32 and dres.is_contiguous()
    if dbias_out is not None:
        assert dbias_out.dtype == F32 and dbias_out.numel() == D and dbias_out.is_contiguous()
        rc = lib.dpfs_rmsnorm_bwd_bias(dy2.data_ptr(), _is_bf16(dy2), x2.data_ptr(), _is_bf16(x2), rstd.data_ptr(),
                                       scale.data_ptr(), _lib.ptr(dres), _lib.ptr(dx), _lib.ptr(dxb),
                                       dscale.data_ptr(), int(accumulate), dbias_out.data_ptr(), scratch.data_ptr(),
                                       M, D, _lib.stream())
        _lib.check(rc, "dpfs_rmsnorm_bwd_bias")
        return dx, dxb, dscale
    rc = lib.dpfs_rmsnorm_bwd(dy2.data_ptr(), _is_bf16(dy2), x2.data_ptr(), _is_bf16(x2), rstd.data_ptr(),
                              scale.data_ptr(), _lib.ptr(dres), _lib.ptr(dx), _lib.ptr(dxb), dscale.data_ptr(),
                              int(accumulate), scratch.data_ptr(), M, D, _lib.stream())
    _lib.check(rc, "dpfs_rmsnorm_bwd")
    return dx, dxb, dscale


# ------------------------------------------------------------------------------------------------ push all-gather
# Sequence-parallel producers that all-gather their bf16 output while writing it (NVLS multicast stores into every
# rank's landing zone, parallel/symm_comm.py:push_block) instead of "write the local block, then all-gather kernel".
# Each returns the gathered [n * Ml, D] bf16 view (valid until the second-next push) after the rank barrier.
def rmsnorm_fwd_push(x_l: torch.Tensor, scale: torch.Tensor, eps: float, symm):
    Ml, D = x_l.shape
    assert x_l.dtype == F32 and x_l.is_contiguous()
    mc, full = symm.push_block(Ml, D)
    rstd = torch.empty(Ml, dtype=F32, device=x_l.device)
    rc = _lib.lib().dpfs_rmsnorm_fwd_mc(x_l.data_ptr(), scale.data_ptr(), mc, rstd.data_ptr(), Ml, D, eps,
                                        _lib.stream())
    _lib.check(rc, "dpfs_rmsnorm_fwd_mc")
    symm.push_done()
    return full, rstd


def add_rmsnorm_fwd_push(x_l: torch.Tensor, y_l: torch.Tensor, bias, scale: torch.Tensor, eps: float, symm):
    """x_out = x + y + bias (fp32, local rows); h = rmsnorm(x_out) * scale pushed to every rank.
    Returns (x_out [Ml, D], h_full [n * Ml, D] bf16, rstd [Ml])."""
    Ml, D = x_l.shape
    assert x_l.dtype == F32 and y_l.dtype == BF16 and x_l.is_contiguous() and y_l.is_contiguous()
    mc, full = symm.push_block(Ml, D)
    x_out = torch.empty_like(x_l)
    rstd = torch.empty(Ml, dtype=F32, device=x_l.device)
    rc = _lib.lib().dpfs_add_rmsnorm_fwd_mc(x_l.data_ptr(), y_l.data_ptr(), _lib.ptr(bias), scale.data_ptr(),
                                            x_out.data_ptr(), mc, rstd.data_ptr(), Ml, D, eps, _lib.stream())
    _lib.check(rc, "dpfs_add_rmsnorm_fwd_mc")
    symm.push_done()
    return x_out, full, rstd


def rmsnorm_bwd_push(dy_l: torch.Tensor, x_l: torch.Tensor, rstd: torch.Tensor, scale: torch.Tensor, dres, symm,
                     dscale_out: torch.Tensor, want_fp32: bool = True):
    """dx = dres + rmsnorm'(dy) on the local rows; its bf16 copy is pushed to every rank.
    Returns (dx fp32 [Ml, D] | None, dx_full bf16 [n * Ml, D])."""
    Ml, D = x_l.shape
    assert dy_l.dtype == BF16 and x_l.dtype == F32 and dy_l.is_contiguous() and x_l.is_contiguous()
    mc, full = symm.push_block(Ml, D)
    dx = torch.empty((Ml, D), dtype=F32, device=x_l.device) if want_fp32 else None
    lib = _lib.lib()
    scratch = _scratch(x_l.device, "norm", lib.dpfs_norm_bwd_blocks(Ml) * D)
    if dres is not None:
        assert dres.dtype == F32 and dres.is_contiguous()
    rc = lib.dpfs_rmsnorm_bwd_mc(dy_l.data_ptr(), x_l.data_ptr(), rstd.data_ptr(), scale.data_ptr(), _lib.ptr(dres),
                                 _lib.ptr(dx), mc, dscale_out.data_ptr(), 0, scratch.data_ptr(), Ml, D, _lib.stream())
    _lib.check(rc, "dpfs_rmsnorm_bwd_mc")
    symm.push_done()
    return dx, full


def embedding_fwd_scatter(ids_flat: torch.Tensor, weight: torch.Tensor, vocab_start: int, vocab_end: int, symm):
    """Vocab-parallel embedding + its collective in one kernel: the owner of each token writes the bf16 row into every
    rank's copy of the output (multicast store), nobody sums anything.  Returns the [M, D] bf16 view (landing zone)."""
    M, D = ids_flat.numel(), weight.shape[1]
    assert ids_flat.dtype == torch.int64 and ids_flat.is_contiguous() and weight.dtype == F32
    mc, out = symm.push_zone(M, D)
    rc = _lib.lib().dpfs_embedding_fwd_mc(ids_flat.data_ptr(), weight.data_ptr(), mc, M, D, vocab_start, vocab_end,
                                          _lib.stream())
    _lib.check(rc, "dpfs_embedding_fwd_mc")
    symm.push_done()
    return out


def cast_to_bf16_push(x_l: torch.Tensor, symm) -> torch.Tensor:
    """bf16(x_l) pushed to every rank: [Ml, D] fp32 -> gathered [n * Ml, D] bf16."""
    Ml, D = x_l.shape
    assert x_l.dtype == F32 and x_l.is_contiguous() and x_l.numel() % 4 == 0
    mc, full = symm.push_block(Ml, D)
    _lib.check(_lib.lib().dpfs_cast_f32_to_bf16_mc(x_l.data_ptr(), mc, x_l.numel(), _lib.stream()),
               "dpfs_cast_f32_to_bf16_mc")
    symm.push_done()
    return full


# ------------------------------------------------------------------------------------------------ SP hand-offs
# Sequence-parallel residual-stream hand-offs as ONE kernel each (csrc/sp_fused.cu): reduce-scatter of the row-parallel
# GEMM's partial products -> row op on this rank's rows -> all-gather of the bf16 result, through the NVSwitch
# (multimem.ld_reduce / multimem.st).  ``part`` is the [M, D] bf16 partial matrix in the symmetric buffer, ``zone`` a
# parallel.symm_comm.SymmZone that receives the gathered [M, D] bf16 result on every rank.
def rs_add_rmsnorm_ag(symm, part: torch.Tensor, x_l: Optional[torch.Tensor], bias: Optional[torch.Tensor],
                      scale: torch.Tensor, eps: float, zone):
    """x_out_l = x_l + reduce_scatter(part) + bias; h = all_gather(rmsnorm(x_out_l) * scale).
    Returns (x_out_l fp32 [M/n, D], h bf16 [M, D] (view of ``zone``), rstd fp32 [M/n])."""
    M, D = part.shape
    n = symm.n
    Ml = M // n
    assert part.dtype == BF16 and part.is_contiguous() and symm._inside(part) and M % n == 0 and symm.mc_base
    if x_l is not None:
        assert x_l.dtype == F32 and x_l.shape == (Ml, D) and x_l.is_contiguous()
    x_out = torch.empty((Ml, D), dtype=F32, device=part.device)
    rstd = torch.empty(Ml, dtype=F32, device=part.device)
    h = zone.view((M, D), BF16)
    mc_part = symm.mc_base + (part.data_ptr() - symm.base)
    rc = _lib.lib().dpfs_rs_add_rmsnorm_ag(mc_part, _lib.ptr(x_l), _lib.ptr(bias), scale.data_ptr(), x_out.data_ptr(),
                                           rstd.data_ptr(), zone.mc, Ml, D, eps, symm.flag_ptrs, symm.rank, n,
                                           _lib.stream())
    _lib.check(rc, "dpfs_rs_add_rmsnorm_ag")
    return x_out, h, rstd


def rs_rmsnorm_bwd_ag(symm, part: torch.Tensor, x_l: torch.Tensor, rstd: torch.Tensor, scale: torch.Tensor,
                      dres_l: Optional[torch.Tensor], zone, dscale_out: torch.Tensor,
                      dbias_out: Optional[torch.Tensor] = None, want_fp32: bool = True):
    """dx_l = dres_l + rmsnorm'(reduce_scatter(part)); its bf16 copy is all-gathered into ``zone``.
    dscale_out / dbias_out receive partial sums over THIS rank's rows.  Returns (dx_l | None, dx bf16 [M, D] view)."""
    M, D = part.shape
    n = symm.n
    Ml = M // n
    assert part.dtype == BF16 and part.is_contiguous() and symm._inside(part) and M % n == 0 and symm.mc_base
    assert x_l.dtype == F32 and x_l.shape == (Ml, D) and x_l.is_contiguous() and rstd.numel() == Ml
    if dres_l is not None:
        assert dres_l.dtype == F32 and dres_l.shape == (Ml, D) and dres_l.is_contiguous()
    lib = _lib.lib()
    dx_l = torch.empty((Ml, D), dtype=F32, device=part.device) if want_fp32 else None
    scratch = _scratch(part.device, "sp_norm", lib.dpfs_sp_blocks(Ml, D) * 2 * D)
    mc_part = symm.mc_base + (part.data_ptr() - symm.base)
    rc = lib.dpfs_rs_rmsnorm_bwd_ag(mc_part, x_l.data_ptr(), rstd.data_ptr(), scale.data_ptr(), _lib.ptr(dres_l),
                                    _lib.ptr(dx_l), zone.mc if zone is not None else 0, _lib.ptr(dscale_out),
                                    _lib.ptr(dbias_out), 0, scratch.data_ptr(), Ml, D, symm.flag_ptrs, symm.rank, n,
                                    _lib.stream())
    _lib.check(rc, "dpfs_rs_rmsnorm_bwd_ag")
    return dx_l, (zone.view((M, D), BF16) if zone is not None else None)


def can_gemm_handoff(symm, a: torch.Tensor, w: torch.Tensor, b_mn_major: bool = False) -> bool:
    """Whether GEMM + hand-off can run as ONE kernel (csrc/gemm_handoff.cu) for a @ w^T (or a @ w)."""
    M = a.shape[0]
    N = w.shape[1] if b_mn_major else w.shape[0]
    tiles = ((M + 127) // 128) * ((N + 255) // 256)
    return (bool(symm.mc_base) and not symm.closed and a.dtype == BF16 and w.dtype == BF16 and N % 8 == 0
            and N <= 1024 and M % symm.n == 0 and M // symm.n >= 128 and a.stride(0) % 8 == 0
            and w.stride(0) % 8 == 0 and a.stride(1) == 1 and w.stride(1) == 1 and tiles <= 16384
            and M * N * 2 <= symm.ag_off)


def gemm_handoff_fwd(symm, a: torch.Tensor, w: torch.Tensor, b_mn_major: bool, x_l: Optional[torch.Tensor],
                     bias: Optional[torch.Tensor], scale: torch.Tensor, eps: float, zone):
    """ONE kernel: x_out_l = x_l + reduce_scatter(a @ w^T) + bias; h = all_gather(rmsnorm(x_out_l) * scale).
    Returns (x_out_l fp32 [M/n, N], h bf16 [M, N] (view of ``zone``), rstd fp32 [M/n])."""
    M, K = a.shape
    N = w.shape[1] if b_mn_major else w.shape[0]
    n = symm.n
    Ml = M // n
    part = symm.buffer((M, N), BF16)
    mc_part = symm.mc_base + (part.data_ptr() - symm.base)
    x_out = torch.empty((Ml, N), dtype=F32, device=a.device)
    rstd = torch.empty(Ml, dtype=F32, device=a.device)
    rc = _lib.lib().dpfs_gemm_handoff_fwd(a.data_ptr(), a.stride(0), w.data_ptr(), w.stride(0), int(b_mn_major),
                                          part.data_ptr(), mc_part, M, N, K, _lib.ptr(x_l), _lib.ptr(bias),
                                          scale.data_ptr(), x_out.data_ptr(), rstd.data_ptr(), zone.mc, eps,
                                          symm.flag_ptrs, symm.rank, n, _lib.stream())
    _lib.check(rc, "dpfs_gemm_handoff_fwd")
    return x_out, zone.view((M, N), BF16), rstd


def gemm_handoff_bwd(symm, a: torch.Tensor, w: torch.Tensor, b_mn_major: bool, x_l: torch.Tensor, rstd: torch.Tensor,
                     scale: torch.Tensor, dres_l: Optional[torch.Tensor], zone, dscale_out: torch.Tensor,
                     dbias_out: Optional[torch.Tensor] = None):
    """ONE kernel: dx_l = dres_l + rmsnorm'(reduce_scatter(a @ w)); bf16(dx) all-gathered into ``zone``.
    Returns (dx_l fp32 [M/n, N], dx bf16 [M, N] view)."""
    M, K = a.shape
    N = w.shape[1] if b_mn_major else w.shape[0]
    n = symm.n
    Ml = M // n
    lib = _lib.lib()
    part = symm.buffer((M, N), BF16)
    mc_part = symm.mc_base + (part.data_ptr() - symm.base)
    dx_l = torch.empty((Ml, N), dtype=F32, device=a.device)
    scratch = _scratch(a.device, "gh_norm", lib.dpfs_gemm_handoff_scratch_rows() * 2 * N)
    rc = lib.dpfs_gemm_handoff_bwd(a.data_ptr(), a.stride(0), w.data_ptr(), w.stride(0), int(b_mn_major),
                                   part.data_ptr(), mc_part, M, N, K, x_l.data_ptr(), rstd.data_ptr(),
                                   scale.data_ptr(), _lib.ptr(dres_l), dx_l.data_ptr(), zone.mc,
                                   _lib.ptr(dscale_out), _lib.ptr(dbias_out), scratch.data_ptr(), symm.flag_ptrs,
                                   symm.rank, n, _lib.stream())
    _lib.check(rc, "dpfs_gemm_handoff_bwd")
    return dx_l, zone.view((M, N), BF16)


# ------------------------------------------------------------------------------------------------ LayerNorm
def layernorm_fwd(x, gamma, beta, eps: float, out_dtype=BF16):
    D = x.shape[-1]
    x2 = x.reshape(-1, D)
    M = x2.shape[0]
    y = torch.empty(x2.shape, dtype=out_dtype, device=x.device)
    mean = torch.empty(M, dtype=F32, device=x.device)
    rstd = torch.empty(M, dtype=F32, device=x.device)
    rc = _lib.lib().dpfs_layernorm_fwd(x2.data_ptr(), _is_bf16(x2), gamma.data_ptr(), _lib.ptr(beta), y.data_ptr(),
                                       _is_bf16(y), mean.data_ptr(), rstd.data_ptr(), M, D, eps, _lib.stream())
    _lib.check(rc, "dpfs_layernorm_fwd")
    return y.view(x.shape), mean, rstd


def layernorm_bwd(dy, x, mean, rstd, gamma, want_dbeta: bool = True):
    D = x.shape[-1]
    x2, dy2 = x.reshape(-1, D), dy.reshape(-1, D).contiguous()
    M = x2.shape[0]
    dx = torch.empty((M, D), dtype=F32, device=x.device)
    dgamma = torch.empty(D, dtype=F32, device=x.device)
    dbeta = torch.empty(D, dtype=F32, device=x.device) if want_dbeta else None
    lib = _lib.lib()
    scratch = _scratch(x.device, "lnorm", (lib.dpfs_norm_bwd_blocks(M) + 1) * 2 * D)
    rc = lib.dpfs_layernorm_bwd(dy2.data_ptr(), _is_bf16(dy2), x2.data_ptr(), _is_bf16(x2), mean.data_ptr(),
                                rstd.data_ptr(), gamma.data_ptr(), dx.data_ptr(), dgamma.data_ptr(),
                                _lib.ptr(dbeta), scratch.data_ptr(), M, D, _lib.stream())
    _lib.check(rc, "dpfs_layernorm_bwd")
    return dx.view(x.shape), dgamma, dbeta


# ------------------------------------------------------------------------------------------------ activations
def swiglu_fwd(gu: torch.Tensor) -> torch.Tensor:
    """gu: bf16 [M, 2F] (gate | up) -> silu(gate) * up, bf16 [M, F]."""
    M, F2 = gu.shape
    assert gu.dtype == BF16 and gu.is_contiguous()
    act = torch.empty((M, F2 // 2), dtype=BF16, device=gu.device)
    rc = _lib.lib().dpfs_swiglu_fwd(gu.data_ptr(), act.data_ptr(), M, F2 // 2, _lib.stream())
    _lib.check(rc, "dpfs_swiglu_fwd")
    return act


def swiglu_bwd(dact: torch.Tensor, gu: torch.Tensor) -> torch.Tensor:
    M, F2 = gu.shape
    assert dact.dtype == BF16 and dact.is_contiguous() and dact.shape == (M, F2 // 2)
    dgu = torch.empty_like(gu)
    rc = _lib.lib().dpfs_swiglu_bwd(dact.data_ptr(), gu.data_ptr(), dgu.data_ptr(), M, F2 // 2, _lib.stream())
    _lib.check(rc, "dpfs_swiglu_bwd")
    return dgu


def swiglu_bwd_bias(dact: torch.Tensor, gu: torch.Tensor, bias_grad: torch.Tensor) -> torch.Tensor:
    """swiglu backward + column sums of dgu into ``bias_grad`` (fp32 [2F]) in the same pass."""
    M, F2 = gu.shape
    assert dact.dtype == BF16 and dact.is_contiguous() and dact.shape == (M, F2 // 2)
    assert bias_grad.dtype == F32 and bias_grad.numel() == F2 and bias_grad.is_contiguous()
    dgu = torch.empty_like(gu)
    lib = _lib.lib()
    scratch = _scratch(gu.device, "colsum", lib.dpfs_swiglu_bwd_scratch_floats(M, F2 // 2))
    rc = lib.dpfs_swiglu_bwd_bias(dact.data_ptr(), gu.data_ptr(), dgu.data_ptr(), bias_grad.data_ptr(), 0,
                                  scratch.data_ptr(), M, F2 // 2, _lib.stream())
    _lib.check(rc, "dpfs_swiglu_bwd_bias")
    return dgu


def cast_colsum(x: torch.Tensor, colsum_out: torch.Tensor) -> torch.Tensor:
    """bf16(x) for fp32 x [M, N] and the column sums of the result into ``colsum_out`` (fp32 [N]), one pass."""
    assert x.dtype == F32 and x.dim() == 2 and x.is_contiguous()
    M, N = x.shape
    assert colsum_out.dtype == F32 and colsum_out.numel() == N and colsum_out.is_contiguous()
    y = torch.empty((M, N), dtype=BF16, device=x.device)
    lib = _lib.lib()
    scratch = _scratch(x.device, "colsum", ((M + 15) // 16) * N)
    rc = lib.dpfs_cast_colsum(x.data_ptr(), y.data_ptr(), colsum_out.data_ptr(), 0, scratch.data_ptr(), M, N,
                              _lib.stream())
    _lib.check(rc, "dpfs_cast_colsum")
    return y


def gelu_fwd(x: torch.Tensor) -> torch.Tensor:
    assert x.dtype == BF16 and x.is_contiguous()
    y = torch.empty_like(x)
    _lib.check(_lib.lib().dpfs_gelu_fwd(x.data_ptr(), y.data_ptr(), x.numel(), _lib.stream()), "dpfs_gelu_fwd")
    return y


def gelu_bwd(dy: torch.Tensor, x: torch.Tensor) -> torch.Tensor:
    assert dy.dtype == BF16 and dy.is_contiguous() and x.is_contiguous()
    dx = torch.empty_like(x)
    _lib.check(_lib.lib().dpfs_gelu_bwd(dy.data_ptr(), x.data_ptr(), dx.data_ptr(), x.numel(), _lib.stream()),
               "dpfs_gelu_bwd")
    return dx


# ------------------------------------------------------------------------------------------------ RoPE
def rope_inplace(buf: torch.Tensor, pos: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor, nseg: int,
                 seg_stride: int, H: int, hd: int, inverse: bool = False):
    """Rotate the first ``nseg`` segments (q, k) of a packed bf16 [M, ld] buffer in place.

    pos: int64 [M]; cos/sin: fp32 [maxlen, hd/2] tables."""
    assert buf.dtype == BF16 and buf.stride(1) == 1 and pos.dtype == torch.int64 and pos.is_contiguous()
    assert cos.dtype == F32 and cos.shape[-1] == hd // 2 and cos.is_contiguous() and sin.is_contiguous()
    rc = _lib.lib().dpfs_rope_inplace(buf.data_ptr(), buf.stride(0), pos.data_ptr(), cos.data_ptr(), sin.data_ptr(),
                                      buf.shape[0], nseg, seg_stride, H, hd, int(inverse), _lib.stream())
    _lib.check(rc, "dpfs_rope_inplace")
    return buf


# ------------------------------------------------------------------------------------------------ reductions
def colsum(x: torch.Tensor, out: Optional[torch.Tensor] = None, accumulate: bool = False) -> torch.Tensor:
    """Column sums of a bf16 [M, N] matrix in fp32 (bias gradients), deterministic two-stage reduction."""
    assert x.dtype == BF16 and x.dim() == 2 and x.stride(1) == 1
    M, N = x.shape
    if out is None:
        out = torch.empty(N, dtype=F32, device=x.device)
    lib = _lib.lib()
    scratch = _scratch(x.device, "colsum", lib.dpfs_colsum_scratch_floats(M, N))
    rc = lib.dpfs_colsum_bf16(x.data_ptr(), x.stride(0), out.data_ptr(), int(accumulate), scratch.data_ptr(), M, N,
                              _lib.stream())
    _lib.check(rc, "dpfs_colsum_bf16")
    return out


def cast_to_bf16(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    assert x.dtype == F32 and x.is_contiguous()
    if out is None:
        out = torch.empty(x.shape, dtype=BF16, device=x.device)
    if x.numel() % 4 != 0:
        out.copy_(x)
        return out
    _lib.check(_lib.lib().dpfs_cast_f32_to_bf16(x.data_ptr(), out.data_ptr(), x.numel(), _lib.stream()),
               "dpfs_cast_f32_to_bf16")
    return out


# ------------------------------------------------------------------------------------------------ embedding
def embedding_fwd(ids: torch.Tensor, weight: torch.Tensor, vocab_start: int, vocab_end: int, out_dtype=F32,
                  out: Optional[torch.Tensor] = None):
    assert ids.dtype == torch.int64 and weight.dtype == F32 and weight.is_contiguous()
    ids_flat = ids.reshape(-1).contiguous()
    M, D = ids_flat.numel(), weight.shape[1]
    if out is None:
        out = torch.empty((M, D), dtype=out_dtype, device=weight.device)
    assert out.shape == (M, D) and out.is_contiguous()
    rc = _lib.lib().dpfs_embedding_fwd(ids_flat.data_ptr(), weight.data_ptr(), out.data_ptr(), _is_bf16(out), M, D,
                                       vocab_start, vocab_end, _lib.stream())
    _lib.check(rc, "dpfs_embedding_fwd")
    return out


def embedding_bwd(ids: torch.Tensor, dout: torch.Tensor, local_vocab: int, vocab_start: int, vocab_end: int,
                  dweight: Optional[torch.Tensor] = None, zero_first: bool = True):
    ids_flat = ids.reshape(-1).contiguous()
    D = dout.shape[-1]
    dout2 = dout.reshape(-1, D)
    assert dout2.dtype in (F32, BF16) and dout2.is_contiguous()
    if dweight is None:
        dweight = torch.empty((local_vocab, D), dtype=F32, device=dout.device)
    fn = _lib.lib().dpfs_embedding_bwd if dout2.dtype == F32 else _lib.lib().dpfs_embedding_bwd_bf16
    rc = fn(ids_flat.data_ptr(), dout2.data_ptr(), dweight.data_ptr(), int(zero_first),
                                       local_vocab, ids_flat.numel(), D, vocab_start, vocab_end, _lib.stream())
    _lib.check(rc, "dpfs_embedding_bwd")
    return dweight


# ------------------------------------------------------------------------------------------------ cross-entropy
def ce_stats(logits: torch.Tensor, targets: torch.Tensor, vocab_start: int, ignore_index: int):
    """Phase 1: per-row (max, sum-exp, target-logit) over the local vocab shard; also counts valid targets."""
    M, Vl = logits.shape
    assert logits.stride(1) == 1 and targets.dtype == torch.int64 and targets.is_contiguous()
    stats = torch.empty((M, 3), dtype=F32, device=logits.device)
    loss_acc = torch.empty(2, dtype=F32, device=logits.device)  # {sum of row losses, number of valid rows}
    rc = _lib.lib().dpfs_ce_stats(logits.data_ptr(), _is_bf16(logits), logits.stride(0), targets.data_ptr(),
                                  stats.data_ptr(), loss_acc.data_ptr(), M, Vl, vocab_start, ignore_index,
                                  _lib.stream())
    _lib.check(rc, "dpfs_ce_stats")
    return stats, loss_acc


def ce_loss(targets: torch.Tensor, all_stats: torch.Tensor, nranks: int, loss_acc: torch.Tensor,
            ignore_index: int = -1, want_row_loss: bool = False):
    """Phase 2a: combine the per-rank stats; loss_acc[0] += sum of row losses (mean = loss_acc[0] / loss_acc[1])."""
    M = targets.numel()
    row_loss = torch.empty(M, dtype=F32, device=targets.device) if want_row_loss else None
    rc = _lib.lib().dpfs_ce_loss(targets.data_ptr(), all_stats.data_ptr(), nranks, loss_acc.data_ptr(),
                                 _lib.ptr(row_loss), M, ignore_index, _lib.stream())
    _lib.check(rc, "dpfs_ce_loss")
    return row_loss



def ce_grad(logits: torch.Tensor, targets: torch.Tensor, all_stats: torch.Tensor, nranks: int,
            loss_acc: torch.Tensor, vocab_start: int, ignore_index: int, upstream: Optional[torch.Tensor] = None):
    """Phase 2b: dlogits = (softmax - onehot) * upstream / n_valid as bf16 [M, Vl]."""
    M, Vl = logits.shape
    dlogits = torch.empty((M, Vl), dtype=BF16, device=logits.device)
    if upstream is not None:
        assert upstream.dtype == F32 and upstream.numel() == 1
    rc = _lib.lib().dpfs_ce_grad(logits.data_ptr(), _is_bf16(logits), logits.stride(0), targets.data_ptr(),
                                 all_stats.data_ptr(), nranks, dlogits.data_ptr(), Vl, loss_acc.data_ptr(),
                                 _lib.ptr(upstream), M, Vl, vocab_start, ignore_index, _lib.stream())
    _lib.check(rc, "dpfs_ce_grad")
    return dlogits


# ------------------------------------------------------------------------------------------------ optimizer
def adam_step(p: torch.Tensor, g: torch.Tensor, m: torch.Tensor, v: torch.Tensor, shadow: Optional[torch.Tensor],
              hyper: torch.Tensor):
    """One fused Adam/AdamW step over flat fp32 arenas; ``hyper`` is a device tensor
    [lr, beta1, beta2, eps, bias_correction1, bias_correction2, weight_decay]."""
    n = p.numel()
    assert p.dtype == F32 and g.dtype == F32 and m.dtype == F32 and v.dtype == F32 and hyper.dtype == F32
    assert g.numel() == n and m.numel() == n and v.numel() == n and n % 4 == 0
    rc = _lib.lib().dpfs_adam_step(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), _lib.ptr(shadow), n,
                                   hyper.data_ptr(), _lib.stream())
    _lib.check(rc, "dpfs_adam_step")
