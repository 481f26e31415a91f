"""bf16 GEMM on tcgen05 tensor cores (csrc/gemm_tcgen05.cu).

``gemm_nt`` computes ``out[M,N] = A[M,K] @ B[N,K]^T`` where each operand may be given K-major (reduction dim
contiguous) or MN-major (stored transposed, output dim contiguous) — the three cases a linear layer needs
(forward, dgrad, wgrad) without materialising a transpose.  Replaces the cuBLAS GEMMs behind ``F.linear`` in the
reference (models/layers.py:49,93).
"""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib

import os

_WS = {}
_WS_RETIRED = []   # outgrown split-K workspaces stay allocated (captured CUDA graphs hold their addresses)
_DEFAULT_BN = int(os.environ.get("DPFS_GEMM_BN", "256"))
_USE_2SM = os.environ.get("DPFS_GEMM_2SM", "1") == "1"


def _workspace(device, nfloats: int) -> torch.Tensor:
    # one workspace per (device, stream): GEMMs on different streams may run concurrently
    key = (device, torch.cuda.current_stream(device).cuda_stream)
    ws = _WS.get(key)
    if ws is None or ws.numel() < nfloats:
        if ws is not None:
            _WS_RETIRED.append(ws)
        ws = torch.empty(max(nfloats, 1 << 22), dtype=torch.float32, device=device)
        _WS[key] = ws
    return ws


def auto_splits(M: int, N: int, K: int, bn: int = 128) -> int:
    """Split-K factor that fills the 148 SMs when the output has few tiles (wgrad: tiny M x N, huge K)."""
    tiles = ((M + 127) // 128) * ((N + bn - 1) // bn)
    kb = (K + 63) // 64
    if tiles >= 120 or kb < 8:
        return 1
    s = max(1, min(kb // 4, (148 * 2) // tiles))
    return s


def gemm_nt(a: torch.Tensor, b: torch.Tensor, *, a_mn_major: bool = False, b_mn_major: bool = False,
            out: Optional[torch.Tensor] = None, out_dtype: torch.dtype = torch.bfloat16,
            bias: Optional[torch.Tensor] = None, residual: Optional[torch.Tensor] = None,
            splits: int = 0, bn: int = 0, two_sm: Optional[bool] = None) -> torch.Tensor:
    """out[M,N] = A[M,K] @ B[N,K]^T (+bias) (+residual).

    a: [M,K] (K-major) or [K,M] (``a_mn_major``); b: [N,K] or [K,N] (``b_mn_major``); 2-D, bf16, last dim
    contiguous (row stride may exceed the row length).
    """
    assert a.dim() == 2 and b.dim() == 2 and a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    assert a.stride(1) == 1 and b.stride(1) == 1
    if a_mn_major:
        K, M = a.shape
    else:
        M, K = a.shape
    if b_mn_major:
        Kb, N = b.shape
    else:
        N, Kb = b.shape
    assert K == Kb, f"reduction dims differ: {K} vs {Kb}"
    if out is None:
        out = torch.empty((M, N), dtype=out_dtype, device=a.device)
    else:
        assert out.shape == (M, N) and out.stride(1) == 1
    out_fp32 = out.dtype == torch.float32
    assert out_fp32 or out.dtype == torch.bfloat16
    if bias is not None:
        assert bias.dtype == torch.float32 and bias.numel() == N and bias.is_contiguous()
    if residual is not None:
        assert residual.dtype == torch.float32 and residual.shape == (M, N) and residual.stride(1) == 1
    if bn == 0:
        # 128x256 tiles halve the smem operand traffic per MMA (measured 5-40% faster than 128x128 on B200)
        bn = _DEFAULT_BN if (_DEFAULT_BN != 256 or N % 256 == 0) else 128
    user_splits = splits
    if splits == 0:
        splits = auto_splits(M, N, K, bn or 128) if residual is None else 1
    if two_sm is None:
        two_sm = _USE_2SM and M >= 256 and N >= 256
    if two_sm:
        if user_splits == 0:   # 74 CTA pairs: split K when the 256x256 tiles alone cannot fill them
            tiles2 = ((M + 255) // 256) * ((N + 255) // 256)
            kb = (K + 63) // 64
            splits = 1
            if residual is None and tiles2 < 60 and kb >= 8:
                splits = max(1, min(kb // 4, 148 // tiles2))
        ws = _workspace(a.device, splits * M * N) if splits > 1 else None
        rc = _lib.lib().dpfs_gemm2sm_bf16(
            a.data_ptr(), a.stride(0), int(a_mn_major), b.data_ptr(), b.stride(0), int(b_mn_major),
            out.data_ptr(), out.stride(0), int(out_fp32), _lib.ptr(bias), _lib.ptr(residual),
            residual.stride(0) if residual is not None else 0, M, N, K, splits, _lib.ptr(ws), _lib.stream())
        _lib.check(rc, "dpfs_gemm2sm_bf16")
        return out
    ws = _workspace(a.device, splits * M * N) if splits > 1 else None
    rc = _lib.lib().dpfs_gemm_bf16(
        a.data_ptr(), a.stride(0), int(a_mn_major), b.data_ptr(), b.stride(0), int(b_mn_major),
        out.data_ptr(), out.stride(0), int(out_fp32), _lib.ptr(bias), _lib.ptr(residual),
        residual.stride(0) if residual is not None else 0, M, N, K, splits, _lib.ptr(ws), bn, _lib.stream())
    _lib.check(rc, "dpfs_gemm_bf16")
    return out


_FUSE_SWIGLU = os.environ.get("DPFS_GEMM_SWIGLU", "1") == "1"


def gemm_swiglu(x: torch.Tensor, w_gate_up: torch.Tensor, bias: Optional[torch.Tensor]):
    """gu = x @ W^T + b (bf16 [M, 2F], gate columns then up columns) and act = silu(gate) * up (bf16 [M, F]).

    One kernel when the 2-SM GEMM's SwiGLU epilogue applies (F % 128 == 0, M >= 256: csrc/gemm2sm_tcgen05.cu); the
    activation is computed in the epilogue registers from the rounded gate / up values, so the stand-alone SwiGLU
    kernel's re-read of the [M, 2F] tensor disappears.  Otherwise GEMM + swiglu_fwd.  ref: models/model.py:95."""
    assert x.dim() == 2 and x.dtype == torch.bfloat16 and w_gate_up.dtype == torch.bfloat16
    M, K = x.shape
    F2 = w_gate_up.shape[0]
    F = F2 // 2
    if (_FUSE_SWIGLU and _USE_2SM and M >= 256 and F % 128 == 0 and x.stride(1) == 1 and w_gate_up.stride(1) == 1
            and x.stride(0) % 8 == 0 and w_gate_up.stride(0) % 8 == 0 and w_gate_up.shape[1] == K):
        gu = torch.empty((M, F2), dtype=torch.bfloat16, device=x.device)
        act = torch.empty((M, F), dtype=torch.bfloat16, device=x.device)
        if bias is not None:
            assert bias.dtype == torch.float32 and bias.numel() == F2 and bias.is_contiguous()
        rc = _lib.lib().dpfs_gemm2sm_swiglu_bf16(x.data_ptr(), x.stride(0), w_gate_up.data_ptr(), w_gate_up.stride(0),
                                                 gu.data_ptr(), act.data_ptr(), _lib.ptr(bias), M, F, K, _lib.stream())
        _lib.check(rc, "dpfs_gemm2sm_swiglu_bf16")
        return gu, act
    from . import fused
    gu = gemm_nt(x, w_gate_up, bias=bias, out_dtype=torch.bfloat16)
    return gu, fused.swiglu_fwd(gu)


def gemm_swiglu_bwd(dy: torch.Tensor, w_down: torch.Tensor, gu: torch.Tensor, bias_grad: torch.Tensor) -> torch.Tensor:
    """d(gate|up) = swiglu'(dy @ W_down, gu) (bf16 [M, 2F]) and its column sums into ``bias_grad`` (fp32 [2F]).

    One kernel when the 2-SM GEMM's SwiGLU-backward epilogue applies (F % 256 == 0, M >= 256): the dgrad accumulator
    dact = dy @ W_down never leaves the SM -- the epilogue reads the saved gate / up values and writes both halves of
    d(gate|up).  Otherwise dgrad GEMM + swiglu_bwd_bias.  ``w_down``: the down_proj weight as stored, [d, F]."""
    from . import fused
    M, K = dy.shape
    F = w_down.shape[1]
    assert gu.shape == (M, 2 * F) and gu.dtype == torch.bfloat16 and gu.is_contiguous()
    assert bias_grad.dtype == torch.float32 and bias_grad.numel() == 2 * F and bias_grad.is_contiguous()
    if (_FUSE_SWIGLU and _USE_2SM and M >= 256 and F % 256 == 0 and dy.dtype == torch.bfloat16
            and w_down.dtype == torch.bfloat16 and dy.stride(1) == 1 and w_down.stride(1) == 1
            and dy.stride(0) % 8 == 0 and w_down.stride(0) % 8 == 0 and w_down.shape[0] == K):
        lib = _lib.lib()
        dgu = torch.empty_like(gu)
        scratch = fused._scratch(dy.device, "swiglu_bwd_epi", lib.dpfs_gemm2sm_swiglu_bwd_rows(M) * 2 * F)
        rc = lib.dpfs_gemm2sm_swiglu_bwd_bf16(dy.data_ptr(), dy.stride(0), w_down.data_ptr(), w_down.stride(0),
                                              gu.data_ptr(), dgu.data_ptr(), bias_grad.data_ptr(), scratch.data_ptr(),
                                              M, F, K, _lib.stream())
        _lib.check(rc, "dpfs_gemm2sm_swiglu_bwd_bf16")
        return dgu
    dact = gemm_nt(dy, w_down, b_mn_major=True, out_dtype=torch.bfloat16)
    return fused.swiglu_bwd_bias(dact, gu, bias_grad)
