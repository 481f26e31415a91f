"""ctypes binding of libdpfs_b200.so (C ABI declared in include/dpfs_b200.h).

The library is looked up in-tree (``distributed_pytorch_from_scratch_b200/lib``).  On a machine with a GPU a missing
library is an error — the framework never silently falls back to PyTorch ops for its CUDA path.
"""
from __future__ import annotations

import ctypes
import os
import threading

import torch

from ..build import LIB_PATH, build_library

_LIB = None
_LOCK = threading.Lock()

c_int = ctypes.c_int
c_ll = ctypes.c_longlong
c_f = ctypes.c_float
c_p = ctypes.c_void_p

# name -> argtypes (all entry points return int: 0 ok, >0 cudaError_t, <0 argument error)
SIGS = {
    "dpfs_gemm_bf16": (c_p, c_ll, c_int, c_p, c_ll, c_int, c_p, c_ll, c_int, c_p, c_p, c_ll, c_int, c_int, c_int,
                       c_int, c_p, c_int, c_p),
    "dpfs_gemm2sm_bf16": (c_p, c_ll, c_int, c_p, c_ll, c_int, c_p, c_ll, c_int, c_p, c_p, c_ll, c_int, c_int, c_int,
                          c_int, c_p, c_p),
    "dpfs_gemm2sm_swiglu_bf16": (c_p, c_ll, c_p, c_ll, c_p, c_p, c_p, c_int, c_int, c_int, c_p),
    "dpfs_gemm2sm_swiglu_bwd_bf16": (c_p, c_ll, c_p, c_ll, c_p, c_p, c_p, c_p, c_int, c_int, c_int, c_p),
    "dpfs_gemm2sm_swiglu_bwd_rows": (c_int,),
    "dpfs_gemm_allreduce_bf16": (c_p, c_ll, c_p, c_ll, c_int, c_p, c_p, c_p, c_ll, c_int, c_int, c_int, c_p, c_int,
                                 c_int, c_int, c_p),
    "dpfs_gemm_reduce_scatter_bf16": (c_p, c_ll, c_p, c_ll, c_int, c_p, c_p, c_p, c_ll, c_p, c_int, c_int, c_int, c_p,
                                      c_int, c_int, c_int, c_p),
    "dpfs_debug_rs_schedule": (c_int, c_int, c_p, c_p, c_p),
    "dpfs_debug_attn_items": (c_int, c_int, c_int, c_int, c_p, c_int),
    "dpfs_allreduce_p2p": (c_p, c_ll, c_ll, c_int, c_p, c_int, c_int, c_p),
    "dpfs_comm_flag_words": (),
    "dpfs_allreduce_nvls": (c_p, c_ll, c_int, c_p, c_int, c_int, c_p),
    "dpfs_reduce_scatter_p2p": (c_p, c_ll, c_ll, c_p, c_p, c_int, c_int, c_p),
    "dpfs_reduce_scatter_nvls": (c_p, c_ll, c_p, c_p, c_int, c_int, c_p),
    "dpfs_allgather_p2p": (c_p, c_p, c_ll, c_ll, c_p, c_int, c_int, c_p),
    "dpfs_allgather_nvls": (c_p, c_p, c_ll, c_p, c_int, c_int, c_p),
    "dpfs_comm_barrier": (c_p, c_int, c_int, c_p),
    "dpfs_allreduce_exp": (c_p, c_p, c_ll, c_ll, c_int, c_int, c_p, c_int, c_int, c_p),
    "dpfs_launch_count": (),
    "dpfs_add_rmsnorm_fwd": (c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_int, c_int, c_f, c_p),
    "dpfs_norm_bwd_blocks": (c_int,),
    "dpfs_rmsnorm_fwd": (c_p, c_int, c_p, c_p, c_int, c_p, c_int, c_int, c_f, c_p),
    "dpfs_rmsnorm_bwd": (c_p, c_int, c_p, c_int, c_p, c_p, c_p, c_p, c_p, c_p, c_int, c_p, c_int, c_int, c_p),
    "dpfs_rmsnorm_bwd_bias": (c_p, c_int, c_p, c_int, c_p, c_p, c_p, c_p, c_p, c_p, c_int, c_p, c_p, c_int, c_int,
                              c_p),
    "dpfs_rmsnorm_fwd_mc": (c_p, c_p, c_p, c_p, c_int, c_int, c_f, c_p),
    "dpfs_add_rmsnorm_fwd_mc": (c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_int, c_int, c_f, c_p),
    "dpfs_rmsnorm_bwd_mc": (c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_int, c_p, c_int, c_int, c_p),
    "dpfs_cast_f32_to_bf16_mc": (c_p, c_p, c_ll, c_p),
    "dpfs_layernorm_fwd": (c_p, c_int, c_p, c_p, c_p, c_int, c_p, c_p, c_int, c_int, c_f, c_p),
    "dpfs_layernorm_bwd": (c_p, c_int, c_p, c_int, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_int, c_int, c_p),
    "dpfs_swiglu_fwd": (c_p, c_p, c_int, c_int, c_p),
    "dpfs_swiglu_bwd": (c_p, c_p, c_p, c_int, c_int, c_p),
    "dpfs_swiglu_bwd_bias": (c_p, c_p, c_p, c_p, c_int, c_p, c_int, c_int, c_p),
    "dpfs_swiglu_bwd_scratch_floats": (c_int, c_int),
    "dpfs_cast_colsum": (c_p, c_p, c_p, c_int, c_p, c_int, c_int, c_p),
    "dpfs_gelu_fwd": (c_p, c_p, c_ll, c_p),
    "dpfs_gelu_bwd": (c_p, c_p, c_p, c_ll, c_p),
    "dpfs_rope_inplace": (c_p, c_ll, c_p, c_p, c_p, c_int, c_int, c_int, c_int, c_int, c_int, c_p),
    "dpfs_colsum_scratch_floats": (c_int, c_int),
    "dpfs_colsum_bf16": (c_p, c_ll, c_p, c_int, c_p, c_int, c_int, c_p),
    "dpfs_cast_f32_to_bf16": (c_p, c_p, c_ll, c_p),
    "dpfs_colsum_bf16_static": (c_p, c_ll, c_p, c_int, c_int, c_int, c_p),
    "dpfs_embedding_fwd": (c_p, c_p, c_p, c_int, c_int, c_int, c_ll, c_ll, c_p),
    "dpfs_embedding_fwd_mc": (c_p, c_p, c_p, c_int, c_int, c_ll, c_ll, c_p),
    "dpfs_embedding_bwd": (c_p, c_p, c_p, c_int, c_ll, c_int, c_int, c_ll, c_ll, c_p),
    "dpfs_embedding_bwd_bf16": (c_p, c_p, c_p, c_int, c_ll, c_int, c_int, c_ll, c_ll, c_p),
    "dpfs_rs_add_rmsnorm_ag": (c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_int, c_int, c_f, c_p, c_int, c_int, c_p),
    "dpfs_rs_rmsnorm_bwd_ag": (c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_int, c_p, c_int, c_int, c_p, c_int,
                               c_int, c_p),
    "dpfs_gemm_handoff_fwd": (c_p, c_ll, c_p, c_ll, c_int, c_p, c_p, c_int, c_int, c_int, c_p, c_p, c_p, c_p, c_p, c_p,
                              c_f, c_p, c_int, c_int, c_p),
    "dpfs_gemm_handoff_bwd": (c_p, c_ll, c_p, c_ll, c_int, c_p, c_p, c_int, c_int, c_int, c_p, c_p, c_p, c_p, c_p, c_p,
                              c_p, c_p, c_p, c_p, c_int, c_int, c_p),
    "dpfs_gemm_handoff_scratch_rows": (),
    "dpfs_debug_gh_schedule": (c_int, c_int, c_int, c_int, c_p, c_p, c_p, c_p, c_p, c_int),
    "dpfs_sp_blocks": (c_int, c_int),
    "dpfs_sp_debug_config": (c_int, c_int),
    "dpfs_comm_barrier_ex": (c_p, c_int, c_int, c_int, c_int, c_int, c_p),
    "dpfs_ce_stats": (c_p, c_int, c_ll, c_p, c_p, c_p, c_int, c_int, c_ll, c_ll, c_p),
    "dpfs_ce_loss": (c_p, c_p, c_int, c_p, c_p, c_int, c_ll, c_p),
    "dpfs_ce_grad": (c_p, c_int, c_ll, c_p, c_p, c_int, c_p, c_ll, c_p, c_p, c_int, c_int, c_ll, c_ll, c_p),
    "dpfs_adam_step": (c_p, c_p, c_p, c_p, c_p, c_ll, c_p, c_p),
    "dpfs_attn_fwd": (c_p, c_p, c_p, c_int, c_int, c_int, c_int, c_f, c_p),
    "dpfs_attn_bwd": (c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_int, c_int, c_int, c_int, c_f, c_p),
}


def _declare(handle):
    for name, argtypes in SIGS.items():
        if hasattr(handle, name):
            fn = getattr(handle, name)
            fn.argtypes = list(argtypes)
            fn.restype = ctypes.c_ulonglong if name == "dpfs_launch_count" else c_int


def available() -> bool:
    return os.path.exists(LIB_PATH)


def lib():
    global _LIB
    if _LIB is None:
        with _LOCK:
            if _LIB is None:
                if not os.path.exists(LIB_PATH):
                    build_library()
                handle = ctypes.CDLL(LIB_PATH)
                _declare(handle)
                _LIB = handle
    return _LIB


def stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def ptr(t) -> int:
    return 0 if t is None else t.data_ptr()


class KernelError(RuntimeError):
    pass


def check(rc: int, what: str):
    if rc != 0:
        raise KernelError(f"{what} failed with code {rc} (negative = library argument error, "
                          f"positive = cudaError_t)")
