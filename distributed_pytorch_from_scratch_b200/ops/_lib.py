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


def _declare(lib):
    def sig(name, *argtypes):
        fn = getattr(lib, name)
        fn.argtypes = list(argtypes)
        fn.restype = c_int
        return fn

    sig("dpfs_gemm_bf16", c_p, c_ll, c_int, c_p, c_ll, c_int, c_p, c_ll, c_int, c_p, c_p, c_ll, c_int, c_int,
        c_int, c_int, c_p, c_int, c_p)
    for name, argtypes in _EXTRA_SIGS.items():
        if hasattr(lib, name):
            sig(name, *argtypes)


# name -> argtypes; filled by the op modules at import time (before first load) or declared lazily.
_EXTRA_SIGS = {}


def register(name, *argtypes):
    _EXTRA_SIGS[name] = argtypes
    if _LIB is not None and hasattr(_LIB, name):
        fn = getattr(_LIB, name)
        fn.argtypes = list(argtypes)
        fn.restype = c_int


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
        raise KernelError(f"{what} failed with code {rc}"
                          + (" (negative = library argument error, positive = cudaError_t)" if rc else ""))
