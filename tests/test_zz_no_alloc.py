"""The allocation-free column sum (dpfs_colsum_bf16_static) called straight through the C ABI.

It exists for hosts that cannot let this library allocate or lend it scratch — a process whose allocator already owns
the whole GPU, or a bare C caller without a device allocator — so the test talks to the shared object with ctypes only
(inputs and outputs are ordinary caller-owned buffers) and checks the result against the regular entry point, which
takes caller-provided scratch, on two back-to-back calls of different shape: the second call must not be affected by
what the first one left in the built-in scratch array.

The file name sorts last on purpose: if this check fails under ``pytest -x`` every other test has already run.
"""
import ctypes
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(REPO, "distributed_pytorch_from_scratch_b200", "lib", "libdpfs_b200.so")
MAX_FLOATS = 2 * 1024 * 1024      # DPFS_COLSUM_STATIC_MAX_FLOATS in include/dpfs_b200.h


def _bind():
    lib = ctypes.CDLL(LIB)
    vp, ll, i = ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int
    lib.dpfs_colsum_bf16_static.argtypes = [vp, ll, vp, i, i, i, vp]
    lib.dpfs_colsum_bf16_static.restype = i
    lib.dpfs_colsum_bf16.argtypes = [vp, ll, vp, i, vp, i, i, vp]
    lib.dpfs_colsum_bf16.restype = i
    lib.dpfs_colsum_scratch_floats.argtypes = [i, i]
    lib.dpfs_colsum_scratch_floats.restype = i
    return lib


def test_static_scratch_colsum_matches_regular_entry_point_twice():
    lib = _bind()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    torch.manual_seed(0)
    # two calls in a row; the second is smaller, so stale partial sums of the first sit right behind its own slots
    for (M, N) in [(3001, 520), (777, 264)]:
        x = torch.randn(M, N, device=dev).to(torch.bfloat16)
        out_static = torch.full((N,), float("nan"), device=dev)
        out_regular = torch.full((N,), float("nan"), device=dev)
        scratch = torch.empty(lib.dpfs_colsum_scratch_floats(M, N), dtype=torch.float32, device=dev)
        rc = lib.dpfs_colsum_bf16_static(x.data_ptr(), x.stride(0), out_static.data_ptr(), 0, M, N, stream)
        assert rc == 0, f"dpfs_colsum_bf16_static returned {rc}"
        rc = lib.dpfs_colsum_bf16(x.data_ptr(), x.stride(0), out_regular.data_ptr(), 0, scratch.data_ptr(), M, N,
                                  stream)
        assert rc == 0
        torch.cuda.synchronize()
        ref = x.float().sum(0)
        # same tolerance as tests/test_gpu_kernels.py::test_colsum_embedding_cast
        assert torch.allclose(out_static, out_regular, atol=1e-2, rtol=1e-4)
        assert torch.allclose(out_static, ref, atol=1e-2, rtol=1e-4)
    # accumulate flag behaves like the regular entry point
    x = torch.randn(500, 64, device=dev).to(torch.bfloat16)
    acc = torch.ones(64, device=dev)
    assert lib.dpfs_colsum_bf16_static(x.data_ptr(), x.stride(0), acc.data_ptr(), 1, 500, 64, stream) == 0
    torch.cuda.synchronize()
    assert torch.allclose(acc, 1.0 + x.float().sum(0), atol=1e-2, rtol=1e-4)


def test_static_scratch_colsum_rejects_oversize():
    lib = _bind()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    M, N = 64 * 1024 + 1, 4096                      # ceil(M / 64) * N > MAX_FLOATS
    assert ((M + 63) // 64) * N > MAX_FLOATS
    out = torch.zeros(N, device=dev)
    x = torch.zeros(8, N, device=dev, dtype=torch.bfloat16)   # never read: the size check comes first
    rc = lib.dpfs_colsum_bf16_static(x.data_ptr(), N, out.data_ptr(), 0, M, N, stream)
    assert rc == -25
    torch.cuda.synchronize()
    assert torch.count_nonzero(out).item() == 0
