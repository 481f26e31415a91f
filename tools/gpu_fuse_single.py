"""Single-GPU timing of the FUSE_AR GEMM variant with every cross-rank action disabled (DPFS_FUSE_DEBUG=15)."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_pytorch_from_scratch_b200.ops import _lib
from distributed_pytorch_from_scratch_b200.ops.gemm import gemm_nt
os.environ["DPFS_FUSE_DEBUG"] = "15"
lib = _lib.lib()
M, N, K = 32000, 512, 256
a = torch.randn(M, K, device="cuda").to(torch.bfloat16); w = torch.randn(N, K, device="cuda").to(torch.bfloat16)
out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
flags = torch.zeros(lib.dpfs_comm_flag_words(), dtype=torch.int32, device="cuda")
data_ptrs = (ctypes.c_ulonglong * 8)(out.data_ptr(), out.data_ptr(), 0, 0, 0, 0, 0, 0)
flag_ptrs = (ctypes.c_ulonglong * 8)(flags.data_ptr(), flags.data_ptr(), 0, 0, 0, 0, 0, 0)
def fused():
    rc = lib.dpfs_gemm_allreduce_bf16(a.data_ptr(), K, w.data_ptr(), K, 0, out.data_ptr(), 0, data_ptrs, 0, M, N, K,
                                      flag_ptrs, 0, 2, 0, _lib.stream())
    assert rc == 0, rc
def plain():
    gemm_nt(a, w, out=out, splits=1)
def timeit(fn, iters=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3
print(f"plain {timeit(plain):.1f} us, fused-variant(no comm) {timeit(fused):.1f} us", flush=True)
ref = a.float() @ w.float().t()
fused(); torch.cuda.synchronize()
print("fused variant GEMM err", (out.float() - ref).abs().max().item())
