"""Multi-GPU bring-up of the symmetric-memory collectives and the fused GEMM + all-reduce (run under torchrun)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
import process_manager as pm
pm.init_pgm(world)
from distributed_pytorch_from_scratch_b200.parallel.symm import try_enable_symm
from distributed_pytorch_from_scratch_b200.ops.gemm import gemm_nt

symm = try_enable_symm(required=True)


def log(*a):
    if rank == 0:
        print(*a, flush=True)


def timeit(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item() * 1e3


torch.manual_seed(100 + rank)
# ---- stand-alone all-reduce / all-gather vs NCCL
for dtype, shape in [(torch.bfloat16, (32000, 512)), (torch.float32, (4096, 512)), (torch.bfloat16, (1000, 64))]:
    x = torch.randn(shape, device="cuda").to(dtype)
    ref = x.clone()
    dist.all_reduce(ref)
    out = symm.all_reduce(x)
    torch.cuda.synchronize()
    err = (out.float() - ref.float()).abs().max().item()
    log(f"[allreduce] {dtype} {shape}: max err vs NCCL {err:.4g} {'OK' if err < 0.1 else 'FAIL'}")
    # in-place on the symmetric buffer, many times in a row (semaphore reuse)
    buf = symm.buffer(shape, dtype)
    for it in range(50):
        buf.copy_(x)
        symm.all_reduce(buf, out=buf)
    torch.cuda.synchronize()
    err = (buf.float() - ref.float()).abs().max().item()
    log(f"[allreduce x50 in place] err {err:.4g} {'OK' if err < 0.1 else 'FAIL'}")
for shape in [(32000, 512), (32000, 2048)]:
    buf = symm.buffer(shape, torch.bfloat16)
    buf.normal_()
    t_own = timeit(lambda: symm.all_reduce(buf, out=buf))
    y = torch.randn(shape, device="cuda").to(torch.bfloat16)
    t_nccl = timeit(lambda: dist.all_reduce(y))
    S = shape[0] * shape[1] * 2
    log(f"[allreduce time] {shape} bf16 ({S/1e6:.1f} MB): NVLS kernel {t_own:.1f} us ({S/t_own/1e3:.0f} GB/s algbw), "
        f"NCCL {t_nccl:.1f} us ({S/t_nccl/1e3:.0f} GB/s)")
st = torch.randn(32000, 3, device="cuda")
g = symm.all_gather_rows(st)
ref = torch.empty(world, 32000, 3, device="cuda")
dist.all_gather_into_tensor(ref.view(-1), st.view(-1))
log(f"[allgather] equal to NCCL: {torch.equal(g, ref)}")

# ---- fused GEMM + all-reduce vs GEMM + NCCL
for (M, N, K, bmn) in [(256, 128, 64, False), (1000, 512, 256, False), (32000, 512, 256, False),
                       (32000, 512, 1024, False), (32000, 512, 768, True), (32000, 512, 2048, True)]:
    a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
    w = (torch.randn(K, N, device="cuda") if bmn else torch.randn(N, K, device="cuda")).to(torch.bfloat16)
    ref = gemm_nt(a, w, b_mn_major=bmn, out_dtype=torch.float32)
    dist.all_reduce(ref)
    out = symm.gemm_allreduce(a, w, b_mn_major=bmn).clone()
    torch.cuda.synchronize()
    err = (out.float() - ref).abs().max().item()
    sc = ref.abs().max().item()
    ok = err < 0.03 * sc
    t_fused = timeit(lambda: symm.gemm_allreduce(a, w, b_mn_major=bmn))

    def unfused():
        y = gemm_nt(a, w, b_mn_major=bmn, out_dtype=torch.bfloat16)
        dist.all_reduce(y)
    t_nccl = timeit(unfused)
    t_gemm = timeit(lambda: gemm_nt(a, w, b_mn_major=bmn, out_dtype=torch.bfloat16))
    S = M * N * 2
    roof_comm = S * (1 + 1 / world) / 770e9 * 1e6
    roof_flop = 2 * M * N * K / 1.656e15 * 1e6
    log(f"[gemm+AR] M={M} N={N} K={K} bmn={bmn}: err {err:.4g}/{sc:.3g} {'OK' if ok else 'FAIL'} | fused {t_fused:.1f} us, "
        f"gemm+NCCL {t_nccl:.1f} us, gemm only {t_gemm:.1f} us | roofline max(comm {roof_comm:.1f}, flop {roof_flop:.1f}) us "
        f"-> {max(roof_comm, roof_flop) / t_fused * 100:.0f}% of roofline")

dist.barrier()
dist.destroy_process_group()
