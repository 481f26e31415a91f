"""Sequence-parallel form of BASELINE.json config 5: RowParallelLinear forward as GEMM + reduce-scatter over rows.

For every (M, K, N), device-timed with CUDA events (max over ranks), both reduce data paths (p2p / nvls):
    fused   : ONE kernel, symm.gemm_reduce_scatter (csrc/gemm_tcgen05.cu, rs mode)
    split   : our 2-SM tcgen05 GEMM into the symmetric buffer + our stand-alone reduce-scatter kernel (today's default)
    ref     : cuBLAS F.linear (bf16) + NCCL reduce_scatter_tensor
and checks the fused result against the split one before timing.

    torchrun --nproc-per-node N --master-addr 127.0.0.1 tools/bench_rowpar_rs.py [--out profiles/rowpar_rs_tpN.md]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import torch.nn.functional as F

ap = argparse.ArgumentParser()
ap.add_argument("--out", default=None)
args = ap.parse_args()
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
import process_manager as pm
pm.init_pgm(world)
from distributed_pytorch_from_scratch_b200.ops.gemm import gemm_nt
from distributed_pytorch_from_scratch_b200.parallel.symm import try_enable_symm
symm = try_enable_symm(required=True)


def timeit(fn, iters=15):
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


# the default model's row-parallel / dgrad shapes at 32000 tokens (K is the full reduction dim, split over the ranks)
shapes = [(32000, 512, 512), (32000, 2048, 512), (32000, 1536, 512), (32000, 4096, 512), (8192, 6144, 2048)]
rows = []
for mode in (["p2p", "nvls"] if symm.mc_base else ["p2p"]):
    symm.mode = mode
    for (M, K, N) in shapes:
        if M % world or M // world < 128 or M * N * 2 > symm.ag_off:
            continue
        Kl = max(64, K // world)
        a = torch.randn(M, Kl, device="cuda").to(torch.bfloat16)
        w = torch.randn(N, Kl, device="cuda").to(torch.bfloat16)
        buf = symm.buffer((M, N), torch.bfloat16)

        def split():
            gemm_nt(a, w, out=buf)
            return symm.reduce_scatter_rows(buf)

        def fused():
            return symm.gemm_reduce_scatter(a, w)

        def ref():
            y = F.linear(a, w)
            o = torch.empty(M // world, N, device="cuda", dtype=torch.bfloat16)
            dist.reduce_scatter_tensor(o, y)
            return o

        want, got = split().float(), fused().float()
        torch.cuda.synchronize()
        err = (want - got).abs().max().item()
        ok = err <= 0.02 * want.abs().max().item() + 1e-3
        t_f, t_s, t_r = timeit(fused), timeit(split), timeit(ref)
        rows.append((mode, M, K, N, t_f, t_s, t_r, ok))
        if rank == 0:
            print(f"[{mode}] M={M} K={K} N={N}: fused {t_f:.1f} us | split {t_s:.1f} us | cuBLAS+NCCL {t_r:.1f} us | "
                  f"{'OK' if ok else f'MISMATCH err={err:.3g}'}", flush=True)
if rank == 0 and args.out:
    with open(args.out, "w") as f:
        f.write(f"# Row-parallel GEMM + reduce-scatter over rows, TP={world}, bf16 (device-timed, max over ranks)\n\n")
        f.write("| reduce path | M | K | N | fused GEMM+RS (us) | 2-SM GEMM + RS kernel (us) | cuBLAS + NCCL (us) | matches |\n"
                "|---|---|---|---|---|---|---|---|\n")
        for (mode, M, K, N, tf, ts, tr, ok) in rows:
            f.write(f"| {mode} | {M} | {K} | {N} | {tf:.1f} | {ts:.1f} | {tr:.1f} | {'yes' if ok else 'NO'} |\n")
dist.barrier()
dist.destroy_process_group()
