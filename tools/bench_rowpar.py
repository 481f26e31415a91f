"""BASELINE.json config 5: RowParallelLinear forward (GEMM + all-reduce) microbench sweep, TP = world size.

For every (M, K, N) it times, device-side with CUDA events (max over ranks):
    fused   : our ONE-kernel tcgen05 GEMM + NVLink reduce (symm.gemm_allreduce)
    split   : our tcgen05 GEMM followed by our stand-alone NVLink all-reduce kernel
    ref     : the reference's path — cuBLAS F.linear (bf16) followed by NCCL all_reduce (models/layers.py:49-51)
and reports achieved fraction of the roofline  max(2*M*N*(K/n)/peak_flops, link_bytes/link_bw)  with the MEASURED
cuBLAS bf16 burst (MEASURED_PEAKS.json) and the NOMINAL NVLink 5 rate of 900 GB/s per direction per GPU (the measured
peer-copy rate on these boxes is ~770 GB/s; the nominal figure is the stricter denominator), link_bytes = S*(1+1/n)
for S = M*N*2.  The fused kernel is timed on both reduce data paths (P2P loads/stores, NVLS multimem).

    torchrun --nproc-per-node N tools/bench_rowpar.py [--out profiles/rowpar_tpN.md] [--mode p2p|nvls]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import torch.nn.functional as F

ap = argparse.ArgumentParser()
ap.add_argument("--out", default=None)
ap.add_argument("--quick", action="store_true")
args = ap.parse_args()
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
import process_manager as pm
pm.init_pgm(world)
from distributed_pytorch_from_scratch_b200.ops.gemm import gemm_nt
from distributed_pytorch_from_scratch_b200.parallel.symm import try_enable_symm
symm = try_enable_symm(required=True)
peaks = {"bf16_tflops": 1656.1}
try:
    peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
except Exception:
    pass
PEAK = peaks["bf16_tflops"] * 1e12
LINK = 900e9


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


Ms = [1024, 8192, 32768] if args.quick else [1024, 4096, 16384, 32768, 65536]
KNs = [(512, 512), (2048, 512), (4096, 1024), (6144, 2048)] if args.quick else \
      [(512, 512), (2048, 512), (1024, 1024), (4096, 1024), (2048, 2048), (6144, 2048), (8192, 8192)]
rows = []
for (K, N) in KNs:
    for M in Ms:
        if M * N * 2 > symm.gather_off or M * N * 8 > 3e9:
            continue
        Kl = K // world
        a = torch.randn(M, Kl, device="cuda").to(torch.bfloat16)
        w = torch.randn(N, Kl, device="cuda").to(torch.bfloat16)
        t_by_mode = {}
        for mode in (["p2p", "nvls"] if symm.mc_base else ["p2p"]):
            symm.mode = mode
            t_by_mode[mode] = timeit(lambda: symm.gemm_allreduce(a, w))
        best = min(t_by_mode, key=t_by_mode.get)
        symm.mode = best
        t_fused = t_by_mode[best]
        buf = symm.buffer((M, N), torch.bfloat16)

        def split():
            gemm_nt(a, w, out=buf)
            symm.all_reduce(buf, out=buf)
        t_split = timeit(split)

        def ref():
            y = F.linear(a, w)
            dist.all_reduce(y)
        t_ref = timeit(ref)
        S = M * N * 2
        roof = max(2 * M * N * Kl / PEAK, S * (1 + 1 / world) / LINK) * 1e6
        rows.append((M, K, N, t_fused, t_split, t_ref, roof, best, t_by_mode))
        if rank == 0:
            print(f"M={M:6d} K={K:5d} N={N:5d} | fused[{best}] {t_fused:8.1f} us | gemm+own AR {t_split:8.1f} us | "
                  f"cuBLAS+NCCL {t_ref:8.1f} us | roofline {roof:7.1f} us -> fused {100*roof/t_fused:5.1f}%, "
                  f"ref {100*roof/t_ref:5.1f}% | speedup vs ref {t_ref/t_fused:4.2f}x", flush=True)
if rank == 0 and args.out:
    with open(args.out, "w") as f:
        f.write(f"# RowParallelLinear forward GEMM + all-reduce, TP={world}, bf16 (device-timed, max over ranks)\n\n")
        f.write(f"roofline = max(2MN(K/n)/{PEAK/1e12:.0f} TFLOP/s (measured cuBLAS burst), S(1+1/n)/900 GB/s (nominal "
                f"NVLink 5 per direction; the measured peer-copy rate here is ~770 GB/s)), S = 2MN bytes.  The fused "
                f"column is the faster of the two reduce data paths (named in the next column).\n\n")
        f.write("| M | K | N | fused GEMM+AR (us) | path (p2p / nvls us) | our GEMM + our AR (us) | cuBLAS + NCCL (us) | roofline (us) | fused % of roofline | ref % | fused vs ref |\n|---|---|---|---|---|---|---|---|---|---|---|\n")
        for (M, K, N, tf, ts, tr, roof, best, tm) in rows:
            both = " / ".join(f"{tm[m]:.1f}" if m in tm else "-" for m in ("p2p", "nvls"))
            f.write(f"| {M} | {K} | {N} | {tf:.1f} | {best} ({both}) | {ts:.1f} | {tr:.1f} | {roof:.1f} | {100*roof/tf:.0f}% | {100*roof/tr:.0f}% | {tr/tf:.2f}x |\n")
dist.barrier()
dist.destroy_process_group()
