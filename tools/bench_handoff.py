"""Micro-benchmark of the sequence-parallel hand-off kernels and of their building blocks at TP = WORLD_SIZE.

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/bench_handoff.py

Device-timed (CUDA events), max over ranks.  Rank 0 writes gpurun_out/handoff_tp{N}.md:
  * the rank barrier alone (per-block flavour at several grid sizes)
  * multimem.ld_reduce only / multimem.st only / both at once on two streams (does the pull overlap the push?)
  * rs_add_rmsnorm_ag / rs_rmsnorm_bwd_ag for barrier flavour x grid cap, landing in one zone or rotating over 25
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
import process_manager as pm
pm.init_pgm(world)
from distributed_pytorch_from_scratch_b200.ops import _lib, fused
from distributed_pytorch_from_scratch_b200.parallel.symm import try_enable_symm

symm = try_enable_symm(required=True)
lib = _lib.lib()
M, D = 32000, 512
Ml = M // world
lines = []


def timeit(fn, iters=20, warm=3):
    for _ in range(warm):
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


def out(s):
    lines.append(s)
    if rank == 0:
        print(s, flush=True)


out(f"# Hand-off micro-benchmarks, TP={world}, [{M}, {D}] bf16 ({M * D * 2 / 1e6:.1f} MB), us per call (max over ranks)")
out("")
# ---- barrier alone
out("| rank barrier (per-block flavour), blocks x threads | us / barrier |")
out("|---|---|")
for blocks, threads in [(1, 32), (64, 512), (125, 512), (148, 512), (296, 256)]:
    reps = 10
    t = timeit(lambda: _lib.check(lib.dpfs_comm_barrier_ex(symm.flag_ptrs, symm.rank, world, blocks, threads, reps,
                                                           _lib.stream()), "barrier_ex"))
    t1 = timeit(lambda: _lib.check(lib.dpfs_comm_barrier_ex(symm.flag_ptrs, symm.rank, world, blocks, threads, 1,
                                                            _lib.stream()), "barrier_ex"))
    out(f"| {blocks} x {threads} | {(t - t1) / (reps - 1):.2f} (kernel with one barrier: {t1:.1f}) |")
out("")
# ---- pull / push alone and together
part = symm.buffer((M, D), torch.bfloat16)
part.normal_()
nbytes = M * D * 2
s2 = torch.cuda.Stream()


def exp(mode, blocks, stream=None):
    st = _lib.stream() if stream is None else stream.cuda_stream
    _lib.check(lib.dpfs_allreduce_exp(symm.mc_base, symm.data_ptrs, 0, nbytes, mode, blocks, symm.flag_ptrs, symm.rank,
                                      world, st), "exp")


out("| data phase only (no barriers), 64 / 128 blocks x 512 threads | us |")
out("|---|---|")
for blocks in (64, 128):
    t_ld = timeit(lambda: exp(1 | 16, blocks))
    t_st = timeit(lambda: exp(2 | 16, blocks))

    def both():
        s2.wait_stream(torch.cuda.current_stream())
        exp(1 | 16, blocks)
        exp(2 | 16, blocks, s2)
        torch.cuda.current_stream().wait_stream(s2)
    t_both = timeit(both)
    out(f"| {blocks} blocks: multimem.ld_reduce of S/n per rank | {t_ld:.1f} |")
    out(f"| {blocks} blocks: multimem.st of S/n per rank | {t_st:.1f} |")
    out(f"| {blocks} blocks: both, on two streams | {t_both:.1f} |")
out("")
# ---- the hand-off kernels
zones = symm.alloc_zones(26, M * D * 2)
x_l = torch.randn(Ml, D, device="cuda")
dres = torch.randn(Ml, D, device="cuda")
bias, scale = torch.randn(D, device="cuda"), torch.rand(D, device="cuda") + 0.5
rstd = torch.rand(Ml, device="cuda") + 0.5
ds, db = torch.empty(D, device="cuda"), torch.empty(D, device="cuda")
out("| kernel | barrier | grid cap | landing zone | us |")
out("|---|---|---|---|---|")
state = {"i": 0}
for bar, bname in [(1, "grid-wide, 1 fence / block"), (3, "grid-wide, fence / thread"), (0, "per block"), (2, "none")]:
    for cap in (0, 74, 37, 24, 16):
        lib.dpfs_sp_debug_config(cap, bar)
        for rot in (False, True):
            def fwd():
                z = zones[state["i"] % 25] if rot else zones[0]
                state["i"] += 1
                fused.rs_add_rmsnorm_ag(symm, part, x_l, bias, scale, 1e-5, z)
            t = timeit(fwd)
            out(f"| rs_add_rmsnorm_ag | {bname} | {cap or 'auto'} | {'rotating over 25' if rot else 'one'} | {t:.1f} |")
        t = timeit(lambda: fused.rs_rmsnorm_bwd_ag(symm, part, x_l, rstd, scale, dres, zones[25], ds, db))
        out(f"| rs_rmsnorm_bwd_ag (+ partial-sum reduce) | {bname} | {cap or 'auto'} | one | {t:.1f} |")
lib.dpfs_sp_debug_config(0, 1)
out("")
# ---- round-1 split kernels for reference
loc = torch.randn(Ml, D, device="cuda").to(torch.bfloat16)
symm.mode = "nvls"
t_rs = timeit(lambda: symm.reduce_scatter_rows(part))
t_ag = timeit(lambda: symm.all_gather_into_buffer(loc))
t_ar = timeit(lambda: symm.all_reduce(part, out=part))
out(f"stand-alone NVLS kernels on the same box: reduce-scatter {t_rs:.1f} us, all-gather {t_ag:.1f} us, "
    f"all-reduce {t_ar:.1f} us")
if rank == 0:
    os.makedirs("gpurun_out", exist_ok=True)
    open(f"gpurun_out/handoff_tp{world}.md", "w").write("\n".join(lines) + "\n")
dist.barrier(); dist.destroy_process_group()
