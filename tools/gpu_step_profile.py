"""Per-kernel time breakdown of one training step of the default model at TP = WORLD_SIZE (1 when run without torchrun).

    python tools/gpu_step_profile.py [--sp on|off|auto] [--eager] [--tag NAME]
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/gpu_step_profile.py --sp on

Uses torch.profiler (CUPTI activity records, a few percent overhead) only to apportion time between kernels; the
step time itself is always taken from bench.py, never from a profiled run.  By default the profiled step is a CUDA
graph replay (what bench.py times): in-kernel waits for peers are then the real skew, not Python launch jitter.
Rank 0 writes gpurun_out/step_profile_tp{N}{tag}.md.
"""
import argparse
import os
import re
import sys
from collections import defaultdict
from dataclasses import asdict

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)

import torch  # noqa: E402
from torch.optim.lr_scheduler import OneCycleLR  # noqa: E402

import process_manager as pm  # noqa: E402
from constants import IGNORE_INDEX, ModelArgumments  # noqa: E402
from distributed_pytorch_from_scratch_b200.engine import TrainStep  # noqa: E402
from distributed_pytorch_from_scratch_b200.ops.optim import FusedAdam  # noqa: E402
from models.model import Transformer  # noqa: E402
from utils import set_seed  # noqa: E402


def short(name: str) -> str:
    name = re.sub(r"^void\s+", "", name)
    name = name.replace("dpfs::", "").replace("(anonymous namespace)::", "")
    name = re.sub(r"\(.*$", "", name)
    name = name.replace("__nv_bfloat16", "bf16").replace("(bool)1", "T").replace("(bool)0", "F").replace("(int)", "")
    return name[:110]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sp", default="auto")
    ap.add_argument("--eager", action="store_true")
    ap.add_argument("--tag", default="")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29617")
    torch.distributed.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    set_seed(0)
    pm.init_pgm(world)
    if world > 1:
        from distributed_pytorch_from_scratch_b200.parallel.symm import try_enable_symm
        try_enable_symm(required=True)
    sp = world > 1 and (args.sp == "on" or (args.sp == "auto" and world >= 8))
    margs = ModelArgumments()
    model = Transformer(**asdict(margs), sequence_parallel=sp).cuda()
    model.reset_parameters()
    model.compute_dtype = torch.bfloat16
    model.train()
    opt = FusedAdam.for_model(model, lr=3e-4)
    sched = OneCycleLR(opt, 3e-4, 10000, pct_start=0.04)
    opt._opt_called = True
    step = TrainStep(model, opt, sched, ignore_index=IGNORE_INDEX, use_graph=not args.eager)
    B, T = 32, 1000
    g = torch.Generator().manual_seed(1234)
    ids = torch.randint(3, margs.vocab_size, (B, T), generator=g)
    pos = torch.arange(T).repeat(B, 1)
    tgt = torch.roll(ids, -1, dims=1)
    shape, bufs = step.load_batch(ids, pos, tgt)
    for _ in range(4):
        step.run_loaded(shape, bufs)
    torch.cuda.synchronize()
    torch.distributed.barrier()
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    nprof = 3
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(nprof):
            step.run_loaded(shape, bufs)
        torch.cuda.synchronize()
    tot = defaultdict(lambda: [0.0, 0])
    t_first, t_last = None, None
    for ev in prof.events():
        if ev.device_type == torch.autograd.DeviceType.CUDA:
            d = getattr(ev, "device_time", None)
            if d is None:
                d = ev.cuda_time
            tot[short(ev.name)][0] += d / nprof
            tot[short(ev.name)][1] += 1
            tr = ev.time_range
            t_first = tr.start if t_first is None else min(t_first, tr.start)
            t_last = tr.end if t_last is None else max(t_last, tr.end)
    # per-rank view of the last profiled step (all ranks): compute time outside the hand-off kernels shows a slow GPU,
    # per-instance hand-off durations show who waits for whom (the rank that arrives last has the shortest kernel)
    evs = sorted((ev for ev in prof.events() if ev.device_type == torch.autograd.DeviceType.CUDA),
                 key=lambda e: e.time_range.start)
    per_step = len(evs) // nprof
    last = evs[-per_step:]
    def dur(e):
        d = getattr(e, "device_time", None)
        return e.cuda_time if d is None else d
    hand = [(short(e.name)[:24], dur(e)) for e in last if "_ag_kernel" in e.name or "handoff_kernel" in e.name]
    other = sum(dur(e) for e in last if "_ag_kernel" not in e.name and "handoff_kernel" not in e.name)
    gathered = [None] * world
    torch.distributed.all_gather_object(gathered, {"rank": rank, "other_us": other, "hand": hand})
    total = sum(v[0] for v in tot.values())
    span = (t_last - t_first) / nprof if t_first is not None else 0.0
    mode = "eager" if args.eager else "CUDA-graph replay"
    lines = [f"# Kernel-time breakdown of one training step (TP={world}{'+SP' if sp else ''}, rank 0, {mode}, default "
             f"model, b32 x t1000, bf16)", "",
             f"CUPTI activity records via torch.profiler, mean of {nprof} steps; sum of kernel time "
             f"{total / 1e3:.2f} ms, first-to-last activity span {span / 1e3:.2f} ms per step, "
             f"{sum(v[1] for v in tot.values()) // nprof} device activities per step.  Environment switches: "
             + (", ".join(f"{k}={v}" for k, v in sorted(os.environ.items()) if k.startswith("DPFS_")) or "none") + ".",
             "", "| time (ms) | share | launches | us / launch | kernel |", "|---|---|---|---|---|"]
    for name, (t, n) in sorted(tot.items(), key=lambda kv: -kv[1][0]):
        n1 = n / nprof
        lines.append(f"| {t / 1e3:.3f} | {100 * t / total:.1f}% | {n1:g} | {t / n1:.1f} | {name} |")
    if world > 1 and gathered[0]["hand"]:
        lines += ["", "## Per-rank view of one step", "",
                  "| rank | kernel time outside the hand-off kernels (ms) | hand-off kernels (ms) |", "|---|---|---|"]
        for g in gathered:
            lines.append(f"| {g['rank']} | {g['other_us'] / 1e3:.3f} | {sum(d for _, d in g['hand']) / 1e3:.3f} |")
        n_inst = min(len(g["hand"]) for g in gathered)
        lines += ["", "Duration (us) of every hand-off kernel of the step, in launch order, on each rank (the rank with the "
                  "shortest one arrived last at the opening barrier):", "",
                  "| # | kernel | " + " | ".join(f"r{g['rank']}" for g in gathered) + " |",
                  "|---|---|" + "---|" * world]
        for i in range(n_inst):
            lines.append(f"| {i} | {gathered[0]['hand'][i][0]} | "
                         + " | ".join(f"{g['hand'][i][1]:.0f}" for g in gathered) + " |")
    out = "\n".join(lines) + "\n"
    if rank == 0:
        os.makedirs(os.path.join(REPO, "gpurun_out"), exist_ok=True)
        with open(os.path.join(REPO, "gpurun_out", f"step_profile_tp{world}{args.tag}.md"), "w") as f:
            f.write(out)
        print(out)
    torch.distributed.barrier()
    torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
