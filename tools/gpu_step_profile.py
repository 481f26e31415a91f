"""Per-kernel time breakdown of one eager (no CUDA graph) training step of the default model at TP=1.

Uses torch.profiler (CUPTI activity records, a few percent overhead) only to apportion time between kernels;
the step time itself is always taken from bench.py, never from a profiled run.
Writes gpurun_out/step_profile.md.
"""
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
    name = name.replace("__nv_bfloat16", "bf16").replace("(bool)1", "MN").replace("(bool)0", "K").replace("(int)", "")
    return name[:90]


def main():
    torch.cuda.set_device(0)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29617")
    torch.distributed.init_process_group("nccl", rank=0, world_size=1)
    set_seed(0)
    pm.init_pgm(1)
    margs = ModelArgumments()
    model = Transformer(**asdict(margs)).cuda()
    model.reset_parameters()
    model.compute_dtype = torch.bfloat16
    model.train()
    opt = FusedAdam.for_model(model, lr=3e-4)
    sched = OneCycleLR(opt, 3e-4, 10000, pct_start=0.04)
    opt._opt_called = True
    step = TrainStep(model, opt, sched, ignore_index=IGNORE_INDEX, use_graph=False)
    B, T = 32, 1000
    g = torch.Generator().manual_seed(1234)
    ids = torch.randint(3, margs.vocab_size, (B, T), generator=g)
    pos = torch.arange(T).repeat(B, 1)
    tgt = torch.roll(ids, -1, dims=1)
    shape, bufs = step.load_batch(ids, pos, tgt)
    for _ in range(3):
        step.run_loaded(shape, bufs)
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        step.run_loaded(shape, bufs)
        torch.cuda.synchronize()
    tot = defaultdict(lambda: [0.0, 0])
    for ev in prof.events():
        if ev.device_type == torch.autograd.DeviceType.CUDA:
            d = getattr(ev, "device_time", None)
            if d is None:
                d = ev.cuda_time
            tot[short(ev.name)][0] += d
            tot[short(ev.name)][1] += 1
    total = sum(v[0] for v in tot.values())
    lines = ["# Kernel-time breakdown of one eager training step (TP=1, default model, b32 x t1000, bf16)", "",
             f"CUPTI activity records via torch.profiler; sum of kernel time {total / 1e3:.2f} ms over "
             f"{sum(v[1] for v in tot.values())} device activities.", "",
             "| time (ms) | share | launches | us / launch | kernel |", "|---|---|---|---|---|"]
    for name, (t, n) in sorted(tot.items(), key=lambda kv: -kv[1][0]):
        lines.append(f"| {t / 1e3:.3f} | {100 * t / total:.1f}% | {n} | {t / n:.1f} | {name} |")
    os.makedirs(os.path.join(REPO, "gpurun_out"), exist_ok=True)
    out = "\n".join(lines) + "\n"
    with open(os.path.join(REPO, "gpurun_out", "step_profile.md"), "w") as f:
        f.write(out)
    print(out)
    torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
