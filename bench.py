"""Headline benchmark: training tokens/s of the reference's default train.py transformer (ModelArgumments:
d512 / ffn2048 / 8 heads / 12 layers / vocab 1024 / maxlen 1000), global batch 32 x 1000 tokens, bf16, at TP = N.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

Both arms run the same step (forward, loss, backward, Adam + OneCycleLR update) on the same synthetic batch of the
named shape with random-init weights; ``--impl reference`` runs the UNMODIFIED reference from baseline/_ref through
its own public API (its Transformer / process_manager / autocast + F.cross_entropy + torch Adam, i.e. train.py's
loop), none of this repo's code on that path.  Timing: CUDA events around exactly K steps, barrier +
synchronize on both sides, max over ranks.  Rank 0 prints ONE JSON line.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from dataclasses import asdict

REPO = os.path.dirname(os.path.abspath(__file__))
BATCH, SEQ = 32, 1000          # reference train.py default batch size, ModelArgumments.maxlen
LR, WARMUP_STEPS, MAX_STEPS = 3e-4, 2000, 20000   # reference train.py defaults (OneCycleLR)
METRIC = "train tokens/sec at TP=N (max over ranks), default train.py transformer"

# BASELINE.json configs 3 and 4 (same step, other shapes); the default (config 2) is the headline the driver runs
CONFIGS = {
    "default": None,
    "gpt2m": dict(name="GPT-2-medium shape (d1024 L24 H16 ffn4096 seq1024 V50304)", batch=8, seq=1024,
                  model=dict(attn_dim=1024, ffn_dim=4096, num_heads=16, num_layers=24, vocab_size=50304, maxlen=1024)),
    "1p3b": dict(name="1.3B (d2048 L24 H16 ffn6144 seq2048 V50304)", batch=4, seq=2048,
                 model=dict(attn_dim=2048, ffn_dim=6144, num_heads=16, num_layers=24, vocab_size=50304, maxlen=2048)),
}


def apply_config(args):
    """Returns (model kwargs, description) and sets the global batch / sequence length of the run."""
    global BATCH, SEQ, METRIC
    from dataclasses import asdict as _asdict
    cfg = CONFIGS[args.config]
    if cfg is None:
        return None, "reference default ModelArgumments (d512 ffn2048 h8 L12 V1024 maxlen1000, 51.5M params)"
    BATCH, SEQ = cfg["batch"], cfg["seq"]
    METRIC = f"train tokens/sec at TP=N (max over ranks), {cfg['name']}"
    return dict(cfg["model"], rope_theta=10000.), cfg["name"]


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """Samples SM clocks / throttle reasons with nvidia-smi while the timed region runs."""

    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int = 0):
        self.rows = []
        self.proc = None
        self.gpu = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0: float, t1: float):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, smmax, reasons = [], [], set()
        for ts, line in self.rows:
            if ts < t0 or ts > t1 + 0.2:
                continue
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1]))
                smmax.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"),
                                 parts[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:  # region shorter than the sampling period: use every sample we have
            for ts, line in self.rows:
                parts = [p.strip() for p in line.split(",")]
                try:
                    sm.append(float(parts[1]))
                    smmax.append(float(parts[2]))
                except (ValueError, IndexError):
                    pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(smmax) if smmax else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------ common
def dist_setup(n_gpus: int):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert world == n_gpus, f"--gpus {n_gpus} but WORLD_SIZE={world}: launch with torch.distributed.run for N > 1"
    torch.cuda.set_device(local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29533")
    dist.init_process_group(backend="nccl", init_method="env://", world_size=world, rank=rank,
                            device_id=torch.device("cuda", local))
    return rank, world, local


def make_batch(vocab: int, seed: int = 1234):
    """Synthetic batch with the layout dataset.collate_fn produces: BOS-prefixed inputs, shifted targets + EOS."""
    import torch
    g = torch.Generator().manual_seed(seed)
    toks = torch.randint(3, vocab, (BATCH, SEQ - 1), generator=g)
    input_ids = torch.cat([torch.zeros(BATCH, 1, dtype=torch.long), toks], dim=1)            # BOS = 0
    target_ids = torch.cat([toks, torch.ones(BATCH, 1, dtype=torch.long)], dim=1)            # EOS = 1
    position_ids = torch.arange(SEQ).unsqueeze(0).repeat(BATCH, 1)
    return input_ids.pin_memory(), position_ids.pin_memory(), target_ids.pin_memory()


def timed(fn_step, steps: int, warmup: int, world: int, sampler):
    """W untimed steps, then exactly K steps between CUDA events with barrier + synchronize on both sides."""
    import torch
    import torch.distributed as dist
    for _ in range(warmup):
        fn_step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time()
    e0.record()
    for _ in range(steps):
        fn_step()
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t1 = time.time()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = t.item()
    return ms / steps, t0, t1


def dbg(msg):
    if os.environ.get("DPFS_DEBUG"):
        print(f"[bench rank {os.environ.get('RANK', '0')}] {msg}", file=sys.stderr, flush=True)


def emit(rank, payload):
    if rank == 0:
        print(json.dumps(payload), flush=True)


# ------------------------------------------------------------------------------------------------ our arm
def run_ours(args):
    sys.path.insert(0, REPO)
    import torch
    import torch.distributed as dist
    from torch.optim.lr_scheduler import OneCycleLR

    import process_manager as pm
    from constants import IGNORE_INDEX, ModelArgumments
    from distributed_pytorch_from_scratch_b200.engine import TrainStep
    from distributed_pytorch_from_scratch_b200.ops import _lib
    from distributed_pytorch_from_scratch_b200.ops.optim import FusedAdam
    from models.model import Transformer
    from utils import set_seed

    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1 or args.gpus > 1:
        rank, world, local = dist_setup(args.gpus)
    else:
        rank, world, local = 0, 1, 0
        torch.cuda.set_device(0)
    set_seed(0)
    pm.init_pgm(world)
    if world > 1 and args.comm != "nccl":
        from distributed_pytorch_from_scratch_b200.parallel.symm import try_enable_symm
        try_enable_symm(required=(args.comm == "fused"))
    mk, desc = apply_config(args)
    margs = ModelArgumments() if mk is None else ModelArgumments(**mk)
    # Sequence parallelism with the one-kernel hand-offs (csrc/sp_fused.cu) beats the all-reduce form at every TP size
    # measured: 15.7 vs 16.3 ms at TP=2, 10.9 vs 13.7 ms at TP=8 (profiles/), so "auto" turns it on for every N > 1.
    sp = world > 1 and (args.sp == "on" or (args.sp == "auto" and pm.pgm.symm is not None))
    model = Transformer(**asdict(margs), sequence_parallel=sp).cuda()
    model.reset_parameters()
    model.compute_dtype = torch.bfloat16
    model.train()
    opt = FusedAdam.for_model(model, lr=LR)
    sched = OneCycleLR(opt, LR, MAX_STEPS, pct_start=WARMUP_STEPS / MAX_STEPS)
    opt._opt_called = True
    step = TrainStep(model, opt, sched, ignore_index=IGNORE_INDEX, use_graph=not args.no_graph)
    ids_h, pos_h, tgt_h = make_batch(margs.vocab_size)
    dbg("model built")

    # ---- device-timed steps (inputs already resident in the static buffers)
    shape, bufs = step.load_batch(ids_h, pos_h, tgt_h)
    torch.cuda.synchronize()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
        time.sleep(0.3)
    last = {}

    def dev_step():
        last["loss"] = step.run_loaded(shape, bufs)
        dbg("step done")

    ms_step, t0, t1 = timed(dev_step, args.steps, args.warmup, world, sampler)
    clocks = sampler.stop(t0, t1) if sampler else None
    loss_dev = float(last["loss"].item())

    # ---- end-to-end steps through the public API: H2D of the batch from pinned memory + D2H of the loss, per step
    def e2e_step():
        loss = step(ids_h, pos_h, tgt_h)
        last["loss_host"] = loss.item()

    ms_e2e, _, _ = timed(e2e_step, args.steps, 1, world, None)
    # ---- parity of the in-kernel NVLink collectives against NCCL on this very box (after the timed regions):
    # loss and this rank's gradient arena of one forward/backward, fused path vs the same model with NCCL collectives
    parity = None
    if world > 1 and pm.pgm.symm is not None and not args.no_check:
        try:
            ids_d, pos_d, tgt_d = ids_h.cuda(), pos_h.cuda(), tgt_h.cuda()
            model.zero_grad(set_to_none=False)
            loss_f = model.loss(ids_d, pos_d, tgt_d, IGNORE_INDEX)
            loss_f.backward()
            g_f = opt.arena.grads.clone()
            symm_saved, pm.pgm.symm = pm.pgm.symm, None          # same kernels, collectives through NCCL
            loss_n = model.loss(ids_d, pos_d, tgt_d, IGNORE_INDEX)
            loss_n.backward()
            g_n = opt.arena.grads.clone()
            pm.pgm.symm = symm_saved
            t = torch.stack([(g_f - g_n).norm() / g_n.norm().clamp_min(1e-30), (loss_f - loss_n).abs().float()])
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            parity = {"loss_abs_diff": float(t[1]), "gradnorm_rel_diff": float(t[0]),
                      "what": "fused NVLink collectives vs NCCL collectives, same weights and batch, max over ranks"}
            del g_f, g_n
        except Exception as e:  # noqa: BLE001  (the check must never take the benchmark down)
            parity = {"error": f"{type(e).__name__}: {e}"}
            pm.pgm.symm = pm.pgm.symm or locals().get("symm_saved")

    tokens = BATCH * SEQ
    payload = {
        "metric": METRIC, "value": tokens / ms_step * 1e3, "unit": "tokens/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic tokens (uniform ids, collate_fn layout), random-init weights",
        "impl": "ours",
        "config": {"model": desc,
                   "global_batch": BATCH, "seq_len": SEQ, "parallelism": f"tp{world}" + ("+sp" if sp else ""),
                   "optimizer": "Adam(3e-4) + OneCycleLR (fused, one launch)",
                   "cuda_graph": bool(step.use_graph),
                   "comm": ("fused-symm" if (pm.pgm.symm is not None) else ("nccl" if world > 1 else "none")),
                   "l2": "per-step working set (~9 GB of activations) >> 126 MB L2; no explicit flush"},
        "clocks": clocks,
        "e2e": {"value": tokens / ms_e2e * 1e3, "unit": "tokens/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": int(3 * BATCH * SEQ * 8), "d2h_bytes_per_step": 4},
        "gpu_launches": int(step.kernels_per_step) * args.steps,
        "kernels_per_step": int(step.kernels_per_step),
        "final_loss": loss_dev,
        "parity": parity,
        "lib": os.path.relpath(_lib.LIB_PATH, REPO),
    }
    emit(rank, payload)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------ reference arm
def run_reference(args):
    ref = os.path.join(REPO, "baseline", "_ref")
    if not os.path.isdir(os.path.join(ref, "models")):
        print(json.dumps({"impl": "reference", "unavailable": "baseline/_ref is missing (reference not installed)"}))
        return
    # the reference's flat module names (models, process_manager, utils, constants) shadow this repo's shims:
    # make sure ONLY baseline/_ref is importable under those names
    sys.path = [ref] + [p for p in sys.path if os.path.abspath(p or ".") != REPO]
    import torch
    import torch.distributed as dist
    import torch.nn.functional as F
    import torch.optim as optim

    try:
        import process_manager as pm
        from constants import IGNORE_INDEX, ModelArgumments
        from models.model import Transformer
        assert os.path.abspath(pm.__file__).startswith(ref), pm.__file__
    except Exception as e:  # pragma: no cover
        print(json.dumps({"impl": "reference", "unavailable": f"import failed: {type(e).__name__}: {e}"}))
        return

    rank, world, local = dist_setup(args.gpus)     # the reference requires an initialised process group
    torch.manual_seed(0)
    torch.cuda.manual_seed_all(0)
    os.environ["DTYPE"] = "bfloat16"               # train.py --bf16 (reference train.py:58-63)
    pm.init_pgm(world)
    mk, desc = apply_config(args)
    margs = ModelArgumments() if mk is None else ModelArgumments(**mk)
    model = Transformer(**asdict(margs))
    model.cuda()
    model.reset_parameters()
    model.train()
    optimizer = optim.Adam(model.parameters(), lr=LR)
    scheduler = optim.lr_scheduler.OneCycleLR(optimizer, LR, MAX_STEPS, pct_start=WARMUP_STEPS / MAX_STEPS)
    ids_h, pos_h, tgt_h = make_batch(margs.vocab_size)
    ids_d, pos_d, tgt_d = ids_h.cuda(), pos_h.cuda(), tgt_h.cuda()
    last = {}

    def train_step(input_ids, position_ids, target_ids):
        # reference train.py:99-109
        with torch.autocast("cuda", enabled=True, dtype=torch.bfloat16):
            logits = model(input_ids, position_ids)
            loss = F.cross_entropy(logits.float().view(-1, logits.size(-1)), target_ids.view(-1),
                                   ignore_index=IGNORE_INDEX, reduction="mean")
        optimizer.zero_grad()
        loss.backward()
        optimizer.step()
        scheduler.step()
        return loss

    def dev_step():
        # its embedding rewrites input_ids in place at TP > 1, so every step gets a fresh copy like a new batch
        last["loss"] = train_step(ids_d.clone(), pos_d, tgt_d)

    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
        time.sleep(0.3)
    ms_step, t0, t1 = timed(dev_step, args.steps, args.warmup, world, sampler)
    clocks = sampler.stop(t0, t1) if sampler else None
    loss_dev = float(last["loss"].item())

    def e2e_step():
        # reference train.py:96-98 + :110 — blocking .cuda() copies of the batch, loss.item() every step
        loss = train_step(ids_h.cuda(), pos_h.cuda(), tgt_h.cuda())
        last["loss_host"] = loss.item()

    ms_e2e, _, _ = timed(e2e_step, args.steps, 1, world, None)
    tokens = BATCH * SEQ
    payload = {
        "metric": METRIC, "value": tokens / ms_step * 1e3, "unit": "tokens/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic tokens (uniform ids, collate_fn layout), random-init weights",
        "impl": "reference",
        "config": {"model": desc,
                   "global_batch": BATCH, "seq_len": SEQ, "parallelism": f"tp{world}",
                   "optimizer": "torch.optim.Adam(3e-4) + OneCycleLR", "cuda_graph": False,
                   "comm": "nccl" if world > 1 else "none",
                   "l2": "per-step working set >> 126 MB L2; no explicit flush"},
        "clocks": clocks,
        "e2e": {"value": tokens / ms_e2e * 1e3, "unit": "tokens/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": int(3 * BATCH * SEQ * 8), "d2h_bytes_per_step": 4},
        "gpu_launches": 0,
        "final_loss": loss_dev,
        "max_mem_gib": torch.cuda.max_memory_allocated() / 2 ** 30,
    }
    emit(rank, payload)
    dist.barrier()
    dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=int(os.environ.get("WORLD_SIZE", "1")))
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--config", choices=list(CONFIGS), default="default",
                    help="default = the reference's train.py model (the headline); gpt2m / 1p3b = BASELINE.json configs 3 / 4")
    ap.add_argument("--comm", choices=["auto", "fused", "nccl"], default="auto",
                    help="TP collectives: our NVLink kernels over symmetric memory (fused), NCCL, or auto")
    ap.add_argument("--no-check", action="store_true",
                    help="skip the one-step fused-vs-NCCL parity check that runs before the timed region at N > 1")
    ap.add_argument("--no-graph", action="store_true", help="run eagerly instead of replaying a CUDA graph")
    ap.add_argument("--sp", choices=["on", "off", "auto"], default="auto",
                    help="Megatron sequence parallelism on top of TP (reduce-scatter / all-gather forms)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
