"""Spawn helpers for the multi-process tests: gloo on CPU (runs everywhere) or NCCL on GPUs.

A failing assertion in any rank propagates as an exception out of ``run_distributed`` (the reference's own tests
only *print* failures and exit 0 — defect D10)."""
import os
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _entry(rank, world, port, backend, fn, args):
    if REPO not in sys.path:
        sys.path.insert(0, REPO)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    os.environ["DPFS_NO_TQDM"] = "1"
    torch.set_num_threads(1)
    kwargs = {}
    if backend == "nccl":
        torch.cuda.set_device(rank)
        kwargs["device_id"] = torch.device("cuda", rank)
    dist.init_process_group(backend=backend, init_method="env://", world_size=world, rank=rank, **kwargs)
    import process_manager as pm
    pm.init_pgm(world)
    try:
        fn(rank, world, *args)
    finally:
        dist.barrier()
        dist.destroy_process_group()


def run_distributed(fn, world: int, port: int, backend: str = "gloo", args=()):
    mp.spawn(_entry, args=(world, port, backend, fn, args), nprocs=world, join=True)
