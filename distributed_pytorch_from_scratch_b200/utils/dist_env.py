"""Seed + distributed bootstrap (reference utils.py:12-24).

``init_dist_env(args, rank)`` keeps the reference contract (``args.master_addr``, ``args.master_port``,
``args.tp_size``; env:// rendezvous; then ``pm.init_pgm``) but the backend/device are selectable
(reference defect D11 hard-codes NCCL and ``cuda:rank``): ``args.backend`` in {"nccl", "gloo"} or the
``DPFS_BACKEND`` env var; default NCCL when CUDA is available, else gloo on the CPU.
"""
from __future__ import annotations

import datetime
import os
import random
from argparse import Namespace

import numpy as np
import torch
import torch.distributed as dist

from ..parallel import process_manager as pm


def set_seed(seed):
    random.seed(seed)
    np.random.seed(seed)
    torch.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed)


def pick_backend(args=None) -> str:
    b = getattr(args, "backend", None) if args is not None else None
    b = b or os.environ.get("DPFS_BACKEND")
    if b:
        return b
    return "nccl" if torch.cuda.is_available() else "gloo"


def init_dist_env(args: Namespace, rank: int):
    addr = args.master_addr
    if addr == "localhost":
        addr = "127.0.0.1"   # the container hostname / localhost may not resolve on the GPU boxes
    os.environ["MASTER_ADDR"] = addr
    os.environ["MASTER_PORT"] = str(args.master_port)
    backend = pick_backend(args)
    if backend == "nccl":
        local = int(os.environ.get("LOCAL_RANK", rank))
        torch.cuda.set_device(local % max(torch.cuda.device_count(), 1))
    if not dist.is_initialized():
        kwargs = {}
        if backend == "nccl":
            kwargs["device_id"] = torch.device("cuda", torch.cuda.current_device())
        dist.init_process_group(backend=backend, init_method="env://", world_size=args.tp_size, rank=rank,
                                timeout=datetime.timedelta(seconds=int(os.environ.get("DPFS_DIST_TIMEOUT", "600"))),
                                **kwargs)
    pm.init_pgm(args.tp_size)
    comm = getattr(args, "comm", None) or os.environ.get("DPFS_COMM", "auto")
    if backend == "nccl" and args.tp_size > 1 and comm in ("fused", "auto"):
        from ..parallel.symm import try_enable_symm
        try_enable_symm(required=(comm == "fused"))
    return pm.pgm
