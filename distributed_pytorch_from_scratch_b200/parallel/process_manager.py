"""Process-group manager: the single source of truth for the tensor-parallel topology.

Same surface as the reference (process_manager.py:8-25): a module-global ``pgm`` set by ``init_pgm(tp_size)`` with
``global_rank, world_size, tp_size, tp_rank, grid, tp_group`` and ``repr() == "TP{n}"``.  Only tensor parallelism
exists (``tp_size == world_size``); ``grid`` keeps a leading axis so a data-parallel dimension can be added without
touching the users of ``tp_group``.

Additions for the B200 runtime: the manager remembers the device/backed in use and owns the symmetric-memory
communicator (``pgm.symm``) that the fused GEMM+collective kernels use over NVLink.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist

pgm: Optional["ProcessGroupManager"] = None


class ProcessGroupManager(object):
    def __init__(self, tp_size: int):
        if dist.is_available() and dist.is_initialized():
            self.global_rank = dist.get_rank()
            self.world_size = dist.get_world_size()
            self.backend = dist.get_backend()
        else:
            # single-process mode (no rendezvous needed for TP=1)
            assert tp_size == 1, "torch.distributed must be initialised before init_pgm() when tp_size > 1"
            self.global_rank, self.world_size, self.backend = 0, 1, "none"
        assert tp_size == self.world_size, (
            f"only tensor parallelism is implemented: tp_size ({tp_size}) must equal world_size ({self.world_size})")
        self.tp_size = tp_size
        self.tp_rank = self.global_rank  # no other parallel axis, so tp_rank == global_rank
        self.grid = torch.arange(self.world_size).view(self.tp_size)
        if self.backend != "none":
            self.tp_group = dist.new_group(self.grid.tolist())
        else:
            self.tp_group = None
        self.symm = None  # parallel.symm.SymmComm once enabled (CUDA + NVLink peers)

    @property
    def device(self) -> torch.device:
        if self.backend == "nccl" or (self.backend == "none" and torch.cuda.is_available()):
            return torch.device("cuda", torch.cuda.current_device())
        return torch.device("cpu")

    def __repr__(self):
        return f"TP{self.tp_size}"


def init_pgm(tp_size: int):
    global pgm
    pgm = ProcessGroupManager(tp_size)
    return pgm


def destroy_pgm():
    global pgm
    if pgm is not None and pgm.symm is not None:
        pgm.symm.close()
    pgm = None
