"""Symmetric-memory communicator for the in-kernel NVLink collectives (filled in by csrc/comm.cu).

``try_enable_symm`` attaches a ``SymmComm`` to ``pm.pgm.symm`` when every TP rank can map its peers' buffers
(torch.distributed._symmetric_memory rendezvous); otherwise the comm ops keep using NCCL.
"""
from __future__ import annotations

import warnings

from . import process_manager as pm


def try_enable_symm(required: bool = False):
    try:
        from .symm_comm import SymmComm
        pm.pgm.symm = SymmComm.create()
    except Exception as e:  # noqa: BLE001
        if required:
            raise
        if pm.pgm.tp_rank == 0:
            warnings.warn(f"symmetric-memory collectives unavailable ({type(e).__name__}: {e}); using NCCL")
        pm.pgm.symm = None
    return pm.pgm.symm
