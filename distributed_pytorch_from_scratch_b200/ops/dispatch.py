"""Decides when the hand-written sm_100a kernels run and keeps the bf16 shadow copies of fp32 parameters.

Policy
    * CUDA tensors in bf16 compute mode (bf16 inputs, or fp32 inputs under ``torch.autocast('cuda', bf16)``, or the
      model-wide native switch) run on libdpfs_b200.so.  A missing/unloadable library on a GPU machine is an error.
    * CPU tensors and fp32-without-autocast run plain PyTorch ops: that is the oracle / parity mode
      (BASELINE.json config 1, the reference's own fp32 unit tests).
"""
from __future__ import annotations

import os

import torch

from . import _lib

_FORCE_OFF = os.environ.get("DPFS_DISABLE_NATIVE", "0") == "1"


def native_available() -> bool:
    return (not _FORCE_OFF) and torch.cuda.is_available()


def bf16_mode(x: torch.Tensor) -> bool:
    if x.dtype == torch.bfloat16:
        return True
    if x.is_cuda and torch.is_autocast_enabled("cuda") and torch.get_autocast_dtype("cuda") == torch.bfloat16:
        return True
    return False


def use_native(x: torch.Tensor) -> bool:
    """True when ``x`` should be processed by the sm_100a kernels."""
    if _FORCE_OFF or not x.is_cuda:
        return False
    if not bf16_mode(x):
        return False
    _lib.lib()  # raises if the extension cannot be loaded: never fall back silently on a GPU box
    return True


def shadow_bf16(p: torch.Tensor) -> torch.Tensor:
    """bf16 copy of an fp32 parameter, cached on the parameter and refreshed when its version counter changes.

    Parameters that live in a model arena carry a *managed* shadow (a view into the bf16 arena which the fused Adam
    kernel rewrites in the same pass as the fp32 update); for those only out-of-band edits (``copy_``,
    ``load_state_dict``) trigger a re-cast here.
    """
    if p.dtype == torch.bfloat16:
        return p.detach()
    st = getattr(p, "_dpfs_shadow", None)
    ver = p._version
    if st is not None and st[0] == ver and st[2] == p.data_ptr():
        return st[1]
    buf = st[1] if (st is not None and st[1].device == p.device and st[1].shape == p.shape) else None
    managed = getattr(p, "_dpfs_shadow_view", None)
    if managed is not None and managed.device == p.device:
        buf = managed
    if buf is None:
        buf = torch.empty(p.shape, dtype=torch.bfloat16, device=p.device)
    with torch.no_grad():
        buf.copy_(p.detach())
    p._dpfs_shadow = (ver, buf, p.data_ptr())
    return buf
