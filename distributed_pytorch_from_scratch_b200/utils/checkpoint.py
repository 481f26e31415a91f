"""Per-rank sharded checkpoints in the reference's exact format (train.py:121-133, test.py:49-55,94-95):

    {save_dir}/tprank-{rank}_iter-{step}_loss-{avg_loss:.4f}.pth  = torch.save(model.state_dict())

i.e. a plain ``OrderedDict[str, Tensor]`` of this rank's fp32 shards (196 keys for the default model), readable by the
reference's ``load_ckpt`` and vice versa.  Parameters living in the flat arena are cloned before saving so the file
holds independent tensors, not views of one big storage.

Extras kept in SEPARATE files so the base format stays untouched: ``resume-tprank-{r}.pt`` with
optimizer / scheduler / step / RNG state for true resume, and ``reshard_checkpoint`` to move a checkpoint between TP
sizes.
"""
from __future__ import annotations

import glob
import os
import re
from collections import OrderedDict
from typing import Dict, List, Optional

import torch

CKPT_RE = re.compile(r"tprank-(\d+)_iter-(\d+)_loss-(.+?)\.pth$")


def ckpt_name(rank: int, step: int, avg_loss: float) -> str:
    return f"tprank-{rank}_iter-{step}_loss-{avg_loss:.4f}.pth"


def list_checkpoints(ckpt_dir: str, rank: int) -> List[str]:
    """This rank's checkpoints sorted by iteration (reference test.py:94-95)."""
    paths = glob.glob(os.path.join(ckpt_dir, f"tprank-{rank}_iter-*_loss-*.pth"))
    return sorted(paths, key=lambda p: int(CKPT_RE.search(os.path.basename(p)).group(2)))


def save_checkpoint(model, save_dir: str, rank: int, step: int, avg_loss: float, keep_last_n: int = -1,
                    verbose: bool = True) -> str:
    os.makedirs(save_dir, exist_ok=True)
    path = os.path.join(save_dir, ckpt_name(rank, step, avg_loss))
    sd = OrderedDict((k, v.detach().clone().contiguous()) for k, v in model.state_dict().items())
    torch.save(sd, path)
    if verbose:
        print(f"[TP rank {rank}]: Model saved to {path}")
    if keep_last_n > 0:
        for old in list_checkpoints(save_dir, rank)[:-keep_last_n]:
            if verbose:
                print(f"[TP rank {rank}]: Remove old checkpoint {old}")
            os.remove(old)
    return path


def load_checkpoint(model, path: str, dtype: Optional[torch.dtype] = None, map_location=None, strict: bool = True):
    """Reference ``load_ckpt`` semantics (test.py:49-55): optional cast of checkpoint + model, strict load."""
    if map_location is None:
        map_location = next(model.parameters()).device
    ckpt = torch.load(path, map_location=map_location, weights_only=True)
    if dtype is not None:
        ckpt = {k: v.to(dtype) for k, v in ckpt.items()}
        model.to(dtype)
    model.load_state_dict(ckpt, strict=strict)
    return model


def save_resume_state(save_dir: str, rank: int, step: int, optimizer, scheduler, accum_loss: float, extra=None):
    os.makedirs(save_dir, exist_ok=True)
    st = {"step": step, "accum_loss": accum_loss,
          "optimizer": optimizer.fused_state() if hasattr(optimizer, "fused_state") else optimizer.state_dict(),
          "scheduler": scheduler.state_dict() if scheduler is not None else None,
          "rng_cpu": torch.get_rng_state(),
          "rng_cuda": torch.cuda.get_rng_state() if torch.cuda.is_available() else None, "extra": extra}
    path = os.path.join(save_dir, f"resume-tprank-{rank}.pt")
    tmp = path + ".tmp"
    torch.save(st, tmp)
    os.replace(tmp, path)       # atomic: a crash mid-write leaves the previous resume file intact
    return path


def load_resume_state(save_dir: str, rank: int, optimizer, scheduler):
    path = os.path.join(save_dir, f"resume-tprank-{rank}.pt")
    if not os.path.exists(path):
        return None
    st = torch.load(path, map_location="cpu", weights_only=False)
    if hasattr(optimizer, "load_fused_state"):
        dev = optimizer.arena.params.device if getattr(optimizer, "arena", None) is not None else None
        o = st["optimizer"]
        if dev is not None:
            o = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in o.items()}
        optimizer.load_fused_state(o)
    else:
        optimizer.load_state_dict(st["optimizer"])
    if scheduler is not None:
        # fast-forward instead of load_state_dict: the resumed run may use a different --max_steps, and the saved
        # state would overwrite the new schedule length
        optimizer._opt_called = True
        for _ in range(int(st["step"])):
            scheduler.step()
    torch.set_rng_state(st["rng_cpu"])
    if st["rng_cuda"] is not None and torch.cuda.is_available():
        torch.cuda.set_rng_state(st["rng_cuda"])
    return st


# ------------------------------------------------------------------------------------------------ resharding
def _shard_dim(key: str) -> Optional[int]:
    """Which dim of a state_dict entry is sharded over TP ranks (None = replicated)."""
    if key.endswith("norm1.scale") or key.endswith("norm2.scale") or key == "norm.scale":
        return None
    if key == "embedding.weight" or key.startswith("lm_head."):
        return 0
    if ".attn.wo." in key or ".ffn.down_proj." in key:
        return 1 if key.endswith("weight") else None       # row-parallel: weight on dim 1, bias replicated
    return 0                                                # column-parallel weights and biases


def merge_shards(shards: List[Dict[str, torch.Tensor]]) -> Dict[str, torch.Tensor]:
    """TP=k per-rank state_dicts (rank order) -> the full TP=1 state_dict."""
    full = OrderedDict()
    for k in shards[0]:
        d = _shard_dim(k)
        full[k] = shards[0][k].clone() if d is None else torch.cat([s[k] for s in shards], dim=d)
    return full


def split_full(full: Dict[str, torch.Tensor], n: int) -> List[Dict[str, torch.Tensor]]:
    out = []
    for r in range(n):
        sd = OrderedDict()
        for k, v in full.items():
            d = _shard_dim(k)
            sd[k] = v.clone() if d is None else v.chunk(n, dim=d)[r].clone().contiguous()
        out.append(sd)
    return out


def reshard_checkpoint(src_dir: str, dst_dir: str, step: int, src_tp: int, dst_tp: int) -> List[str]:
    """Convert the step-``step`` checkpoint written at TP=src_tp into one loadable at TP=dst_tp."""
    shards, loss_tag = [], "0.0000"
    for r in range(src_tp):
        cands = [p for p in list_checkpoints(src_dir, r) if int(CKPT_RE.search(os.path.basename(p)).group(2)) == step]
        assert cands, f"no checkpoint for rank {r} at iter {step} in {src_dir}"
        loss_tag = CKPT_RE.search(os.path.basename(cands[0])).group(3)
        shards.append(torch.load(cands[0], map_location="cpu", weights_only=True))
    new = split_full(merge_shards(shards), dst_tp)
    os.makedirs(dst_dir, exist_ok=True)
    paths = []
    for r, sd in enumerate(new):
        p = os.path.join(dst_dir, f"tprank-{r}_iter-{step}_loss-{loss_tag}.pth")
        torch.save(sd, p)
        paths.append(p)
    return paths
