"""Fused Adam / AdamW (csrc/train_ops.cu:adam_kernel) behind the torch.optim.Optimizer interface.

The reference trains with ``optim.Adam(lr)`` driven by ``OneCycleLR`` (train.py:83-84), which rewrites ``lr`` AND
``betas[0]`` every step.  ``FusedAdam`` keeps ``param_groups`` so the stock scheduler works unchanged, but the update
itself is ONE kernel launch over the model's flat parameter arena; ``lr``/``beta1``/bias corrections are read from a
small device buffer (refreshed by an async H2D copy), which makes the launch CUDA-graph capturable.  The same launch
rewrites the bf16 shadow weights the tcgen05 GEMMs consume.

The H2D copy reads its pinned source when the copy EXECUTES on the stream, not when it is enqueued, and the host runs
many steps ahead of the GPU under graph replay.  Every step therefore writes its hyper-parameters into its own pinned
slot of a ring; a slot is only rewritten after the event recorded behind its previous copy has completed.

Off the fused path (CPU / parameters not in an arena) it runs the identical math with torch ops, so CPU tests can
compare it against ``torch.optim.Adam`` directly.
"""
from __future__ import annotations

import math
from typing import Optional

import torch

from . import fused


class FusedAdam(torch.optim.Optimizer):
    _RING = 64      # pinned hyper-parameter slots (steps the host may run ahead before it has to wait)

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, arena=None):
        """``arena``: a models.fused_transformer.ParamArena (or None for the generic per-tensor torch fallback).
        ``weight_decay`` is decoupled (AdamW); with the reference's wd = 0 Adam == AdamW exactly."""
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
        super().__init__(params, defaults)
        self.arena = arena
        self.step_count = 0
        self._hyper_host: Optional[torch.Tensor] = None   # [_RING, 8] pinned: one slot per in-flight step
        self._hyper_dev: Optional[torch.Tensor] = None
        self._hyper_events = [None] * self._RING
        self._m = self._v = None
        if arena is not None:
            self._m = torch.zeros_like(arena.params)
            self._v = torch.zeros_like(arena.params)
            self._hyper_host = torch.zeros(self._RING, 8, dtype=torch.float32)
            if arena.params.is_cuda:
                self._hyper_host = self._hyper_host.pin_memory()
            self._hyper_dev = torch.zeros(8, dtype=torch.float32, device=arena.params.device)

    @classmethod
    def for_model(cls, model, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
        """Fused optimizer over the model's arena when the fused path is active, else the generic fallback."""
        arena = None
        p0 = next(model.parameters())
        if p0.is_cuda and hasattr(model, "fused_runtime"):
            from ..models.fused_transformer import fused_supported
            from .dispatch import native_available
            if native_available() and fused_supported(model):
                arena = model.fused_runtime().arena
        return cls(model.parameters(), lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, arena=arena)

    def zero_grad(self, set_to_none: bool = False):
        if self.arena is not None:
            return  # every gradient slot is overwritten by the wgrad kernels each step
        super().zero_grad(set_to_none=set_to_none)

    # -- split in two so a CUDA graph can capture only the launch
    def update_hyper(self):
        """Advance the step counter and upload {lr, beta1, beta2, eps, bc1, bc2, wd} (async, pinned)."""
        self.step_count += 1
        g = self.param_groups[0]
        b1, b2 = g["betas"]
        t = self.step_count
        vals = [g["lr"], b1, b2, g["eps"], 1.0 - b1 ** t, 1.0 - b2 ** t, g["weight_decay"], 0.0]
        if self._hyper_host is not None:
            i = t % self._RING
            ev = self._hyper_events[i]
            if ev is not None:
                ev.synchronize()          # the copy that last read this slot has executed (a no-op unless the host
                                          # is more than _RING steps ahead of the GPU)
            slot = self._hyper_host[i]
            slot.copy_(torch.tensor(vals, dtype=torch.float32))
            self._hyper_dev.copy_(slot, non_blocking=True)
            if self._hyper_dev.is_cuda:
                ev = ev or torch.cuda.Event()
                ev.record()
                self._hyper_events[i] = ev
        return vals

    def launch(self):
        a = self.arena
        fused.adam_step(a.params, a.grads, self._m, self._v, a.shadow, self._hyper_dev)

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        if self.arena is not None:
            self.update_hyper()
            self.launch()
            return loss
        self.step_count += 1
        t = self.step_count
        for g in self.param_groups:
            b1, b2 = g["betas"]
            lr, eps, wd = g["lr"], g["eps"], g["weight_decay"]
            bc1, bc2 = 1.0 - b1 ** t, 1.0 - b2 ** t
            for p in g["params"]:
                if p.grad is None:
                    continue
                st = self.state[p]
                if not st:
                    st["exp_avg"] = torch.zeros_like(p)
                    st["exp_avg_sq"] = torch.zeros_like(p)
                m, v = st["exp_avg"], st["exp_avg_sq"]
                if wd != 0.0:
                    p.mul_(1.0 - lr * wd)
                m.mul_(b1).add_(p.grad, alpha=1.0 - b1)
                v.mul_(b2).addcmul_(p.grad, p.grad, value=1.0 - b2)
                denom = (v.sqrt() / math.sqrt(bc2)).add_(eps)
                p.addcdiv_(m, denom, value=-lr / bc1)
        return loss

    # -- resume support (extra files next to the reference-format checkpoints)
    def fused_state(self):
        if self.arena is None:
            return {"step": self.step_count, "state": self.state_dict()}
        # "layout": the arena's (offset, numel) list -- m / v are flat arenas, only meaningful for the same layout
        return {"step": self.step_count, "m": self._m.clone(), "v": self._v.clone(), "layout": self._layout_sig()}

    def _layout_sig(self):
        return [(off, p.numel()) for p, off in self.arena.entries] if self.arena is not None else None

    def load_fused_state(self, st):
        self.step_count = st["step"]
        if self.arena is None:
            self.load_state_dict(st["state"])
        else:
            lay = st.get("layout")
            if lay is not None and [tuple(x) for x in lay] != self._layout_sig():
                raise RuntimeError("resume file was written with a different parameter-arena layout (another version "
                                   "of the framework or another model): its Adam moments cannot be applied")
            self._m.copy_(st["m"])
            self._v.copy_(st["v"])
