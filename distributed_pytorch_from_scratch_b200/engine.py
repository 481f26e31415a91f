"""Training-step executor: H2D prefetch of the batch, fused forward/backward/optimizer, optional CUDA-graph replay.

``TrainStep`` is the public "one training step" API used by train.py and bench.py:

    step = TrainStep(model, optimizer, scheduler, use_graph=True)
    loss = step(batch["input_ids"], batch["position_ids"], batch["target_ids"])   # host (pinned) or device tensors

With ``use_graph`` the whole step (every kernel of forward, backward and the fused Adam launch, plus the TP
collectives) is captured once per batch shape into a CUDA graph and replayed, so the ~400 launches of a step cost one
``cudaGraphLaunch`` — the step is launch-bound otherwise (reference train.py:94-110 is plain eager PyTorch).
Per-step hyper-parameters (lr, beta1 from OneCycleLR, bias corrections) live in a device buffer updated by a tiny
async H2D copy before each replay.  Inputs are copied into static device buffers with non-blocking copies from
pinned memory (reference K17: three blocking ``.cuda()`` copies).
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch

from .ops import _lib
from .ops.optim import FusedAdam


class TrainStep:
    def __init__(self, model, optimizer, scheduler=None, ignore_index: int = -1, use_graph: bool = False,
                 autocast_bf16: bool = False, max_graphs: int = 4, bucket: int = 0):
        """``bucket`` > 0: sequence lengths are rounded up to a multiple of ``bucket`` (capped at the model's maxlen)
        before the step runs: the extra positions repeat the last input id, continue the position ids and carry
        ``ignore_index`` targets.  With causal attention and an ignore-aware mean loss that changes neither the loss nor
        any gradient, but the loader's ragged batches collapse onto a handful of shapes, so every step is a CUDA-graph
        replay instead of a few hundred eager launches (measured at TP=2: 59 ms / step eager on ragged batches against
        ~14 ms replayed)."""
        self.model, self.optimizer, self.scheduler = model, optimizer, scheduler
        self.ignore_index = ignore_index
        self.device = next(model.parameters()).device
        self.autocast_bf16 = autocast_bf16
        fused_opt = isinstance(optimizer, FusedAdam) and optimizer.arena is not None
        self.use_graph = bool(use_graph and self.device.type == "cuda" and fused_opt)
        from .parallel import process_manager as pm
        if pm.pgm is not None and pm.pgm.tp_size > 1 and pm.pgm.symm is None:
            # NCCL collectives inside a captured step hang on this stack; only our own NVLink kernels are captured
            self.use_graph = False
        self.fused_opt = fused_opt
        self.max_graphs = max_graphs
        self.bucket = int(bucket)
        self.maxlen = int(getattr(model, "maxlen", 0) or 0)
        self._graphs: Dict[Tuple[int, int], dict] = {}
        self._pool = None
        self._cap_stream = None
        self._static: Dict[Tuple[int, int], tuple] = {}
        self.kernels_per_step = 0
        self.steps = 0

    # ------------------------------------------------------------------------------------------- helpers
    def _buffers(self, shape):
        st = self._static.get(shape)
        if st is None:
            st = tuple(torch.zeros(shape, dtype=torch.int64, device=self.device) for _ in range(3))
            self._static[shape] = st
        return st

    def _loss_backward(self, ids, pos, tgt):
        if self.autocast_bf16 and self.device.type == "cuda":
            with torch.autocast("cuda", dtype=torch.bfloat16):
                loss = self.model.loss(ids, pos, tgt, self.ignore_index)
        else:
            loss = self.model.loss(ids, pos, tgt, self.ignore_index)
        loss.backward()
        return loss

    def _eager(self, ids, pos, tgt):
        self.optimizer.zero_grad()
        loss = self._loss_backward(ids, pos, tgt)
        self.optimizer.step()
        return loss.detach()

    def _capture(self, shape, bufs):
        ids, pos, tgt = bufs
        opt = self.optimizer
        # warm-up on a side stream: allocates every cached scratch buffer and configures every kernel
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(2):
                self._loss_backward(ids, pos, tgt)
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        before = _lib.lib().dpfs_launch_count() if _lib.available() else 0
        # thread_local: NCCL's watchdog thread may poll events while we capture (it never touches our stream)
        kw = {"capture_error_mode": "thread_local"}
        # capture on a high-priority stream: kernels that the model forks onto its (default-priority) side stream -- the
        # weight-gradient GEMMs of the sequence-parallel hand-off path -- then only take SMs the critical path leaves idle
        if self._cap_stream is None:
            self._cap_stream = torch.cuda.Stream(priority=-1)
        kw["stream"] = self._cap_stream
        if self._pool is not None:
            kw["pool"] = self._pool
        with torch.cuda.graph(g, **kw):
            loss = self._loss_backward(ids, pos, tgt)
            opt.launch()
            loss = loss.detach()
        if self._pool is None:
            self._pool = g.pool()
        self.kernels_per_step = (_lib.lib().dpfs_launch_count() - before) if _lib.available() else 0
        return {"graph": g, "loss": loss}

    # ------------------------------------------------------------------------------------------- public
    def _padded_len(self, T: int) -> int:
        if self.bucket <= 0 or T % self.bucket == 0:
            return T
        Tp = (T + self.bucket - 1) // self.bucket * self.bucket
        if self.maxlen and T <= self.maxlen:
            Tp = min(Tp, self.maxlen)          # position ids must stay inside the RoPE table
        return Tp

    def load_batch(self, input_ids, position_ids, target_ids):
        """Async copy of a batch (host pinned or device) into the static device buffers; returns them."""
        B, T = input_ids.shape
        Tp = self._padded_len(T)
        shape = (B, Tp)
        bufs = self._buffers(shape)
        if Tp == T:
            for dst, src in zip(bufs, (input_ids, position_ids, target_ids)):
                dst.copy_(src, non_blocking=True)
            return shape, bufs
        ids, pos, tgt = bufs
        ids[:, :T].copy_(input_ids, non_blocking=True)
        pos[:, :T].copy_(position_ids, non_blocking=True)
        tgt[:, :T].copy_(target_ids, non_blocking=True)
        ids[:, T:] = ids[:, T - 1:T]                                             # any valid id: never a target
        pos[:, T:] = pos[:, T - 1:T] + torch.arange(1, Tp - T + 1, device=pos.device, dtype=pos.dtype)
        tgt[:, T:] = self.ignore_index
        return shape, bufs

    def run_loaded(self, shape, bufs) -> torch.Tensor:
        """One step on a batch that is already in the static buffers."""
        ids, pos, tgt = bufs
        if self.use_graph:
            entry = self._graphs.get(shape)
            if entry is None and len(self._graphs) < self.max_graphs:
                entry = self._capture(shape, bufs)
                self._graphs[shape] = entry
            if entry is not None:
                self.optimizer.update_hyper()
                entry["graph"].replay()
                loss = entry["loss"]
            else:
                loss = self._eager(ids, pos, tgt)
        else:
            before = _lib.lib().dpfs_launch_count() if (self.device.type == "cuda" and _lib.available()) else 0
            loss = self._eager(ids, pos, tgt)
            if self.device.type == "cuda" and _lib.available():
                self.kernels_per_step = _lib.lib().dpfs_launch_count() - before
        if self.scheduler is not None:
            self.scheduler.step()
        self.steps += 1
        return loss

    def __call__(self, input_ids, position_ids, target_ids) -> torch.Tensor:
        shape, bufs = self.load_batch(input_ids, position_ids, target_ids)
        return self.run_loaded(shape, bufs)
