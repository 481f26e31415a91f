"""Optional TensorBoard writer (tensorboardX -> torch.utils.tensorboard -> no-op) and CUDA-event step timers.

The reference hard-depends on tensorboardX (train.py:15, defect D11) and has no timers at all."""
from __future__ import annotations

import time
from typing import Optional

import torch


class _NullWriter:
    def add_scalar(self, *a, **k):
        pass

    def close(self):
        pass


def make_summary_writer(log_dir: str):
    try:
        from tensorboardX import SummaryWriter  # type: ignore
        return SummaryWriter(log_dir=log_dir)
    except Exception:  # noqa: BLE001
        pass
    try:
        from torch.utils.tensorboard import SummaryWriter
        return SummaryWriter(log_dir=log_dir)
    except Exception:  # noqa: BLE001
        return _NullWriter()


class StepTimer:
    """Device-timed step intervals (CUDA events on the current stream); wall-clock on CPU."""

    def __init__(self, device: torch.device):
        self.cuda = device.type == "cuda"
        self._start = None
        self._t0 = 0.0

    def start(self):
        if self.cuda:
            self._start = torch.cuda.Event(enable_timing=True)
            self._start.record()
        else:
            self._t0 = time.perf_counter()

    def stop_ms(self) -> float:
        if self.cuda:
            end = torch.cuda.Event(enable_timing=True)
            end.record()
            end.synchronize()
            return self._start.elapsed_time(end)
        return (time.perf_counter() - self._t0) * 1e3


def max_over_ranks(value: float, device: torch.device, group=None) -> float:
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return value
    t = torch.tensor([value], dtype=torch.float64, device=device if device.type == "cuda" else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())


class nvtx_range:
    """``with nvtx_range("layer3"):`` — NVTX ranges for nsys / ncu windows; free when CUDA is absent."""

    def __init__(self, name: str):
        self.name = name
        self.on = torch.cuda.is_available()

    def __enter__(self):
        if self.on:
            torch.cuda.nvtx.range_push(self.name)

    def __exit__(self, *exc):
        if self.on:
            torch.cuda.nvtx.range_pop()
