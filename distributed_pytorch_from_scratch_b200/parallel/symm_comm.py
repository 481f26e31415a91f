"""Symmetric-memory communicator: NVLink peer / NVLS multicast buffers + the kernels that use them.

torch.distributed._symmetric_memory is used ONLY to allocate and exchange the buffers (rendezvous over the existing
process group: peer pointers + one multicast address through the NVSwitch); all data movement is done by our own
kernels (csrc/comm.cu, the FUSE_AR variant in csrc/gemm_tcgen05.cu):

    all_reduce(t)                      two-shot NVLS all-reduce (multimem.ld_reduce + multimem.st), in place when ``t``
                                       already lives in the symmetric buffer, staged through it otherwise
    all_gather_rows(t)                 multimem.st broadcast of each rank's block
    gemm_allreduce(a, w, b_mn_major)   ONE kernel: tcgen05 GEMM whose tiles are reduced across ranks through the
                                       switch as they complete (RowParallelLinear fwd, ColumnParallelLinear dgrad)
    gemm_reduce_scatter(a, w, ...)     ONE kernel: the sequence-parallel form -- every rank keeps only its rows of the
                                       sum (opt-in with DPFS_FUSE_RS=1 until it has been timed against the split path)

Synchronisation uses binary semaphores in a peer-mapped flag buffer (no epochs -> safe under CUDA-graph replay).
All ranks must issue the same sequence of collectives (SPMD), on one stream.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional

import torch
import torch.distributed as dist

from ..ops import _lib
from . import process_manager as pm

BF16, F32 = torch.bfloat16, torch.float32


class SymmZone:
    """One region of a symmetric allocation: this rank's bytes, their multicast alias (0 without NVLS) and every
    rank's address of the same region."""

    def __init__(self, local_u8: torch.Tensor, mc_ptr: int, peer_ptrs):
        self.local, self.mc, self.peer_ptrs = local_u8, mc_ptr, peer_ptrs
        self.nbytes = local_u8.numel()

    def view(self, shape, dtype=BF16) -> torch.Tensor:
        numel = 1
        for d in shape:
            numel *= d
        nbytes = numel * torch.empty((), dtype=dtype).element_size()
        assert nbytes <= self.nbytes, "symmetric zone too small"
        return self.local[:nbytes].view(dtype).view(*shape)


class SymmComm:
    def __init__(self, data, data_hdl, flags, flags_hdl, rank, n):
        self.data, self.hdl = data, data_hdl
        self.flags, self.flags_hdl = flags, flags_hdl
        self.rank, self.n = rank, n
        self.base = data.data_ptr()
        self.capacity = data.numel()
        self.mc_base = int(getattr(data_hdl, "multicast_ptr", 0) or 0)
        self.flag_ptrs = (ctypes.c_ulonglong * 8)(*([int(p) for p in flags_hdl.buffer_ptrs] + [0] * (8 - n)))
        self.data_ptrs = (ctypes.c_ulonglong * 8)(*([int(p) for p in data_hdl.buffer_ptrs] + [0] * (8 - n)))
        # layout: [0, ag_off) collective inputs / GEMM partials | [ag_off, gather_off) sequence-parallel all-gather
        # landing zone | [gather_off, capacity) small all-gathers (cross-entropy statistics)
        self.gather_off = self.capacity - self.capacity // 16
        self.ag_off = (self.capacity // 2) // 256 * 256
        self.ag_bytes = self.gather_off - self.ag_off
        # "push" all-gather (needs the multicast mapping; DPFS_FUSE_AG=0 turns it off): producers write their bf16 block
        # straight into every rank's landing zone.  The landing area is then split: first half for the stand-alone
        # all-gather kernels, second half = two alternating push zones (see push_block).
        self.push_enabled = bool(self.mc_base) and os.environ.get("DPFS_FUSE_AG", "1") == "1"
        if self.push_enabled:
            half = (self.ag_bytes // 2) // 256 * 256
            self.push_off = self.ag_off + half
            self.push_zone_bytes = ((self.ag_bytes - half) // 2) // 256 * 256
            self.ag_bytes = half
            self._push_zone = 0
        self.closed = False
        self._zone_allocs = []
        self._poison = os.environ.get("DPFS_SYMM_POISON", "0") == "1"
        self._poison_half = (self.ag_off // 2) // 256 * 256
        self._poison_side, self._poison_used = 0, 0
        # data path of the reduce: "p2p" (plain NVLink loads/stores, rank-ordered fp32 sums) or "nvls" (multimem.*)
        # Measured (profiles/comm_paths_r1.md): P2P wins at TP=2 (100 vs 116 us for 32.8 MB), NVLS wins at TP=8 (123 vs
        # 143 us; NCCL 158 us) because the switch-side reduction moves S(1+1/n) bytes instead of 2S(n-1)/n.
        default_mode = "nvls" if (n >= 8 and self.mc_base) else "p2p"
        self.mode = os.environ.get("DPFS_AR_MODE", default_mode)
        if n not in (2, 4, 8):
            self.mode = "nvls"
        if self.mode == "nvls" and not self.mc_base:
            raise RuntimeError("NVLS multicast is not available for this process group")

    # ------------------------------------------------------------------------------------------ construction
    @classmethod
    def create(cls, data_bytes: Optional[int] = None) -> "SymmComm":
        import torch.distributed._symmetric_memory as symm_mem
        n, rank, group = pm.pgm.tp_size, pm.pgm.tp_rank, pm.pgm.tp_group
        assert n > 1 and dist.get_backend(group) == "nccl"
        dev = torch.device("cuda", torch.cuda.current_device())
        if data_bytes is None:
            data_bytes = int(os.environ.get("DPFS_SYMM_MB", "512")) << 20
        name = group.group_name
        try:
            symm_mem.enable_symm_mem_for_group(name)
        except Exception:  # noqa: BLE001  (not needed / deprecated on newer torch)
            pass
        data = symm_mem.empty(data_bytes, dtype=torch.uint8, device=dev)
        hdl = symm_mem.rendezvous(data, name)
        words = _lib.lib().dpfs_comm_flag_words()
        flags = symm_mem.empty(words, dtype=torch.int32, device=dev)
        flags.zero_()
        fh = symm_mem.rendezvous(flags, name)
        torch.cuda.synchronize()
        dist.barrier(group=group)
        return cls(data, hdl, flags, fh, rank, n)

    def close(self):
        self.closed = True

    # ------------------------------------------------------------------------------------------ activation zones
    def alloc_zones(self, count: int, zone_bytes: int):
        """``count`` persistent symmetric regions of ``zone_bytes`` each (one extra symmetric allocation, rendezvoused
        over the TP group: every rank must call this at the same point with the same arguments).  Returns a list of
        ``SymmZone``.  The sequence-parallel runtime keeps one zone per gathered activation that the backward pass
        needs again (h1 / h2 of every layer), so nothing is cloned out of a recycled landing buffer."""
        import torch.distributed._symmetric_memory as symm_mem
        zone_bytes = (zone_bytes + 255) // 256 * 256
        dev = self.data.device
        buf = symm_mem.empty(count * zone_bytes, dtype=torch.uint8, device=dev)
        hdl = symm_mem.rendezvous(buf, pm.pgm.tp_group.group_name)
        mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
        ptrs = [int(x) for x in hdl.buffer_ptrs]
        self._zone_allocs.append((buf, hdl))     # keep the allocation (and its peer mappings) alive
        return [SymmZone(buf[i * zone_bytes:(i + 1) * zone_bytes], mc + i * zone_bytes if mc else 0,
                         [q + i * zone_bytes for q in ptrs]) for i in range(count)]

    # ------------------------------------------------------------------------------------------ buffers
    def buffer(self, shape, dtype, offset_bytes: int = 0) -> torch.Tensor:
        """A tensor view into this rank's symmetric buffer (producers write collective inputs here directly)."""
        numel = 1
        for s in shape:
            numel *= s
        nbytes = numel * torch.empty((), dtype=dtype).element_size()
        assert offset_bytes % 256 == 0 and offset_bytes + nbytes <= self.ag_off, "symmetric buffer too small"
        if self._poison and offset_bytes == 0 and nbytes <= self._poison_half:
            # Debug mode (DPFS_SYMM_POISON=1): results handed out as views of this buffer are only valid until the next
            # collective.  Alternate between the two halves and fill the one retired by this call with NaNs, so a
            # consumer that reads a stale view gets NaNs (and a NaN loss) instead of silently reading newer data.
            self._poison_side ^= 1
            offset_bytes = self._poison_side * self._poison_half
            old = (1 - self._poison_side) * self._poison_half
            if self._poison_used:
                self.data[old:old + self._poison_used].view(torch.int16).fill_(0x7FC0)      # bf16 NaN pattern
            self._poison_used = (nbytes + 1) // 2 * 2
        return self.data[offset_bytes:offset_bytes + nbytes].view(dtype).view(*shape)

    def _inside(self, t: torch.Tensor) -> bool:
        p = t.data_ptr()
        return self.base <= p and p + t.numel() * t.element_size() <= self.base + self.capacity

    def supports(self, t: torch.Tensor) -> bool:
        nbytes = t.numel() * t.element_size()
        return (not self.closed and t.is_cuda and t.dtype in (BF16, F32) and nbytes % 16 == 0
                and nbytes <= self.ag_off and t.is_contiguous())

    # ------------------------------------------------------------------------------------------ collectives
    def all_reduce(self, t: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        nbytes = t.numel() * t.element_size()
        if self._inside(t):
            work = t
        else:
            work = self.buffer(t.shape, t.dtype)
            work.copy_(t)
        off = work.data_ptr() - self.base
        if self.mode == "nvls":
            rc = _lib.lib().dpfs_allreduce_nvls(self.mc_base + off, nbytes, int(t.dtype == BF16), self.flag_ptrs,
                                                self.rank, self.n, _lib.stream())
        else:
            rc = _lib.lib().dpfs_allreduce_p2p(self.data_ptrs, off, nbytes, int(t.dtype == BF16), self.flag_ptrs,
                                               self.rank, self.n, _lib.stream())
        _lib.check(rc, "dpfs_allreduce")
        if out is None:
            return work if work is t else work.clone()
        if out.data_ptr() != work.data_ptr():
            out.copy_(work)
        return out

    def can_gather(self, t: torch.Tensor) -> bool:
        nbytes = t.numel() * t.element_size()
        return (not self.closed and nbytes > 0 and nbytes % 16 == 0
                and nbytes * self.n <= self.capacity - self.gather_off)

    def all_gather_rows(self, t: torch.Tensor) -> torch.Tensor:
        """(...) -> (n, ...), rank order; the result is a private copy."""
        t = t.contiguous()
        nbytes = t.numel() * t.element_size()
        pad = (-nbytes) % 16
        assert pad == 0 and nbytes * self.n <= self.capacity - self.gather_off
        if self.mc_base and self.mode == "nvls":
            rc = _lib.lib().dpfs_allgather_nvls(t.data_ptr(), self.mc_base + self.gather_off, nbytes, self.flag_ptrs,
                                                self.rank, self.n, _lib.stream())
        else:
            rc = _lib.lib().dpfs_allgather_p2p(t.data_ptr(), self.data_ptrs, self.gather_off, nbytes, self.flag_ptrs,
                                               self.rank, self.n, _lib.stream())
        _lib.check(rc, "dpfs_allgather")
        view = self.data[self.gather_off:self.gather_off + nbytes * self.n].view(t.dtype).view((self.n,) + t.shape)
        return view.clone()

    def can_fuse_shape(self, M: int, N: int) -> bool:
        """Whether an [M, N] bf16 output can go through the fused GEMM + all-reduce kernel."""
        if self.closed or self.n not in (2, 4, 8) or (self.mode == "nvls" and not self.mc_base):
            return False
        tiles = ((M + 127) // 128) * ((N + 127) // 128)
        return M > 0 and N % 8 == 0 and M * N * 2 <= self.ag_off and tiles <= 16384

    def can_fuse_gemm_ar(self, a: torch.Tensor, w: torch.Tensor, b_mn_major: bool = False) -> bool:
        if a.dtype != BF16 or w.dtype != BF16 or a.stride(0) % 8 or w.stride(0) % 8:
            return False
        return self.can_fuse_shape(a.shape[0], w.shape[1] if b_mn_major else w.shape[0])

    def gemm_allreduce(self, a: torch.Tensor, w: torch.Tensor, b_mn_major: bool = False, bn: int = 0) -> torch.Tensor:
        """all_reduce(a @ w^T) (or a @ w when ``b_mn_major``) in ONE kernel; the result is a view of the symmetric
        buffer, valid until the next collective — consume it right away."""
        M, K = a.shape
        N = w.shape[1] if b_mn_major else w.shape[0]
        out = self.buffer((M, N), BF16)
        off = out.data_ptr() - self.base
        mc = self.mc_base + off if self.mode == "nvls" else 0
        rc = _lib.lib().dpfs_gemm_allreduce_bf16(a.data_ptr(), a.stride(0), w.data_ptr(), w.stride(0), int(b_mn_major),
                                                 out.data_ptr(), mc, self.data_ptrs, off, M, N, K, self.flag_ptrs,
                                                 self.rank, self.n, bn, _lib.stream())
        _lib.check(rc, "dpfs_gemm_allreduce_bf16")
        return out

    # ------------------------------------------------------------------------------------------ sequence parallel
    def can_fuse_gemm_rs(self, a: torch.Tensor, w: torch.Tensor, b_mn_major: bool = False) -> bool:
        """GEMM -> reduce-scatter in one kernel (both data paths: P2P pulls or NVLS multimem.ld_reduce)."""
        if self.closed or a.dtype != BF16 or w.dtype != BF16 or self.n not in (2, 4, 8):
            return False
        M = a.shape[0]
        N = w.shape[1] if b_mn_major else w.shape[0]
        tiles = ((M + 127) // 128) * ((N + 127) // 128)
        return (N % 8 == 0 and M % self.n == 0 and M // self.n >= 128 and M * N * 2 <= self.ag_off
                and tiles <= 16384 and a.stride(0) % 8 == 0 and w.stride(0) % 8 == 0)

    def gemm_reduce_scatter(self, a: torch.Tensor, w: torch.Tensor, b_mn_major: bool = False,
                            bn: int = 0) -> torch.Tensor:
        """This rank's rows of sum_r (a_r @ w_r^T) (or a_r @ w_r when ``b_mn_major``): bf16 [M / n, N], private memory.
        ONE kernel: the partial tiles land in the symmetric buffer and the comm warps of the same CTAs reduce the rows
        their rank owns while the tensor cores compute the next tiles."""
        M, K = a.shape
        N = w.shape[1] if b_mn_major else w.shape[0]
        part = self.buffer((M, N), BF16)
        out = torch.empty((M // self.n, N), dtype=BF16, device=a.device)
        off = part.data_ptr() - self.base
        mc = self.mc_base + off if self.mode == "nvls" else 0
        rc = _lib.lib().dpfs_gemm_reduce_scatter_bf16(a.data_ptr(), a.stride(0), w.data_ptr(), w.stride(0),
                                                      int(b_mn_major), part.data_ptr(), mc, self.data_ptrs, off,
                                                      out.data_ptr(), M, N, K, self.flag_ptrs, self.rank, self.n, bn,
                                                      _lib.stream())
        _lib.check(rc, "dpfs_gemm_reduce_scatter_bf16")
        return out

    def reduce_scatter_rows(self, partial: torch.Tensor) -> torch.Tensor:
        """partial: bf16 [M, N] living in the symmetric buffer (each rank's partial sums) -> this rank's reduced block
        of rows, bf16 [M / n, N] in private memory."""
        assert self._inside(partial) and partial.dtype == BF16 and partial.is_contiguous()
        M, N = partial.shape
        assert M % self.n == 0
        out = torch.empty((M // self.n, N), dtype=BF16, device=partial.device)
        off = partial.data_ptr() - self.base
        if self.mode == "nvls":
            rc = _lib.lib().dpfs_reduce_scatter_nvls(self.mc_base + off, M * N * 2, out.data_ptr(), self.flag_ptrs,
                                                     self.rank, self.n, _lib.stream())
        else:
            assert self.n in (2, 4, 8), "the P2P reduce-scatter kernel is instantiated for 2, 4 and 8 ranks"
            rc = _lib.lib().dpfs_reduce_scatter_p2p(self.data_ptrs, off, M * N * 2, out.data_ptr(), self.flag_ptrs,
                                                    self.rank, self.n, _lib.stream())
        _lib.check(rc, "dpfs_reduce_scatter")
        return out

    def all_gather_into_buffer(self, local: torch.Tensor) -> torch.Tensor:
        """local: [Ml, ...] (any dtype) -> view [n * Ml, ...] of the symmetric landing zone holding every rank's block
        in rank order.  The view is overwritten by the next all-gather: consume (or clone) it first."""
        local = local.contiguous()
        nbytes = local.numel() * local.element_size()
        assert nbytes % 16 == 0 and nbytes * self.n <= self.ag_bytes, "all-gather landing zone too small"
        if self.mode == "nvls":
            rc = _lib.lib().dpfs_allgather_nvls(local.data_ptr(), self.mc_base + self.ag_off, nbytes, self.flag_ptrs,
                                                self.rank, self.n, _lib.stream())
        else:
            rc = _lib.lib().dpfs_allgather_p2p(local.data_ptr(), self.data_ptrs, self.ag_off, nbytes, self.flag_ptrs,
                                               self.rank, self.n, _lib.stream())
        _lib.check(rc, "dpfs_allgather")
        shape = (self.n * local.shape[0],) + tuple(local.shape[1:])
        return self.data[self.ag_off:self.ag_off + nbytes * self.n].view(local.dtype).view(shape)

    # ------------------------------------------------------------------------------------------ push all-gather
    def can_push(self, block_rows: int, row_bytes: int) -> bool:
        nbytes = block_rows * row_bytes
        return (self.push_enabled and not self.closed and nbytes % 16 == 0
                and nbytes * self.n <= self.push_zone_bytes)

    def push_block(self, block_rows: int, cols: int, dtype=BF16):
        """Next landing zone for a pushed all-gather of [block_rows, cols] blocks: returns (multicast address of THIS
        rank's block, view [n * block_rows, cols] of the whole zone in this rank's copy).  The producer kernel writes
        through the multicast address; call ``push_done()`` before reading the view.

        Two zones alternate.  A rank can only be writing gather k+2 into a zone after it has passed the barrier of
        gather k+1, which every rank enters only after it has finished (in stream order) reading gather k from that
        zone -- so one barrier per gather is enough and no rank overwrites data a peer still reads."""
        esz = torch.empty((), dtype=dtype).element_size()
        nbytes = block_rows * cols * esz
        assert self.can_push(block_rows, cols * esz), "push all-gather zone too small (or DPFS_FUSE_AG off)"
        self._push_zone ^= 1
        off = self.push_off + self._push_zone * self.push_zone_bytes
        view = self.data[off:off + nbytes * self.n].view(dtype).view(self.n * block_rows, cols)
        return self.mc_base + off + self.rank * nbytes, view

    def push_zone(self, rows: int, cols: int, dtype=BF16):
        """Like push_block but for producers that address the whole [rows, cols] zone themselves (the vocab-parallel
        embedding scatters token rows): returns (multicast address of the zone, view of this rank's copy)."""
        esz = torch.empty((), dtype=dtype).element_size()
        nbytes = rows * cols * esz
        assert self.push_enabled and nbytes % 16 == 0 and nbytes <= self.push_zone_bytes
        self._push_zone ^= 1
        off = self.push_off + self._push_zone * self.push_zone_bytes
        return self.mc_base + off, self.data[off:off + nbytes].view(dtype).view(rows, cols)

    def push_done(self):
        """Rank-wide barrier after a pushed all-gather: every rank's block is complete and visible."""
        self.barrier()

    def barrier(self):
        _lib.check(_lib.lib().dpfs_comm_barrier(self.flag_ptrs, self.rank, self.n, _lib.stream()), "dpfs_comm_barrier")
