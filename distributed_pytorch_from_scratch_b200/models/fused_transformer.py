"""Hand-scheduled forward / backward of the tensor-parallel Transformer on the sm_100a kernel library.

Memory layout (per rank): all parameters live in ONE flat fp32 arena, their gradients in a second arena with the
same offsets, and a bf16 "shadow" arena (what the tcgen05 GEMMs read) in a third.  The arena order places
``wq|wk|wv`` and ``gate|up`` of a layer back to back, so a layer's QKV projection and its gate/up projection are
each ONE GEMM over a contiguous ``[3*Dl, d]`` / ``[2*Fl, d]`` weight — while ``state_dict()`` still exposes the
separate reference keys (the nn.Parameters are views into the arena; checkpoint format unchanged).
The fused Adam kernel updates the fp32 arena and rewrites the bf16 shadow in the same pass.

One decoder layer forward (M = b*t rows, per rank; ref models/model.py:116-121 and SURVEY §3.2):
    h1   = rmsnorm(x)                               fused norm kernel, bf16 out, saves rstd
    qkv  = h1 @ [Wq;Wk;Wv]^T + b                    tcgen05 GEMM, bias in the epilogue
    qkv  = rope(qkv)                                in place on the q,k segments
    a    = flash_attn(qkv)                          tcgen05 flash attention, heads split by TMA maps
    x2   = x + allreduce(a @ Wo^T) + bo             GEMM (+ residual/bias epilogue at TP=1) / fused GEMM+AR
    h2   = rmsnorm(x2)
    gu   = h2 @ [Wg;Wu]^T + b ; act = silu(g) * u
    x3   = x2 + allreduce(act @ Wd^T) + bd
Backward mirrors it with dgrad (MN-major weight operand, no transposes), wgrad (split-K over tokens, fp32 straight
into the grad arena), ONE all-reduce per column-parallel group (q/k/v share one, gate/up share one: the reference
does five per layer, defect D7).
"""
from __future__ import annotations

import math
import os
from typing import List, Optional

import torch

from ..ops import fused
from ..ops.attention import attn_bwd, attn_fwd
from ..ops.gemm import gemm_nt, gemm_swiglu, gemm_swiglu_bwd
from ..parallel import comm_ops
from ..parallel import process_manager as pm
from ..utils.logging import nvtx_range

BF16, F32 = torch.bfloat16, torch.float32


def _nvtx(label):
    """NVTX range around one phase of the hand-scheduled step (shows up in nsys / as ncu --nvtx-include windows).
    ``label`` may contain ``{i}``: the layer index of the LayerPack found among the call's arguments / ctx."""
    def deco(fn):
        def wrapped(*a, **k):
            idx = ""
            if "{i}" in label:
                pack = next((x for x in a if isinstance(x, LayerPack)), None) or getattr(a[0], "pack", None)
                idx = getattr(pack, "index", "")
            with nvtx_range(label.format(i=idx)):
                return fn(*a, **k)
        wrapped.__name__ = fn.__name__
        return wrapped
    return deco


_ALIGN = 64  # elements; keeps every group 256 B (fp32) / 128 B (bf16) aligned for TMA and vector access


def fused_unsupported_reason(model) -> Optional[str]:
    """None when the hand-scheduled sm_100a path can run this model, else a sentence naming the limit it hits."""
    n = pm.pgm.tp_size
    hd = model.head_dim
    d, f, V = model.attn_dim, model.ffn_dim, model.vocab_size
    if hd not in (64, 128):
        return f"head_dim {hd} is not 64 or 128 (the tcgen05 flash-attention kernels are built for those two)"
    if d % 8:
        return f"attn_dim {d} is not a multiple of 8 (16-byte vector access / TMA alignment)"
    if d > 2048:
        return f"attn_dim {d} > 2048 (the fused norm kernels keep one row of the residual stream in registers)"
    if (d // n) % 8 or (f // n) % 8:
        return f"attn_dim / tp ({d // n}) and ffn_dim / tp ({f // n}) must be multiples of 8"
    if V % n or (V // n) % 8:
        return f"vocab_size {V} must split into multiples of 8 over tp_size {n}"
    return None


_WARNED = set()


def fused_supported(model) -> bool:
    """Whether the fused path runs.  A model that misses a limit drops to the module path (PyTorch attention with a
    materialised t x t mask, cuBLAS GEMMs: the reference's own path) -- never silently: the reason is printed once."""
    why = fused_unsupported_reason(model)
    if why is not None and why not in _WARNED and model.attn_dim > 0:
        _WARNED.add(why)
        if next(model.parameters()).is_cuda:
            print(f"[dpfs] WARNING: the fused sm_100a Transformer path is OFF for this model: {why}. "
                  f"Falling back to the PyTorch module path (much slower).", flush=True)
    return why is None


# ================================================================================================ parameter arena
class ParamArena:
    """Re-homes every parameter of a Transformer into flat fp32 / grad / bf16-shadow arenas (views keep the API)."""

    def __init__(self, model):
        groups = []
        for layer in model.layers:
            a, f = layer.attn, layer.ffn
            groups += [[a.wq.weight, a.wk.weight, a.wv.weight], [a.wq.bias, a.wk.bias, a.wv.bias],
                       [a.wo.weight], [f.gate_proj.weight, f.up_proj.weight],
                       [f.gate_proj.bias, f.up_proj.bias], [f.down_proj.weight]]
        groups += [[model.embedding.weight], [model.lm_head.weight], [model.lm_head.bias]]
        # every REPLICATED parameter (RMSNorm scales, the biases of the row-parallel wo / down_proj) in ONE contiguous
        # block: under sequence parallelism their gradients are partial sums over the local rows and are all-reduced
        # with a single small collective at the end of the backward pass
        norm_group = [s_ for layer in model.layers for s_ in (layer.norm1.scale, layer.norm2.scale)]
        norm_group.append(model.norm.scale)
        norm_group += [b_ for layer in model.layers for b_ in (layer.attn.wo.bias, layer.ffn.down_proj.bias)]
        groups.append(norm_group)
        seen = set()
        for g in groups:
            for p in g:
                assert p is not None and p.dtype == F32, "the fused path keeps fp32 master parameters"
                seen.add(id(p))
        assert len(seen) == sum(1 for _ in model.parameters()), "unexpected parameter set for the fused path"
        self.device = groups[0][0].device
        offs, total = [], 0
        for g in groups:
            total = (total + _ALIGN - 1) // _ALIGN * _ALIGN
            for p in g:
                offs.append(total)
                total += p.numel()
        total = (total + _ALIGN - 1) // _ALIGN * _ALIGN
        self.numel = total
        self.params = torch.zeros(total, dtype=F32, device=self.device)
        self.grads = torch.zeros(total, dtype=F32, device=self.device)
        self.shadow = torch.zeros(total, dtype=BF16, device=self.device)
        self.entries = []  # (param, offset)
        flat = [p for g in groups for p in g]
        with torch.no_grad():
            for p, off in zip(flat, offs):
                n = p.numel()
                view = self.params[off:off + n].view(p.shape)
                view.copy_(p.data)
                p.data = view
                p.grad = self.grads[off:off + n].view(p.shape)
                p._dpfs_shadow_view = self.shadow[off:off + n].view(p.shape)
                p._dpfs_arena_off = off
                self.entries.append((p, off))
        first, last = norm_group[0], norm_group[-1]
        self.norm_grads = self.grads[first._dpfs_arena_off:last._dpfs_arena_off + last.numel()]
        self._sig = None
        self.sync_shadow(force=True)

    def view(self, which: torch.Tensor, first, *shape):
        off = first._dpfs_arena_off
        n = 1
        for s in shape:
            n *= s
        return which[off:off + n].view(*shape)

    def valid(self) -> bool:
        base = self.params.data_ptr()
        for p, off in (self.entries[0], self.entries[-1]):
            if p.data_ptr() != base + off * 4 or p.dtype != F32:
                return False
        return True

    def _signature(self):
        return sum(p._version for p, _ in self.entries)

    def sync_shadow(self, force: bool = False):
        """Re-cast fp32 -> bf16 if anything other than the fused optimizer touched the parameters."""
        sig = self._signature()
        if force or sig != self._sig:
            fused.cast_to_bf16(self.params, out=self.shadow)
            self._sig = sig

    def reattach_grads(self):
        for p, off in self.entries:
            if p.grad is None or p.grad.data_ptr() != self.grads.data_ptr() + off * 4:
                p.grad = self.grads[off:off + p.numel()].view(p.shape)


class LayerPack:
    """Fused views of one decoder layer's parameters / grads (no copies)."""

    def __init__(self, arena: ParamArena, layer, d: int, Dl: int, Fl: int):
        a, f = layer.attn, layer.ffn
        P, G, S = arena.params, arena.grads, arena.shadow
        v = arena.view
        self.wqkv, self.g_wqkv = v(S, a.wq.weight, 3 * Dl, d), v(G, a.wq.weight, 3 * Dl, d)
        self.bqkv, self.g_bqkv = v(P, a.wq.bias, 3 * Dl), v(G, a.wq.bias, 3 * Dl)
        self.wo, self.g_wo = v(S, a.wo.weight, d, Dl), v(G, a.wo.weight, d, Dl)
        self.bo, self.g_bo = v(P, a.wo.bias, d), v(G, a.wo.bias, d)
        self.wgu, self.g_wgu = v(S, f.gate_proj.weight, 2 * Fl, d), v(G, f.gate_proj.weight, 2 * Fl, d)
        self.bgu, self.g_bgu = v(P, f.gate_proj.bias, 2 * Fl), v(G, f.gate_proj.bias, 2 * Fl)
        self.wd, self.g_wd = v(S, f.down_proj.weight, d, Fl), v(G, f.down_proj.weight, d, Fl)
        self.bd, self.g_bd = v(P, f.down_proj.bias, d), v(G, f.down_proj.bias, d)
        self.s1, self.g_s1 = v(P, layer.norm1.scale, d), v(G, layer.norm1.scale, d)
        self.s2, self.g_s2 = v(P, layer.norm2.scale, d), v(G, layer.norm2.scale, d)
        self.params = [a.wq.weight, a.wk.weight, a.wv.weight, a.wq.bias, a.wk.bias, a.wv.bias, a.wo.weight,
                       a.wo.bias, f.gate_proj.weight, f.up_proj.weight, f.gate_proj.bias, f.up_proj.bias,
                       f.down_proj.weight, f.down_proj.bias, layer.norm1.scale, layer.norm2.scale]


class Meta:
    """Static shape info of one forward call."""

    def __init__(self, B, T, d, Hl, hd, Fl, n):
        self.B, self.T, self.M = B, T, B * T
        self.d, self.Hl, self.hd, self.Dl, self.Fl, self.n = d, Hl, hd, Hl * hd, Fl, n
        self.scale = 1.0 / math.sqrt(hd)


# ================================================================================================ building blocks
def _rowpar_residual(a_bf16, w_bf16, bias, residual, norm_scale=None, eps=1e-5):
    """x_out = residual + all_reduce(a @ W^T) + bias  (row-parallel linear folded into the residual stream) and,
    when ``norm_scale`` is given, h = rmsnorm(x_out) for the next sub-block.  Returns (x_out, h | None, rstd | None).

    TP=1: residual + bias ride in the GEMM epilogue.  TP>1: ONE kernel does GEMM + all-reduce over NVLink (tiles are
    summed through the NVSwitch as they complete), then residual + bias + norm are one fused pass."""
    n = pm.pgm.tp_size
    if n == 1:
        if norm_scale is None:
            # residual + bias ride in the GEMM epilogue
            return gemm_nt(a_bf16, w_bf16, bias=bias, residual=residual, out_dtype=F32), None, None
        # a norm follows: bf16 GEMM output through the TMA-store epilogue, then residual + bias + norm in ONE pass
        # (measured 25 + 50 us vs 78 us residual-epilogue GEMM + 23 us norm for the wo projection)
        y = gemm_nt(a_bf16, w_bf16, out_dtype=BF16)
        return fused.add_rmsnorm_fwd(residual, y, bias, norm_scale, eps)
    symm = pm.pgm.symm
    if symm is not None and symm.can_fuse_gemm_ar(a_bf16, w_bf16):
        y = symm.gemm_allreduce(a_bf16, w_bf16)
    else:
        y = gemm_nt(a_bf16, w_bf16, out_dtype=BF16)
        comm_ops.all_reduce_sum_(y)
    return fused.add_rmsnorm_fwd(residual, y, bias, norm_scale, eps, want_norm=norm_scale is not None)


def _colpar_dgrad(dy_bf16, w_bf16):
    """dx = all_reduce(dy @ W)  (column-parallel backward, one collective for the whole fused group)."""
    n = pm.pgm.tp_size
    if n == 1:
        return gemm_nt(dy_bf16, w_bf16, b_mn_major=True, out_dtype=BF16)
    symm = pm.pgm.symm
    if symm is not None and symm.can_fuse_gemm_ar(dy_bf16, w_bf16, b_mn_major=True):
        return symm.gemm_allreduce(dy_bf16, w_bf16, b_mn_major=True)
    dx = gemm_nt(dy_bf16, w_bf16, b_mn_major=True, out_dtype=BF16)
    comm_ops.all_reduce_sum_(dx)
    return dx


def _wgrad(dy_bf16, x_bf16, out):
    gemm_nt(dy_bf16, x_bf16, a_mn_major=True, b_mn_major=True, out=out)


class FusedLayerFn(torch.autograd.Function):
    @staticmethod
    @_nvtx("layer{i}.fwd")
    def forward(ctx, x, pos, rt, pack: LayerPack, meta: Meta, *params):
        m = meta
        h1, rstd1 = fused.rmsnorm_fwd(x, pack.s1, rt.eps)
        qkv = gemm_nt(h1, pack.wqkv, bias=pack.bqkv, out_dtype=BF16)
        fused.rope_inplace(qkv, pos, rt.cos_half, rt.sin_half, 2, m.Dl, m.Hl, m.hd, inverse=False)
        attn, lse = attn_fwd(qkv, m.B, m.T, m.Hl, m.hd, m.scale)
        x2, h2, rstd2 = _rowpar_residual(attn, pack.wo, pack.bo, x, pack.s2, rt.eps)
        gu, act = gemm_swiglu(h2, pack.wgu, pack.bgu)      # SwiGLU in the GEMM epilogue when the shape allows
        x3, _, _ = _rowpar_residual(act, pack.wd, pack.bd, x2)
        ctx.save_for_backward(x, pos, h1, rstd1, qkv, attn, lse, x2, h2, rstd2, gu, act)
        ctx.pack, ctx.meta, ctx.rt = pack, meta, rt
        return x3

    @staticmethod
    @_nvtx("layer{i}.bwd")
    def backward(ctx, dx3):
        x, pos, h1, rstd1, qkv, attn, lse, x2, h2, rstd2, gu, act = ctx.saved_tensors
        pack, m, rt = ctx.pack, ctx.meta, ctx.rt
        dx3 = dx3.contiguous()
        dx3_b = rt.take_stash(dx3)
        if dx3_b is None:
            # dx3 did not come from the fused RMSNorm backward above us (e.g. the module-path head): cast it and take
            # the down_proj bias gradient with a pass of its own
            dx3_b = fused.cast_to_bf16(dx3)
            fused.colsum(dx3_b, out=pack.g_bd)
        # ---- FFN
        _wgrad(dx3_b, act, pack.g_wd)
        # down_proj dgrad with the SwiGLU backward (+ gate|up bias grads) in its epilogue: dact never reaches memory
        dgu = gemm_swiglu_bwd(dx3_b, pack.wd, gu, pack.g_bgu)
        dh2 = _colpar_dgrad(dgu, pack.wgu)                                            # [M, d]
        _wgrad(dgu, h2, pack.g_wgu)
        del dgu
        # dx2 is the output gradient of wo's residual add: its column sums are the wo bias gradient (same pass)
        dx2, dx2_b, _ = fused.rmsnorm_bwd(dh2, x2, rstd2, pack.s2, dres=dx3, want_fp32=True, want_bf16=True,
                                          dscale_out=pack.g_s2, dbias_out=pack.g_bo)
        # ---- attention
        dattn = gemm_nt(dx2_b, pack.wo, b_mn_major=True, out_dtype=BF16)              # [M, Dl]
        _wgrad(dx2_b, attn, pack.g_wo)
        dqkv = attn_bwd(qkv, attn, dattn, lse, m.B, m.T, m.Hl, m.hd, m.scale)
        fused.rope_inplace(dqkv, pos, rt.cos_half, rt.sin_half, 2, m.Dl, m.Hl, m.hd, inverse=True)
        dh1 = _colpar_dgrad(dqkv, pack.wqkv)
        _wgrad(dqkv, h1, pack.g_wqkv)
        fused.colsum(dqkv, out=pack.g_bqkv)
        if pack.prev is not None:
            # dx is the output gradient of the layer below: hand it its bf16 copy and its down_proj bias gradient
            dx, dx_b, _ = fused.rmsnorm_bwd(dh1, x, rstd1, pack.s1, dres=dx2, want_fp32=True, want_bf16=True,
                                            dscale_out=pack.g_s1, dbias_out=pack.prev.g_bd)
            rt.put_stash(dx, dx_b)
        else:
            dx, _, _ = fused.rmsnorm_bwd(dh1, x, rstd1, pack.s1, dres=dx2, want_fp32=True, want_bf16=False,
                                         dscale_out=pack.g_s1)
        return (dx, None, None, None, None) + (None,) * len(pack.params)


class FusedEmbedFn(torch.autograd.Function):
    @staticmethod
    @_nvtx("embed.fwd")
    def forward(ctx, ids_flat, rt, weight):
        emb = rt.model.embedding
        if pm.pgm.tp_size == 1:
            out = fused.embedding_fwd(ids_flat, weight.detach(), emb.vocab_st_idx, emb.vocab_ed_idx, out_dtype=F32)
        else:
            # exactly one rank owns each row, so summing bf16 rows is exact; the reference also casts the embedding
            # output to bf16 before the first layer (models/model.py:153-154)
            symm = pm.pgm.symm
            D = weight.shape[1]
            if symm is not None and symm.push_enabled and ids_flat.numel() * D * 2 <= symm.push_zone_bytes \
                    and (ids_flat.numel() * D * 2) % 16 == 0:
                # opt-in (DPFS_FUSE_AG=1): owner-scatter through the multicast mapping, no reduction at all
                out = fused.embedding_fwd_scatter(ids_flat, weight.detach(), emb.vocab_st_idx, emb.vocab_ed_idx,
                                                  symm).float()
            elif symm is not None and ids_flat.numel() * D * 2 <= symm.ag_off:
                part = symm.buffer((ids_flat.numel(), D), BF16)
                fused.embedding_fwd(ids_flat, weight.detach(), emb.vocab_st_idx, emb.vocab_ed_idx, out=part)
                out = symm.all_reduce(part, out=part).float()
            else:
                part = fused.embedding_fwd(ids_flat, weight.detach(), emb.vocab_st_idx, emb.vocab_ed_idx,
                                           out_dtype=BF16)
                out = comm_ops.all_reduce_sum_(part).float()
        ctx.save_for_backward(ids_flat)
        ctx.rt = rt
        return out

    @staticmethod
    @_nvtx("embed.bwd")
    def backward(ctx, dx):
        (ids_flat,) = ctx.saved_tensors
        rt = ctx.rt
        emb = rt.model.embedding
        fused.embedding_bwd(ids_flat, dx.contiguous(), emb.weight.shape[0], emb.vocab_st_idx, emb.vocab_ed_idx,
                            dweight=rt.g_emb, zero_first=True)
        return None, None, None


class FusedHeadLossFn(torch.autograd.Function):
    """final RMSNorm -> vocab-parallel lm_head GEMM -> vocab-parallel softmax cross-entropy (mean over valid)."""

    @staticmethod
    @_nvtx("head.fwd")
    def forward(ctx, x, targets_flat, rt, ignore_index, *params):
        n = pm.pgm.tp_size
        head = rt.model.lm_head
        Vl = head.odim_partition
        vocab_start = pm.pgm.tp_rank * Vl
        h, rstd = fused.rmsnorm_fwd(x, rt.sN, rt.eps)
        logits = gemm_nt(h, rt.wlm, bias=rt.blm, out_dtype=F32)
        stats, loss_acc = fused.ce_stats(logits, targets_flat, vocab_start, ignore_index)
        if n > 1:
            symm = pm.pgm.symm
            all_stats = symm.all_gather_rows(stats) if symm is not None else comm_ops.all_gather_rows(stats)
        else:
            all_stats = stats
        fused.ce_loss(targets_flat, all_stats, n, loss_acc, ignore_index)
        loss = loss_acc[0] / loss_acc[1]
        ctx.save_for_backward(x, targets_flat, h, rstd, logits, all_stats, loss_acc)
        ctx.rt, ctx.ignore_index, ctx.vocab_start = rt, ignore_index, vocab_start
        return loss

    @staticmethod
    @_nvtx("head.bwd")
    def backward(ctx, gout):
        x, targets_flat, h, rstd, logits, all_stats, loss_acc = ctx.saved_tensors
        rt = ctx.rt
        n = pm.pgm.tp_size
        up = gout.detach().reshape(1).float().contiguous()
        dlogits = fused.ce_grad(logits, targets_flat, all_stats, n, loss_acc, ctx.vocab_start, ctx.ignore_index,
                                upstream=up)
        dh = _colpar_dgrad(dlogits, rt.wlm)
        _wgrad(dlogits, h, rt.g_wlm)
        fused.colsum(dlogits, out=rt.g_blm)
        dx, dx_b, _ = fused.rmsnorm_bwd(dh, x, rstd, rt.sN, dres=None, want_fp32=True, want_bf16=True,
                                        dscale_out=rt.g_sN, dbias_out=rt.packs[-1].g_bd)
        rt.put_stash(dx, dx_b)
        return (dx, None, None, None) + (None,) * 3


# ================================================================================================ runtime
class FusedRuntime:
    def __init__(self, model):
        self.model = model
        n = pm.pgm.tp_size
        self.n = n
        self.eps = model.norm.eps
        d, hd = model.attn_dim, model.head_dim
        self.d, self.hd = d, hd
        self.Hl = model.num_heads // n
        self.Dl = self.Hl * hd
        self.Fl = model.ffn_dim // n
        self.arena = ParamArena(model)
        A = self.arena
        self.packs: List[LayerPack] = [LayerPack(A, layer, d, self.Dl, self.Fl) for layer in model.layers]
        for i, pack in enumerate(self.packs):
            pack.prev = self.packs[i - 1] if i > 0 else None
            pack.index = i
        self._stash = None
        self._fwd_handoff = None      # (x_l, gathered normalised input of the next block, its rstd)
        self._zones = None            # symmetric activation zones of the hand-off path (allocated on first use)
        self._zone_rows = 0
        self._bz, self._bz_events = 0, [None, None]
        self._side, self._side_keep, self._side_dirty = None, [], False
        head, emb = model.lm_head, model.embedding
        Vl = head.odim_partition
        self.wlm, self.g_wlm = A.view(A.shadow, head.weight, Vl, d), A.view(A.grads, head.weight, Vl, d)
        self.blm, self.g_blm = A.view(A.params, head.bias, Vl), A.view(A.grads, head.bias, Vl)
        self.sN, self.g_sN = A.view(A.params, model.norm.scale, d), A.view(A.grads, model.norm.scale, d)
        self.g_emb = A.view(A.grads, emb.weight, emb.weight.shape[0], d)
        self.head_params = [model.norm.scale, head.weight, head.bias]
        half = hd // 2
        self.cos_half = model.rope_cos_f32[:, :half].contiguous().to(A.device)
        self.sin_half = model.rope_sin_f32[:, :half].contiguous().to(A.device)

    def valid(self) -> bool:
        return self.arena.valid() and self.n == pm.pgm.tp_size

    # The fused RMSNorm backward of one block produces, in the same pass, the bf16 copy of the residual-stream gradient
    # and the down_proj bias gradient that the block below needs.  They travel outside autograd (which only carries the
    # fp32 gradient); the consumer checks that the gradient it received is the very tensor the stash was made for.
    def put_stash(self, dx: torch.Tensor, dx_bf16: torch.Tensor) -> None:
        self._stash = (dx, dx_bf16)

    def take_stash(self, dx: torch.Tensor) -> Optional[torch.Tensor]:
        st, self._stash = self._stash, None
        if st is not None and st[0].data_ptr() == dx.data_ptr() and st[0].shape == dx.shape:
            return st[1]
        return None

    # ---------------------------------------------------------------------------------------------- forward
    def forward_hidden(self, input_ids: torch.Tensor, position_ids: torch.Tensor) -> torch.Tensor:
        assert input_ids.dim() == 2
        B, T = input_ids.shape
        self.arena.sync_shadow()
        self._stash = None
        if torch.is_grad_enabled():
            self.arena.reattach_grads()
        meta = Meta(B, T, self.d, self.Hl, self.hd, self.Fl, self.n)
        ids = input_ids.reshape(-1).contiguous()
        pos = position_ids.reshape(-1).contiguous()
        if self.sp_active(B * T):
            if self.handoff_ready(B * T, B):
                # ONE kernel per residual-stream hand-off (reduce-scatter -> norm -> all-gather through the switch)
                x = HandoffEmbedFn.apply(ids, self, self.model.embedding.weight)
                for pack in self.packs:
                    x = HandoffLayerFn.apply(x, pos, self, pack, meta, *pack.params)
                return x
            x = FusedEmbedSPFn.apply(ids, self, self.model.embedding.weight)
            for pack in self.packs:
                x = FusedLayerSPFn.apply(x, pos, self, pack, meta, *pack.params)
            return x  # fp32 [M / n, d]: this rank's rows
        x = FusedEmbedFn.apply(ids, self, self.model.embedding.weight)
        for pack in self.packs:
            x = FusedLayerFn.apply(x, pos, self, pack, meta, *pack.params)
        return x  # fp32 [M, d]

    # ---------------------------------------------------------------------------------------------- SP hand-off path
    def handoff_ready(self, M: int, B: Optional[int] = None) -> bool:
        """True when the one-kernel hand-offs (csrc/sp_fused.cu) can run this batch: NVLS multicast available, model
        width within the kernels' register budget, partial matrix fits the symmetric buffer.  Allocates the symmetric
        activation zones on first use (a collective: every rank takes this branch for the same batch shapes)."""
        symm = pm.pgm.symm
        if symm is None or not symm.mc_base or symm.closed or os.environ.get("DPFS_SP_HANDOFF", "1") != "1":
            return False
        if self.d % 8 or self.d > 1024 or M * self.d * 2 > symm.ag_off:
            return False
        if self._zones is None or M > self._zone_rows:
            # h1 / h2 of every layer and the final norm's output stay resident for the backward pass (2L + 1 zones);
            # two more zones alternate as the landing buffer of the gathered bf16 residual-stream gradient (two, so a
            # weight-gradient GEMM on the side stream may still read one while the next hand-off fills the other)
            # Sized for the longest sequence the model accepts at this batch size, not for this batch: the loader's
            # batches differ in length, and a zone set that had to grow would be a second ~GB allocation (the old one
            # stays alive for the CUDA graphs captured on it).
            rows = max(M, (B or 0) * int(getattr(self.model, "maxlen", 0) or 0))
            self._zones = symm.alloc_zones(2 * len(self.packs) + 3, rows * self.d * 2)
            self._zone_rows = rows
            self._bz = 0
        return True

    # Weight-gradient GEMMs never feed the rest of the backward pass; on the hand-off path they run on a side stream so
    # they fill the SMs that the NVLink-bound hand-off kernels leave idle (those run on a fraction of the SMs).
    #   * next_bzone(): the landing zone for the next gathered gradient; before handing a zone out again, the main
    #     stream waits for the side-stream work that read it two hand-offs ago.
    #   * on_side(fn, keep): run fn on the side stream after everything enqueued so far; ``keep`` are the tensors it
    #     reads, held until join_side() so the caching allocator cannot hand their memory to a later main-stream tensor.
    def side_enabled(self) -> bool:
        return os.environ.get("DPFS_SP_SIDE_WGRAD", "1") == "1"

    def next_bzone(self):
        self._bz ^= 1
        ev = self._bz_events[self._bz]
        if ev is not None:
            torch.cuda.current_stream().wait_event(ev)
            self._bz_events[self._bz] = None
        return self._zones[-1 - self._bz]

    def on_side(self, fn, keep=(), reads_bzone: bool = False) -> None:
        if not self.side_enabled():
            fn()
            return
        if self._side is None:
            self._side = torch.cuda.Stream()
        main = torch.cuda.current_stream()
        self._side.wait_stream(main)
        with torch.cuda.stream(self._side):
            fn()
            if reads_bzone:
                ev = torch.cuda.Event()
                ev.record(self._side)
                self._bz_events[self._bz] = ev
        self._side_keep.extend(keep)
        self._side_dirty = True

    def join_side(self) -> None:
        if self._side_dirty:
            torch.cuda.current_stream().wait_stream(self._side)
            self._side_dirty = False
        self._side_keep.clear()
        self._bz_events = [None, None]

    def put_handoff(self, x_l, h, rstd) -> None:
        self._fwd_handoff = (x_l, h, rstd)

    def take_handoff(self, x_l):
        st, self._fwd_handoff = self._fwd_handoff, None
        assert st is not None and st[0].data_ptr() == x_l.data_ptr(), "hand-off state does not belong to this input"
        return st[1], st[2]

    def sp_active(self, M: int) -> bool:
        """Sequence parallelism is on when the model asks for it, TP > 1 and the rows split evenly (16 B units)."""
        return bool(getattr(self.model, "sequence_parallel", False)) and self.n > 1 and M % self.n == 0 \
            and ((M // self.n) * self.d * 2) % 16 == 0

    def forward_loss(self, input_ids, position_ids, target_ids, ignore_index: int = -1) -> torch.Tensor:
        x = self.forward_hidden(input_ids, position_ids)
        tg = target_ids.reshape(-1).contiguous()
        if self.sp_active(tg.numel()):
            if self._fwd_handoff is not None:
                return HandoffHeadLossFn.apply(x, tg, self, ignore_index, *self.head_params)
            return FusedHeadLossSPFn.apply(x, tg, self, ignore_index, *self.head_params)
        return FusedHeadLossFn.apply(x, tg, self, ignore_index, *self.head_params)

    def forward_logits(self, input_ids, position_ids) -> torch.Tensor:
        B, T = input_ids.shape
        x = self.forward_hidden(input_ids, position_ids)
        self._fwd_handoff = None
        if self.sp_active(B * T):
            x = _all_gather_rows(x, keep=True)      # gather the rows back for the (replicated) logits path
        x = x.view(B, T, self.d)
        with torch.autocast("cuda", dtype=BF16):
            logits = self.model.lm_head(self.model.norm(x))
        return logits


# ================================================================================================ sequence parallel
# Megatron-style sequence parallelism on top of TP (north-star config 4; the reference has none):
# the residual stream, the norms and every elementwise op work on this rank's M/n rows; each all-reduce becomes
#     row-parallel GEMM  -> reduce-scatter over rows        (forward of wo / down, backward of q|k|v / gate|up / lm_head)
#     all-gather of rows -> column-parallel GEMM            (forward of q|k|v / gate|up / lm_head, backward of wo / down)
# Same bytes on the wire, 1/n of the norm / residual work and of the fp32 residual-stream memory.  RMSNorm scale
# gradients become partial sums over local rows and are all-reduced once per step (ParamArena.norm_grads).
def _reduce_scatter_rows(part_bf16):
    """Sum the ranks' partial [M, N] matrices and keep this rank's block of rows -> [M / n, N] bf16."""
    symm = pm.pgm.symm
    if symm is not None and symm._inside(part_bf16) and (symm.mode == "nvls" or symm.n in (2, 4, 8)):
        return symm.reduce_scatter_rows(part_bf16)
    import torch.distributed as dist
    M, N = part_bf16.shape
    out = torch.empty((M // pm.pgm.tp_size, N), dtype=part_bf16.dtype, device=part_bf16.device)
    dist.reduce_scatter_tensor(out, part_bf16.contiguous(), group=pm.pgm.tp_group)
    return out


def _all_gather_rows(local, keep: bool = False):
    """[Ml, ...] -> [n * Ml, ...] in rank order.  ``keep``: return a private copy (saved for backward)."""
    symm = pm.pgm.symm
    if symm is not None and local.numel() * local.element_size() * symm.n <= symm.ag_bytes:
        full = symm.all_gather_into_buffer(local)
        return full.clone() if keep else full
    import torch.distributed as dist
    local = local.contiguous()
    out = torch.empty((pm.pgm.tp_size * local.shape[0],) + tuple(local.shape[1:]), dtype=local.dtype,
                      device=local.device)
    dist.all_gather_into_tensor(out, local, group=pm.pgm.tp_group)
    return out


def _push_comm(x_l):
    """The symmetric communicator when pushed all-gathers are enabled (DPFS_FUSE_AG=1 + NVLS multicast) and a block of
    x_l's shape fits a push zone; None -> the split path (local block + all-gather kernel)."""
    symm = pm.pgm.symm
    if symm is not None and x_l.dim() == 2 and symm.can_push(x_l.shape[0], x_l.shape[1] * 2):
        return symm
    return None


def _partial_out(M, N):
    """Where a row-parallel GEMM writes its partial sums: straight into the symmetric buffer when there is one."""
    symm = pm.pgm.symm
    if symm is not None and M * N * 2 <= symm.ag_off:
        return symm.buffer((M, N), BF16)
    return torch.empty((M, N), dtype=BF16, device=torch.device("cuda", torch.cuda.current_device()))


# GEMM -> reduce-scatter as ONE kernel (csrc/gemm_tcgen05.cu, rs mode).  Opt-in: the split path (2-SM GEMM + NVLS
# reduce-scatter kernel) is the measured default until the fused kernel has been timed against it at TP=8.
_FUSE_RS = os.environ.get("DPFS_FUSE_RS", "0") == "1"


def _rowpar_rs(a_bf16, w_bf16, b_mn_major=False):
    """reduce_scatter_rows(a @ W^T)  (or a @ W when b_mn_major)."""
    M = a_bf16.shape[0]
    N = w_bf16.shape[1] if b_mn_major else w_bf16.shape[0]
    symm = pm.pgm.symm
    if _FUSE_RS and symm is not None and symm.can_fuse_gemm_rs(a_bf16, w_bf16, b_mn_major):
        return symm.gemm_reduce_scatter(a_bf16, w_bf16, b_mn_major=b_mn_major)
    part = _partial_out(M, N)
    gemm_nt(a_bf16, w_bf16, b_mn_major=b_mn_major, out=part)
    return _reduce_scatter_rows(part)


class FusedLayerSPFn(torch.autograd.Function):
    @staticmethod
    @_nvtx("layer{i}.sp.fwd")
    def forward(ctx, x_l, pos, rt, pack: LayerPack, meta: Meta, *params):
        m = meta
        push = _push_comm(x_l)
        if push is not None:   # the norm kernel all-gathers h as it writes it (multicast stores)
            h1, rstd1 = fused.rmsnorm_fwd_push(x_l, pack.s1, rt.eps, push)
            h1 = h1.clone()    # saved for the backward wgrad; the landing zone is recycled two pushes later
        else:
            h1_l, rstd1 = fused.rmsnorm_fwd(x_l, pack.s1, rt.eps)
            h1 = _all_gather_rows(h1_l, keep=True)
        qkv = gemm_nt(h1, pack.wqkv, bias=pack.bqkv, out_dtype=BF16)
        fused.rope_inplace(qkv, pos, rt.cos_half, rt.sin_half, 2, m.Dl, m.Hl, m.hd, inverse=False)
        attn, lse = attn_fwd(qkv, m.B, m.T, m.Hl, m.hd, m.scale)
        y_l = _rowpar_rs(attn, pack.wo)
        if push is not None:
            x2_l, h2, rstd2 = fused.add_rmsnorm_fwd_push(x_l, y_l, pack.bo, pack.s2, rt.eps, push)
            h2 = h2.clone()
        else:
            x2_l, h2_l, rstd2 = fused.add_rmsnorm_fwd(x_l, y_l, pack.bo, pack.s2, rt.eps)
            h2 = _all_gather_rows(h2_l, keep=True)
        gu, act = gemm_swiglu(h2, pack.wgu, pack.bgu)      # SwiGLU in the GEMM epilogue when the shape allows
        y_l = _rowpar_rs(act, pack.wd)
        x3_l, _, _ = fused.add_rmsnorm_fwd(x2_l, y_l, pack.bd, None, rt.eps, want_norm=False)
        ctx.save_for_backward(x_l, pos, h1, rstd1, qkv, attn, lse, x2_l, h2, rstd2, gu, act)
        ctx.pack, ctx.meta, ctx.rt = pack, meta, rt
        return x3_l

    @staticmethod
    @_nvtx("layer{i}.sp.bwd")
    def backward(ctx, dx3_l):
        x_l, pos, h1, rstd1, qkv, attn, lse, x2_l, h2, rstd2, gu, act = ctx.saved_tensors
        pack, m, rt = ctx.pack, ctx.meta, ctx.rt
        dx3_l = dx3_l.contiguous()
        push = _push_comm(dx3_l)
        if push is not None:
            dx3 = fused.cast_to_bf16_push(dx3_l, push)                                   # [M, d] bf16 (landing zone)
        else:
            dx3 = _all_gather_rows(fused.cast_to_bf16(dx3_l))                            # [M, d] bf16
        _wgrad(dx3, act, pack.g_wd)
        Ml, r = dx3_l.shape[0], pm.pgm.tp_rank
        fused.colsum(dx3[r * Ml:(r + 1) * Ml], out=pack.g_bd)    # partial over this rank's rows (see norm_grads)
        dgu = gemm_swiglu_bwd(dx3, pack.wd, gu, pack.g_bgu)
        dh2_l = _rowpar_rs(dgu, pack.wgu, b_mn_major=True)                               # [Ml, d]
        _wgrad(dgu, h2, pack.g_wgu)
        del dgu
        if push is not None:
            dx2_l, dx2 = fused.rmsnorm_bwd_push(dh2_l, x2_l, rstd2, pack.s2, dx3_l, push, dscale_out=pack.g_s2)
        else:
            dx2_l, dx2_lb, _ = fused.rmsnorm_bwd(dh2_l, x2_l, rstd2, pack.s2, dres=dx3_l, want_fp32=True,
                                                 want_bf16=True, dscale_out=pack.g_s2)
            dx2 = _all_gather_rows(dx2_lb)
        dattn = gemm_nt(dx2, pack.wo, b_mn_major=True, out_dtype=BF16)
        _wgrad(dx2, attn, pack.g_wo)
        fused.colsum(dx2[r * Ml:(r + 1) * Ml], out=pack.g_bo)
        dqkv = attn_bwd(qkv, attn, dattn, lse, m.B, m.T, m.Hl, m.hd, m.scale)
        fused.rope_inplace(dqkv, pos, rt.cos_half, rt.sin_half, 2, m.Dl, m.Hl, m.hd, inverse=True)
        dh1_l = _rowpar_rs(dqkv, pack.wqkv, b_mn_major=True)
        _wgrad(dqkv, h1, pack.g_wqkv)
        fused.colsum(dqkv, out=pack.g_bqkv)
        dx_l, _, _ = fused.rmsnorm_bwd(dh1_l, x_l, rstd1, pack.s1, dres=dx2_l, want_fp32=True, want_bf16=False,
                                       dscale_out=pack.g_s1)
        return (dx_l, None, None, None, None) + (None,) * len(pack.params)


class FusedEmbedSPFn(torch.autograd.Function):
    @staticmethod
    @_nvtx("embed.sp.fwd")
    def forward(ctx, ids_flat, rt, weight):
        emb = rt.model.embedding
        M, D = ids_flat.numel(), weight.shape[1]
        part = _partial_out(M, D)
        fused.embedding_fwd(ids_flat, weight.detach(), emb.vocab_st_idx, emb.vocab_ed_idx, out=part)
        x_l = _reduce_scatter_rows(part).float()       # exact: one owner per row
        ctx.save_for_backward(ids_flat)
        ctx.rt = rt
        return x_l

    @staticmethod
    @_nvtx("embed.sp.bwd")
    def backward(ctx, dx_l):
        (ids_flat,) = ctx.saved_tensors
        rt = ctx.rt
        emb = rt.model.embedding
        dx = _all_gather_rows(dx_l.contiguous())       # fp32 [M, d]
        fused.embedding_bwd(ids_flat, dx, emb.weight.shape[0], emb.vocab_st_idx, emb.vocab_ed_idx, dweight=rt.g_emb,
                            zero_first=True)
        # last node of the backward pass: complete the RMSNorm scale gradients (partial sums over local rows)
        comm_ops.all_reduce_sum_(rt.arena.norm_grads)
        return None, None, None


class FusedHeadLossSPFn(torch.autograd.Function):
    @staticmethod
    @_nvtx("head.sp.fwd")
    def forward(ctx, x_l, targets_flat, rt, ignore_index, *params):
        n = pm.pgm.tp_size
        head = rt.model.lm_head
        Vl = head.odim_partition
        vocab_start = pm.pgm.tp_rank * Vl
        h_l, rstd = fused.rmsnorm_fwd(x_l, rt.sN, rt.eps)
        h = _all_gather_rows(h_l, keep=True)
        logits = gemm_nt(h, rt.wlm, bias=rt.blm, out_dtype=F32)
        stats, loss_acc = fused.ce_stats(logits, targets_flat, vocab_start, ignore_index)
        symm = pm.pgm.symm
        all_stats = symm.all_gather_rows(stats) if symm is not None else comm_ops.all_gather_rows(stats)
        fused.ce_loss(targets_flat, all_stats, n, loss_acc, ignore_index)
        loss = loss_acc[0] / loss_acc[1]
        ctx.save_for_backward(x_l, targets_flat, h, rstd, logits, all_stats, loss_acc)
        ctx.rt, ctx.ignore_index, ctx.vocab_start = rt, ignore_index, vocab_start
        return loss

    @staticmethod
    @_nvtx("head.sp.bwd")
    def backward(ctx, gout):
        x_l, targets_flat, h, rstd, logits, all_stats, loss_acc = ctx.saved_tensors
        rt = ctx.rt
        n = pm.pgm.tp_size
        up = gout.detach().reshape(1).float().contiguous()
        dlogits = fused.ce_grad(logits, targets_flat, all_stats, n, loss_acc, ctx.vocab_start, ctx.ignore_index,
                                upstream=up)
        dh_l = _rowpar_rs(dlogits, rt.wlm, b_mn_major=True)
        _wgrad(dlogits, h, rt.g_wlm)
        fused.colsum(dlogits, out=rt.g_blm)
        dx_l, _, _ = fused.rmsnorm_bwd(dh_l, x_l, rstd, rt.sN, dres=None, want_fp32=True, dscale_out=rt.g_sN)
        return (dx_l, None, None, None) + (None,) * 3


# ================================================================================================ SP hand-off path
# The same sequence-parallel schedule with every residual-stream hand-off as ONE kernel (csrc/sp_fused.cu):
#     partial = row-parallel GEMM (symmetric buffer)   ->   rs_add_rmsnorm_ag / rs_rmsnorm_bwd_ag
# i.e. reduce-scatter, residual + bias + RMSNorm (or RMSNorm backward) on this rank's rows and the all-gather of the
# bf16 result run as one pass through the NVSwitch (multimem.ld_reduce in, multimem.st out, both link directions busy).
# The norm that OPENS a block is therefore computed by the kernel that CLOSES the previous one: autograd carries only
# the fp32 local rows x_l between the Functions; the gathered activations travel through the runtime (put_handoff /
# take_handoff forward, put_stash / take_stash backward) and are checked against the tensor they belong to.
# GEMM + hand-off as ONE kernel (csrc/gemm_handoff.cu: tile flags instead of the opening barrier, the reduction of the
# first row-blocks under the GEMM of the later ones) or as 2-SM GEMM + stand-alone hand-off kernel.
def _fused_gemm_handoff() -> bool:
    """GEMM + hand-off as one kernel?  DPFS_SP_FUSED_GEMM=0/1 forces the choice; by default the measured winner per TP
    size (profiles/bench_tp_r2.md): one kernel at TP=2 (13.83 vs 14.12 ms / step), 2-SM GEMM + stand-alone hand-off kernel at
    TP=4 (10.53 vs 12.62 ms) and TP=8 (9.06 vs 10.43 ms) -- the fused kernel's GEMM is the 1-SM one and shares its SMs with
    the comm warps, which costs more than the overlap returns once the per-rank GEMMs are short."""
    env = os.environ.get("DPFS_SP_FUSED_GEMM")
    if env is not None:
        return env == "1"
    return pm.pgm.tp_size <= 2


def _handoff_fwd(a, w, x_l, bias, scale, eps, zone):
    """(x_out_l, h gathered, rstd) of x_out_l = x_l + reduce_scatter(a @ w^T) + bias, h = rmsnorm(x_out_l) * scale."""
    symm = pm.pgm.symm
    if _fused_gemm_handoff() and fused.can_gemm_handoff(symm, a, w):
        return fused.gemm_handoff_fwd(symm, a, w, False, x_l, bias, scale, eps, zone)
    part = symm.buffer((a.shape[0], w.shape[0]), BF16)
    gemm_nt(a, w, out=part)
    return fused.rs_add_rmsnorm_ag(symm, part, x_l, bias, scale, eps, zone)


def _handoff_bwd(dy, w, x_l, rstd, scale, dres_l, zone, dscale_out, dbias_out):
    """(dx_l, dx gathered) of dx_l = dres_l + rmsnorm'(reduce_scatter(dy @ w)) (column-parallel dgrad + hand-off)."""
    symm = pm.pgm.symm
    if _fused_gemm_handoff() and fused.can_gemm_handoff(symm, dy, w, True):
        return fused.gemm_handoff_bwd(symm, dy, w, True, x_l, rstd, scale, dres_l, zone, dscale_out, dbias_out)
    part = symm.buffer((dy.shape[0], w.shape[1]), BF16)
    gemm_nt(dy, w, b_mn_major=True, out=part)
    return fused.rs_rmsnorm_bwd_ag(symm, part, x_l, rstd, scale, dres_l, zone, dscale_out, dbias_out)


class HandoffEmbedFn(torch.autograd.Function):
    @staticmethod
    @_nvtx("embed.handoff.fwd")
    def forward(ctx, ids_flat, rt, weight):
        emb = rt.model.embedding
        symm = pm.pgm.symm
        M, D = ids_flat.numel(), weight.shape[1]
        part = symm.buffer((M, D), BF16)
        fused.embedding_fwd(ids_flat, weight.detach(), emb.vocab_st_idx, emb.vocab_ed_idx, out=part)
        # exact: one owner per row, the others contribute zeros; the first layer's norm rides in the same kernel
        x_l, h1, rstd1 = fused.rs_add_rmsnorm_ag(symm, part, None, None, rt.packs[0].s1, rt.eps, rt._zones[0])
        rt.put_handoff(x_l, h1, rstd1)
        ctx.save_for_backward(ids_flat)
        ctx.rt = rt
        return x_l

    @staticmethod
    @_nvtx("embed.handoff.bwd")
    def backward(ctx, dx_l):
        (ids_flat,) = ctx.saved_tensors
        rt = ctx.rt
        emb = rt.model.embedding
        dx = rt.take_stash(dx_l)          # gathered bf16 [M, d], produced by layer 0's closing hand-off
        assert dx is not None
        fused.embedding_bwd(ids_flat, dx, emb.weight.shape[0], emb.vocab_st_idx, emb.vocab_ed_idx, dweight=rt.g_emb,
                            zero_first=True)
        # last node of the backward pass: complete the gradients of the replicated parameters (partial sums over rows)
        comm_ops.all_reduce_sum_(rt.arena.norm_grads)
        rt.join_side()                    # the weight gradients computed on the side stream are complete from here on
        return None, None, None


class HandoffLayerFn(torch.autograd.Function):
    @staticmethod
    @_nvtx("layer{i}.handoff.fwd")
    def forward(ctx, x_l, pos, rt, pack: LayerPack, meta: Meta, *params):
        m = meta
        symm = pm.pgm.symm
        li = pack.index
        h1, rstd1 = rt.take_handoff(x_l)
        qkv = gemm_nt(h1, pack.wqkv, bias=pack.bqkv, out_dtype=BF16)
        fused.rope_inplace(qkv, pos, rt.cos_half, rt.sin_half, 2, m.Dl, m.Hl, m.hd, inverse=False)
        attn, lse = attn_fwd(qkv, m.B, m.T, m.Hl, m.hd, m.scale)
        x2_l, h2, rstd2 = _handoff_fwd(attn, pack.wo, x_l, pack.bo, pack.s2, rt.eps, rt._zones[2 * li + 1])
        gu, act = gemm_swiglu(h2, pack.wgu, pack.bgu)      # SwiGLU in the GEMM epilogue when the shape allows
        nxt = rt.packs[li + 1].s1 if li + 1 < len(rt.packs) else rt.sN      # the norm that opens the next block
        x3_l, hn, rstdn = _handoff_fwd(act, pack.wd, x2_l, pack.bd, nxt, rt.eps, rt._zones[2 * li + 2])
        rt.put_handoff(x3_l, hn, rstdn)
        ctx.save_for_backward(x_l, pos, h1, rstd1, qkv, attn, lse, x2_l, h2, rstd2, gu, act)
        ctx.pack, ctx.meta, ctx.rt = pack, meta, rt
        return x3_l

    @staticmethod
    @_nvtx("layer{i}.handoff.bwd")
    def backward(ctx, dx3_l):
        x_l, pos, h1, rstd1, qkv, attn, lse, x2_l, h2, rstd2, gu, act = ctx.saved_tensors
        pack, m, rt = ctx.pack, ctx.meta, ctx.rt
        symm = pm.pgm.symm
        dx3_l = dx3_l.contiguous()
        dx3 = rt.take_stash(dx3_l)        # gathered bf16 [M, d] from the hand-off that closed the block above
        assert dx3 is not None
        rt.on_side(lambda: _wgrad(dx3, act, pack.g_wd), keep=(dx3, act), reads_bzone=True)
        dgu = gemm_swiglu_bwd(dx3, pack.wd, gu, pack.g_bgu)    # dgrad of down_proj + SwiGLU backward in one kernel
        rt.on_side(lambda: _wgrad(dgu, h2, pack.g_wgu), keep=(dgu, h2))
        # dx2 is the output gradient of wo's residual add: its (local) column sums are the wo bias gradient
        dx2_l, dx2 = _handoff_bwd(dgu, pack.wgu, x2_l, rstd2, pack.s2, dx3_l, rt.next_bzone(), pack.g_s2, pack.g_bo)
        dattn = gemm_nt(dx2, pack.wo, b_mn_major=True, out_dtype=BF16)
        rt.on_side(lambda: _wgrad(dx2, attn, pack.g_wo), keep=(dx2, attn), reads_bzone=True)
        dqkv = attn_bwd(qkv, attn, dattn, lse, m.B, m.T, m.Hl, m.hd, m.scale)
        fused.rope_inplace(dqkv, pos, rt.cos_half, rt.sin_half, 2, m.Dl, m.Hl, m.hd, inverse=True)
        rt.on_side(lambda: _wgrad(dqkv, h1, pack.g_wqkv), keep=(dqkv, h1))
        fused.colsum(dqkv, out=pack.g_bqkv)
        # dx is the output gradient of the block below (its down_proj bias gradient comes from the same pass)
        prev_bd = pack.prev.g_bd if pack.prev is not None else None
        dx_l, dx = _handoff_bwd(dqkv, pack.wqkv, x_l, rstd1, pack.s1, dx2_l, rt.next_bzone(), pack.g_s1, prev_bd)
        rt.put_stash(dx_l, dx)
        return (dx_l, None, None, None, None) + (None,) * len(pack.params)


class HandoffHeadLossFn(torch.autograd.Function):
    @staticmethod
    @_nvtx("head.handoff.fwd")
    def forward(ctx, x_l, targets_flat, rt, ignore_index, *params):
        n = pm.pgm.tp_size
        head = rt.model.lm_head
        Vl = head.odim_partition
        vocab_start = pm.pgm.tp_rank * Vl
        h, rstd = rt.take_handoff(x_l)        # final RMSNorm, computed and gathered by the last layer's hand-off
        logits = gemm_nt(h, rt.wlm, bias=rt.blm, out_dtype=F32)
        stats, loss_acc = fused.ce_stats(logits, targets_flat, vocab_start, ignore_index)
        symm = pm.pgm.symm
        all_stats = symm.all_gather_rows(stats)
        fused.ce_loss(targets_flat, all_stats, n, loss_acc, ignore_index)
        loss = loss_acc[0] / loss_acc[1]
        ctx.save_for_backward(x_l, targets_flat, h, rstd, logits, all_stats, loss_acc)
        ctx.rt, ctx.ignore_index, ctx.vocab_start = rt, ignore_index, vocab_start
        return loss

    @staticmethod
    @_nvtx("head.handoff.bwd")
    def backward(ctx, gout):
        x_l, targets_flat, h, rstd, logits, all_stats, loss_acc = ctx.saved_tensors
        rt = ctx.rt
        n = pm.pgm.tp_size
        symm = pm.pgm.symm
        up = gout.detach().reshape(1).float().contiguous()
        dlogits = fused.ce_grad(logits, targets_flat, all_stats, n, loss_acc, ctx.vocab_start, ctx.ignore_index,
                                upstream=up)
        rt.on_side(lambda: _wgrad(dlogits, h, rt.g_wlm), keep=(dlogits, h))
        fused.colsum(dlogits, out=rt.g_blm)
        dx_l, dx = _handoff_bwd(dlogits, rt.wlm, x_l, rstd, rt.sN, None, rt.next_bzone(), rt.g_sN, rt.packs[-1].g_bd)
        rt.put_stash(dx_l, dx)
        return (dx_l, None, None, None) + (None,) * 3
