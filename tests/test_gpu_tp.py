"""Multi-GPU tensor-parallel tests (NCCL + our NVLink kernels).  Skipped unless the box has >= 2 GPUs."""
import os

import pytest
import torch

from conftest import free_port
from dist_utils import run_distributed

pytestmark = pytest.mark.gpu


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _comm_worker(rank, world):
    import torch.distributed as dist
    from distributed_pytorch_from_scratch_b200.ops.gemm import gemm_nt
    from distributed_pytorch_from_scratch_b200.parallel.symm import try_enable_symm
    symm = try_enable_symm(required=True)
    torch.manual_seed(100 + rank)
    for dtype, shape in [(torch.bfloat16, (4000, 512)), (torch.float32, (1024, 256))]:
        x = torch.randn(shape, device="cuda").to(dtype)
        ref = x.clone()
        dist.all_reduce(ref)
        out = symm.all_reduce(x)
        assert (out.float() - ref.float()).abs().max().item() <= (0.07 if dtype == torch.bfloat16 else 1e-4)
    st = torch.randn(1000, 4, device="cuda")
    g = symm.all_gather_rows(st)
    ref = torch.empty(world, 1000, 4, device="cuda")
    dist.all_gather_into_tensor(ref.view(-1), st.view(-1))
    assert torch.equal(g, ref)
    # both data paths of the sequence-parallel halves and of the all-reduce (P2P loads/stores vs NVLS multimem)
    part_ref = torch.randn(2048, 512, device="cuda").to(torch.bfloat16)
    full = part_ref.float().clone()
    dist.all_reduce(full)
    loc = torch.randn(2048 // world, 256, device="cuda").to(torch.bfloat16)
    ag_ref = torch.empty(2048, 256, device="cuda", dtype=torch.bfloat16)
    dist.all_gather_into_tensor(ag_ref, loc)
    for mode in (["p2p", "nvls"] if symm.mc_base else ["p2p"]):
        symm.mode = mode
        buf = symm.buffer((2048, 512), torch.bfloat16)
        buf.copy_(part_ref)
        rs = symm.reduce_scatter_rows(buf)
        rows = 2048 // world
        assert (rs.float() - full[rank * rows:(rank + 1) * rows]).abs().max().item() < 0.07, mode
        assert torch.equal(symm.all_gather_into_buffer(loc), ag_ref), mode
        buf.copy_(part_ref)
        ar = symm.all_reduce(buf, out=buf)
        assert (ar.float() - full).abs().max().item() < 0.07, mode
    symm.mode = "p2p"
    for (M, N, K, bmn) in [(1000, 512, 256, False), (4096, 512, 768, True)]:
        a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
        w = (torch.randn(K, N, device="cuda") if bmn else torch.randn(N, K, device="cuda")).to(torch.bfloat16)
        # oracle: plain PyTorch fp32 matmul of the same bf16 operands + NCCL sum
        ref = a.float() @ (w.float() if bmn else w.float().t())
        dist.all_reduce(ref)
        for mode in (["p2p", "nvls"] if symm.mc_base else ["p2p"]):
            symm.mode = mode
            assert symm.can_fuse_gemm_ar(a, w, bmn), mode
            for _ in range(3):      # repeated calls reuse the same semaphores
                out = symm.gemm_allreduce(a, w, b_mn_major=bmn).clone()
            assert (out.float() - ref).abs().max().item() < 0.03 * ref.abs().max().item(), mode
        symm.mode = "p2p"


def _fused_rs_worker(rank, world):
    """GEMM -> reduce-scatter in one kernel vs fp32 GEMM + NCCL all-reduce + row slice; row ranges that straddle the
    128-row tiles (1000 rows per rank) and aligned ones, K-major and MN-major B, both reduce data paths, repeated calls."""
    import torch.distributed as dist
    from distributed_pytorch_from_scratch_b200.ops.gemm import gemm_nt
    from distributed_pytorch_from_scratch_b200.parallel.symm import try_enable_symm
    symm = try_enable_symm(required=True)
    torch.manual_seed(200 + rank)
    for mode in (["p2p", "nvls"] if symm.mc_base else ["p2p"]):
        symm.mode = mode
        for (M, N, K, bmn) in [(1000 * world, 512, 256, False), (2048 * world, 768, 320, True),
                               (128 * world, 256, 64, False)]:
            a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
            w = (torch.randn(K, N, device="cuda") if bmn else torch.randn(N, K, device="cuda")).to(torch.bfloat16)
            ref = a.float() @ (w.float() if bmn else w.float().t())     # fp32 PyTorch oracle
            dist.all_reduce(ref)
            rows = M // world
            ref = ref[rank * rows:(rank + 1) * rows]
            assert symm.can_fuse_gemm_rs(a, w, bmn)
            for _ in range(3):      # repeated calls reuse the same semaphores
                out = symm.gemm_reduce_scatter(a, w, b_mn_major=bmn)
            torch.cuda.synchronize()
            assert out.shape == (rows, N)
            err = (out.float() - ref).abs().max().item()
            assert err < 0.03 * ref.abs().max().item(), (mode, M, N, K, bmn, err)
    symm.mode = "p2p"


def _push_worker(rank, world):
    """Producers that all-gather their bf16 output through the multicast mapping vs the same producers followed by the
    stand-alone all-gather; several rounds so both landing zones are recycled."""
    import os
    os.environ["DPFS_FUSE_AG"] = "1"
    import torch.distributed as dist
    from distributed_pytorch_from_scratch_b200.ops import fused
    from distributed_pytorch_from_scratch_b200.parallel.symm import try_enable_symm
    symm = try_enable_symm(required=True)
    if not symm.push_enabled:
        return      # no NVLS multicast on this box: nothing to test
    torch.manual_seed(300 + rank)
    Ml, D = 1000, 512

    def gather(t):
        out = torch.empty((world * t.shape[0],) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        dist.all_gather_into_tensor(out, t.contiguous())
        return out

    for rnd in range(5):
        x = torch.randn(Ml, D, device="cuda")
        y = torch.randn(Ml, D, device="cuda").to(torch.bfloat16)
        bias, scale = torch.randn(D, device="cuda"), torch.rand(D, device="cuda") + 0.5
        h_ref, rstd_ref = fused.rmsnorm_fwd(x, scale, 1e-5)
        h, rstd = fused.rmsnorm_fwd_push(x, scale, 1e-5, symm)
        assert torch.equal(h, gather(h_ref)) and torch.equal(rstd, rstd_ref), rnd
        xo_ref, h2_ref, r2_ref = fused.add_rmsnorm_fwd(x, y, bias, scale, 1e-5)
        xo, h2, r2 = fused.add_rmsnorm_fwd_push(x, y, bias, scale, 1e-5, symm)
        assert torch.equal(xo, xo_ref) and torch.equal(h2, gather(h2_ref)) and torch.equal(r2, r2_ref), rnd
        assert torch.equal(fused.cast_to_bf16_push(x, symm), gather(fused.cast_to_bf16(x))), rnd
        dy, dres = torch.randn(Ml, D, device="cuda").to(torch.bfloat16), torch.randn(Ml, D, device="cuda")
        ds_ref, ds = torch.empty(D, device="cuda"), torch.empty(D, device="cuda")
        dx_ref, dxb_ref, _ = fused.rmsnorm_bwd(dy, x, rstd_ref, scale, dres=dres, want_fp32=True, want_bf16=True,
                                               dscale_out=ds_ref)
        dx, dxb = fused.rmsnorm_bwd_push(dy, x, rstd_ref, scale, dres, symm, dscale_out=ds)
        assert torch.equal(dx, dx_ref) and torch.equal(dxb, gather(dxb_ref)) and torch.equal(ds, ds_ref), rnd
        # vocab-parallel embedding: owner-scatter == masked lookup + all-reduce (exact: one owner per token)
        V, Vl = 1024, 1024 // world
        w = torch.randn(Vl, D, device="cuda")
        ids = torch.randint(0, V, (Ml * 2,), device="cuda")
        dist.broadcast(ids, 0)
        part = fused.embedding_fwd(ids, w, rank * Vl, (rank + 1) * Vl, out_dtype=torch.bfloat16)
        dist.all_reduce(part)
        assert torch.equal(fused.embedding_fwd_scatter(ids, w, rank * Vl, (rank + 1) * Vl, symm), part), rnd
    torch.cuda.synchronize()


def _handoff_worker(rank, world):
    """One-kernel sequence-parallel hand-offs (reduce-scatter -> row op -> all-gather, csrc/sp_fused.cu) against the
    same maths in plain PyTorch fp32 on the NCCL-summed partials; several rounds re-use the barrier semaphores."""
    import torch.distributed as dist
    from distributed_pytorch_from_scratch_b200.ops import fused
    from distributed_pytorch_from_scratch_b200.parallel.symm import try_enable_symm
    symm = try_enable_symm(required=True)
    if not symm.mc_base:
        return      # no NVLS multicast on this box
    torch.manual_seed(400 + rank)

    def gather(t):
        out = torch.empty((world * t.shape[0],) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        dist.all_gather_into_tensor(out, t.contiguous())
        return out

    for (Ml, D) in [(250, 512), (129, 256), (64, 1024)]:
        M = Ml * world
        zones = symm.alloc_zones(2, M * D * 2)
        for rnd in range(3):
            part = symm.buffer((M, D), torch.bfloat16)
            mine = torch.randn(M, D, device="cuda").to(torch.bfloat16)
            part.copy_(mine)
            total = mine.float()
            dist.all_reduce(total)
            y_l = total[rank * Ml:(rank + 1) * Ml].to(torch.bfloat16).float()    # the switch rounds the sum to bf16
            x_l = torch.randn(Ml, D, device="cuda")
            bias, scale = torch.randn(D, device="cuda"), torch.rand(D, device="cuda") + 0.5
            # ---- forward
            xo, h, rstd = fused.rs_add_rmsnorm_ag(symm, part, x_l, bias, scale, 1e-5, zones[0])
            xo_ref = x_l + y_l + bias
            rstd_ref = torch.rsqrt(xo_ref.pow(2).mean(-1) + 1e-5)
            h_ref = gather((xo_ref * rstd_ref[:, None] * scale).to(torch.bfloat16))
            # the in-switch fp32 sum may round differently from NCCL's: allow one bf16 ulp of the partial sum
            assert (xo - xo_ref).abs().max().item() < 0.07, (Ml, D, rnd)
            assert (rstd - rstd_ref).abs().max().item() < 2e-3 * rstd_ref.abs().max().item()
            assert (h.float() - h_ref.float()).abs().max().item() < 0.06, (Ml, D, rnd)
            assert torch.equal(h, gather(h[rank * Ml:(rank + 1) * Ml])), "ranks disagree on the gathered rows"
            # embedding hand-off: no residual, no bias
            part.copy_(mine)
            xo2, h2, _ = fused.rs_add_rmsnorm_ag(symm, part, None, None, scale, 1e-5, zones[1])
            assert (xo2 - y_l).abs().max().item() < 0.07
            # ---- backward
            part.copy_(mine)
            dres = torch.randn(Ml, D, device="cuda")
            ds, db = torch.empty(D, device="cuda"), torch.empty(D, device="cuda")
            dx_l, dx = fused.rs_rmsnorm_bwd_ag(symm, part, x_l, rstd_ref, scale, dres, zones[1], ds, db)
            xh = x_l * rstd_ref[:, None]
            g = y_l * scale
            dx_ref = dres + rstd_ref[:, None] * (g - xh * (g * xh).mean(-1, keepdim=True))
            assert (dx_l - dx_ref).abs().max().item() < 0.05 * max(1.0, dx_ref.abs().max().item()), (Ml, D, rnd)
            assert (dx.float() - gather(dx_ref.to(torch.bfloat16)).float()).abs().max().item() \
                < 0.05 * max(1.0, dx_ref.abs().max().item())
            ds_ref, db_ref = (y_l * xh).sum(0), dx_ref.sum(0)
            assert (ds - ds_ref).abs().max().item() < 0.02 * ds_ref.abs().max().item() + 0.05
            assert (db - db_ref).abs().max().item() < 0.02 * db_ref.abs().max().item() + 0.05
    torch.cuda.synchronize()


def _gemm_handoff_worker(rank, world):
    """GEMM + hand-off in ONE kernel (csrc/gemm_handoff.cu) against fp32 PyTorch matmul + NCCL sum + the row op in
    plain PyTorch: row ranges that straddle the 128-row tiles, all three row widths, K-major and MN-major weights,
    repeated calls (the tile flags carry a launch epoch that has to advance in step on every rank)."""
    import torch.distributed as dist
    from distributed_pytorch_from_scratch_b200.ops import fused
    from distributed_pytorch_from_scratch_b200.parallel.symm import try_enable_symm
    symm = try_enable_symm(required=True)
    if not symm.mc_base:
        return
    torch.manual_seed(500 + rank)

    def gather(t):
        out = torch.empty((world * t.shape[0],) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        dist.all_gather_into_tensor(out, t.contiguous())
        return out

    for (Ml, N, K) in [(500, 512, 256), (256, 1024, 128), (130, 256, 64), (4000, 512, 64)]:
        M = Ml * world
        zones = symm.alloc_zones(2, M * N * 2)
        amp = float(K * world) ** -0.25         # the reduced product has unit scale at every TP size
        a = (torch.randn(M, K, device="cuda") * amp).to(torch.bfloat16)
        w = (torch.randn(N, K, device="cuda") * amp).to(torch.bfloat16)          # forward: [N, K], K-major
        wt = w.t().contiguous()                                                    # backward: [K, N], MN-major
        assert fused.can_gemm_handoff(symm, a, w) and fused.can_gemm_handoff(symm, a, wt, True)
        # oracle with the kernel's rounding points: each rank's partial product is stored as bf16, the switch sums
        # the n partials in fp32 and returns a bf16 row
        total = (a.float() @ w.float().t()).to(torch.bfloat16).float()
        dist.all_reduce(total)
        y_l = total[rank * Ml:(rank + 1) * Ml].to(torch.bfloat16).float()
        sc_ = max(1.0, y_l.abs().max().item())
        x_l = torch.randn(Ml, N, device="cuda")
        bias, scale = torch.randn(N, device="cuda"), torch.rand(N, device="cuda") + 0.5
        for rnd in range(3):
            xo, h, rstd = fused.gemm_handoff_fwd(symm, a, w, False, x_l, bias, scale, 1e-5, zones[0])
            xo_ref = x_l + y_l + bias
            rstd_ref = torch.rsqrt(xo_ref.pow(2).mean(-1) + 1e-5)
            h_ref = gather((xo_ref * rstd_ref[:, None] * scale).to(torch.bfloat16))
            assert (xo - xo_ref).abs().max().item() < 0.02 * sc_, (Ml, N, K, rnd, (xo - xo_ref).abs().max().item())
            assert (h.float() - h_ref.float()).abs().max().item() < 0.08, (Ml, N, K, rnd)
            assert torch.equal(h, gather(h[rank * Ml:(rank + 1) * Ml])), "ranks disagree on the gathered rows"
            dres = torch.randn(Ml, N, device="cuda")
            ds, db = torch.empty(N, device="cuda"), torch.empty(N, device="cuda")
            dx_l, dx = fused.gemm_handoff_bwd(symm, a, wt, True, x_l, rstd_ref, scale, dres, zones[1], ds, db)
            xh = x_l * rstd_ref[:, None]
            g = y_l * scale
            dx_ref = dres + rstd_ref[:, None] * (g - xh * (g * xh).mean(-1, keepdim=True))
            tol = 0.03 * max(1.0, dx_ref.abs().max().item())
            assert (dx_l - dx_ref).abs().max().item() < tol, (Ml, N, K, rnd)
            assert (dx.float() - gather(dx_ref.to(torch.bfloat16)).float()).abs().max().item() < tol
            ds_ref, db_ref = (y_l * xh).sum(0), dx_ref.sum(0)
            assert (ds - ds_ref).abs().max().item() < 0.02 * ds_ref.abs().max().item() + 0.5
            assert (db - db_ref).abs().max().item() < 0.02 * db_ref.abs().max().item() + 0.5
    torch.cuda.synchronize()


def _layers_worker(rank, world):
    """A toy model built ONLY from the public layer classes (no Transformer.loss): with the symmetric communicator on,
    RowParallelLinear.forward and ColumnParallelLinear's backward run the one-kernel GEMM + all-reduce and the
    embedding runs the native lookup (+ owner scatter); results match the same layers over NCCL collectives."""
    import process_manager as pm
    from torch.profiler import ProfilerActivity, profile
    from distributed_pytorch_from_scratch_b200.parallel.symm import try_enable_symm
    from models.layers import ColumnParallelLinear, ParallelVocabularyEmbedding, RowParallelLinear

    torch.manual_seed(7)
    torch.cuda.manual_seed(7)
    V, D, F_ = 1024, 512, 1024
    emb, col, row = ParallelVocabularyEmbedding(V, D).cuda(), ColumnParallelLinear(D, F_, gather_output=False).cuda(), \
        RowParallelLinear(F_, D, split_input=False).cuda()
    for m in (emb, col, row):
        m.reset_parameters()
    with torch.no_grad():
        col.bias.normal_(); row.bias.normal_()
    ids = torch.randint(0, V, (8, 250), device="cuda")
    torch.distributed.broadcast(ids, 0)

    def run():
        for m in (emb, col, row):
            m.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            y = row(torch.nn.functional.silu(col(emb(ids).to(torch.bfloat16))))
        loss = y.float().pow(2).mean()
        loss.backward()
        return y.detach().float().clone(), [p.grad.detach().float().clone() for m in (emb, col, row)
                                            for p in m.parameters()]

    y_ref, g_ref = run()                      # NCCL collectives (pgm.symm is None)
    symm = try_enable_symm(required=True)
    run()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        y, g = run()
        torch.cuda.synchronize()
    names = [e.name for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    fused_ar = [n for n in names if "gemm_bf16_kernel" in n and "true, false>" in n.replace("(bool)1", "true")
                .replace("(bool)0", "false")]
    assert len(fused_ar) >= 2, names          # row-parallel forward + column-parallel dgrad
    assert any("embedding_fwd" in n for n in names) and any("embedding_bwd" in n for n in names), names
    assert not any("nccl" in n.lower() and "AllReduce" in n for n in names), names
    assert (y - y_ref).abs().max().item() < 0.05 * max(1.0, y_ref.abs().max().item())
    for a, b in zip(g, g_ref):
        assert (a - b).norm().item() <= 0.03 * b.norm().item() + 1e-6
    pm.pgm.symm = None


def _model_worker(rank, world, comm, sp=False, poison=False, fused_gemm=False):
    import process_manager as pm
    if fused_gemm:   # GEMM + hand-off as one kernel (read at import of models.fused_transformer)
        os.environ["DPFS_SP_FUSED_GEMM"] = "1"
    if poison:      # debug mode: views of the symmetric buffer that outlive the next collective read as NaN
        os.environ["DPFS_SYMM_POISON"] = "1"
    from distributed_pytorch_from_scratch_b200.engine import TrainStep
    from distributed_pytorch_from_scratch_b200.ops.optim import FusedAdam
    from distributed_pytorch_from_scratch_b200.parallel.symm import try_enable_symm
    from models.model import Transformer, VallinaTransformer
    if comm == "fused":
        try_enable_symm(required=True)
    cfg = dict(attn_dim=512, ffn_dim=1024, num_heads=8, num_layers=2, vocab_size=1024, maxlen=256)
    torch.manual_seed(0)
    torch.cuda.manual_seed(0)
    state, cstate = torch.get_rng_state(), torch.cuda.get_rng_state()
    model = Transformer(**cfg, sequence_parallel=sp).cuda()
    torch.set_rng_state(state); torch.cuda.set_rng_state(cstate)
    model.reset_parameters()
    van = VallinaTransformer(**cfg).cuda()
    torch.set_rng_state(state); torch.cuda.set_rng_state(cstate)
    van.reset_parameters()
    g = torch.Generator().manual_seed(1)
    B = 4 if world <= 4 else 8          # >= 128 rows per rank so the fused GEMM -> reduce-scatter path is eligible
    ids = torch.randint(3, 1024, (B, 200), generator=g).cuda()
    tgt = torch.randint(3, 1024, (B, 200), generator=g).cuda()
    pos = torch.arange(200, device="cuda").unsqueeze(0).repeat(B, 1)
    ref_logits = van(ids, pos)
    ref = torch.nn.functional.cross_entropy(ref_logits.view(-1, 1024), tgt.view(-1))
    ref.backward()
    model.compute_dtype = torch.bfloat16
    loss = model.loss(ids, pos, tgt)
    loss.backward()
    assert abs(loss.item() - ref.item()) < 5e-3, (loss.item(), ref.item())
    wq = model.layers[0].attn.wq
    op = wq.odim_partition
    gr = van.layers[0].attn.wq.weight.grad[op * rank:op * (rank + 1)]
    assert (wq.weight.grad - gr).norm().item() / gr.norm().item() < 0.05
    dn = model.layers[1].ffn.down_proj
    ip = dn.idim_partition
    gr = van.layers[1].ffn.down_proj.weight.grad[:, ip * rank:ip * (rank + 1)]
    assert (dn.weight.grad - gr).norm().item() / gr.norm().item() < 0.05
    assert (model.norm.scale.grad - van.norm.scale.grad).norm().item() / van.norm.scale.grad.norm().item() < 0.05
    # replicated biases of the row-parallel layers: under sequence parallelism their gradients are row-partial sums
    # completed by the once-per-step all-reduce of the replicated-parameter block
    for got, want in ((model.layers[0].attn.wo.bias.grad, van.layers[0].attn.wo.bias.grad),
                      (model.layers[1].ffn.down_proj.bias.grad, van.layers[1].ffn.down_proj.bias.grad),
                      (model.layers[0].norm2.scale.grad, van.layers[0].norm2.scale.grad)):
        assert (got - want).norm().item() / want.norm().item() < 0.05
    with torch.no_grad():
        lg = model(ids, pos).float()
    assert lg.shape == (B, 200, 1024) and (lg - ref_logits).abs().max().item() < 0.08
    # a few optimizer steps through the engine (CUDA graph when our own collectives are active); a fresh model, as
    # in train.py: the step is captured before any eager backward has run on the default stream
    del model
    torch.set_rng_state(state); torch.cuda.set_rng_state(cstate)
    model = Transformer(**cfg, sequence_parallel=sp).cuda()
    model.reset_parameters()
    model.compute_dtype = torch.bfloat16
    opt = FusedAdam.for_model(model, lr=1e-3)
    step = TrainStep(model, opt, None, use_graph=True)
    assert step.use_graph == (pm.pgm.symm is not None)
    losses = [step(ids, pos, tgt).item() for _ in range(6)]
    assert losses[-1] < losses[0] - 0.2, losses


def _worlds():
    """TP sizes to run: every power of two the box has GPUs for (the driver's test lease has one GPU -> all skipped
    there; the committed logs under profiles/ are from 2- and 8-GPU boxes)."""
    return [pytest.param(n, marks=pytest.mark.skipif(_ngpu() < n, reason=f"needs >= {n} GPUs")) for n in (2, 4, 8)]


@pytest.mark.parametrize("world", _worlds())
def test_symm_collectives_and_fused_gemm_allreduce(world):
    run_distributed(_comm_worker, world, free_port(), backend="nccl")


@pytest.mark.parametrize("world", _worlds())
def test_fused_gemm_reduce_scatter(world):
    run_distributed(_fused_rs_worker, world, free_port(), backend="nccl")


@pytest.mark.parametrize("world", _worlds())
def test_push_allgather_producers(world):
    run_distributed(_push_worker, world, free_port(), backend="nccl")


@pytest.mark.parametrize("world", _worlds())
def test_gemm_handoff_one_kernel(world):
    run_distributed(_gemm_handoff_worker, world, free_port(), backend="nccl")


@pytest.mark.parametrize("world", _worlds())
def test_public_layer_classes_run_fused_collectives(world):
    run_distributed(_layers_worker, world, free_port(), backend="nccl")


@pytest.mark.parametrize("world", _worlds())
def test_sp_handoff_kernels(world):
    run_distributed(_handoff_worker, world, free_port(), backend="nccl")


@pytest.mark.parametrize("world", _worlds())
@pytest.mark.parametrize("comm", ["nccl", "fused"])
def test_fused_transformer_matches_oracle(world, comm):
    run_distributed(_model_worker, world, free_port(), backend="nccl", args=(comm,))


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("sp", [False, True])
def test_no_stale_symmetric_views_under_poison_mode(sp):
    """DPFS_SYMM_POISON=1 NaN-fills every symmetric-buffer view as soon as the next collective retires it: the model
    still matches the oracle and trains, so no consumer reads a result past its lifetime."""
    run_distributed(_model_worker, 2, free_port(), backend="nccl", args=("fused", sp, True))


@pytest.mark.parametrize("world", _worlds())
def test_sequence_parallel_one_kernel_gemm_handoff_matches_oracle(world):
    run_distributed(_model_worker, world, free_port(), backend="nccl", args=("fused", True, False, True))


@pytest.mark.parametrize("world", _worlds())
@pytest.mark.parametrize("comm", ["nccl", "fused"])
def test_sequence_parallel_matches_oracle(world, comm):
    run_distributed(_model_worker, world, free_port(), backend="nccl", args=(comm, True))
