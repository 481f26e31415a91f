"""Fused sm_100a Transformer path vs the fp32 module-path oracle on the same weights; engine / graph / checkpoint."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"
CFG = dict(attn_dim=256, ffn_dim=1024, num_heads=4, num_layers=2, vocab_size=1024, maxlen=256)


def _model(seed=0, **over):
    import process_manager as pm
    from models.model import Transformer
    pm.init_pgm(1)
    torch.manual_seed(seed)
    m = Transformer(**{**CFG, **over}).to(DEV)
    m.reset_parameters()
    return m


def _batch(B=4, T=200, V=1024, seed=1):
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(3, V, (B, T), generator=g).to(DEV)
    tgt = torch.randint(3, V, (B, T), generator=g).to(DEV)
    tgt[:, -5:] = -1
    pos = torch.arange(T, device=DEV).unsqueeze(0).repeat(B, 1)
    return ids, pos, tgt


def test_fused_loss_and_grads_match_fp32_oracle():
    from distributed_pytorch_from_scratch_b200.ops import _lib
    m = _model()
    ids, pos, tgt = _batch()
    m.compute_dtype = torch.float32
    ref = m.loss(ids, pos, tgt)
    ref.backward()
    ref_grads = {k: p.grad.clone() for k, p in m.named_parameters()}
    for p in m.parameters():
        p.grad = None
    m.compute_dtype = torch.bfloat16
    before = _lib.lib().dpfs_launch_count()
    loss = m.loss(ids, pos, tgt)
    loss.backward()
    assert _lib.lib().dpfs_launch_count() - before > 50, "the native sm_100a path did not run"
    assert abs(loss.item() - ref.item()) < 2e-3
    for k, p in m.named_parameters():
        r = ref_grads[k]
        rel = (p.grad - r).norm().item() / (r.norm().item() + 1e-12)
        assert rel < 0.03, (k, rel)
    with torch.no_grad():
        lg = m(ids, pos).float()
        m.compute_dtype = torch.float32
        lr_ = m(ids, pos).float()
    assert lg.shape == (4, 200, 1024) and (lg - lr_).abs().max().item() < 0.05


def test_head_dim_128_and_autocast_entry():
    m = _model(attn_dim=256, num_heads=2)
    ids, pos, tgt = _batch(B=2, T=130)
    m.compute_dtype = torch.float32
    ref = m.loss(ids, pos, tgt).item()
    m.compute_dtype = None
    with torch.autocast("cuda", dtype=torch.bfloat16):      # the reference's way of asking for bf16 (train.py:99)
        loss = m.loss(ids, pos, tgt)
    loss.backward()
    assert abs(loss.item() - ref) < 2e-3
    assert m.layers[0].attn.wq.weight.grad is not None


def test_train_step_engine_graph_matches_eager_and_learns():
    from distributed_pytorch_from_scratch_b200.engine import TrainStep
    from distributed_pytorch_from_scratch_b200.ops.optim import FusedAdam
    ids, pos, tgt = _batch()
    losses = {}
    for mode in ("eager", "graph"):
        m = _model(seed=3)
        m.compute_dtype = torch.bfloat16
        opt = FusedAdam.for_model(m, lr=1e-3)
        assert opt.arena is not None
        sched = torch.optim.lr_scheduler.OneCycleLR(opt, 1e-3, 50, pct_start=0.2)
        opt._opt_called = True
        step = TrainStep(m, opt, sched, use_graph=(mode == "graph"))
        ls = [step(ids.cpu().pin_memory(), pos.cpu().pin_memory(), tgt.cpu().pin_memory()).item() for _ in range(8)]
        losses[mode] = ls
        assert step.kernels_per_step > 50
    assert losses["eager"][-1] < losses["eager"][0] - 0.3
    for a, b in zip(losses["eager"], losses["graph"]):
        assert abs(a - b) < 2e-2, (losses["eager"], losses["graph"])


def test_fused_adam_equals_torch_adam_on_model():
    """One fused step == one torch.optim.Adam step on the same grads (and the bf16 shadow follows)."""
    from distributed_pytorch_from_scratch_b200.ops.optim import FusedAdam
    m = _model(seed=5)
    m.compute_dtype = torch.bfloat16
    ids, pos, tgt = _batch()
    opt = FusedAdam.for_model(m, lr=1e-3)
    m.loss(ids, pos, tgt).backward()
    ref_p = [p.detach().clone() for p in m.parameters()]
    ref_params = [torch.nn.Parameter(p.clone()) for p in ref_p]
    for rp, p in zip(ref_params, m.parameters()):
        rp.grad = p.grad.clone()
    ropt = torch.optim.Adam(ref_params, lr=1e-3)
    ropt.step()
    opt.step()
    for rp, p in zip(ref_params, m.parameters()):
        assert torch.allclose(p, rp, atol=1e-6, rtol=1e-5)
    arena = opt.arena
    assert torch.equal(arena.shadow, arena.params.to(torch.bfloat16))


def test_fused_adam_no_host_sync_matches_torch_adam_onecycle():
    """48 optimizer steps enqueued WITHOUT any host synchronisation (the host runs far ahead of the GPU, as it does
    under CUDA-graph replay) must use each step's own lr / beta1 / bias corrections: compare with torch.optim.Adam +
    OneCycleLR (reference train.py:83-84,106-109) fed the same gradients.  A single reused pinned staging buffer fails
    this (the async H2D copy reads the host memory when it executes, not when it is enqueued)."""
    from distributed_pytorch_from_scratch_b200.ops.optim import FusedAdam
    m = _model(seed=7)
    m.compute_dtype = torch.bfloat16
    opt = FusedAdam.for_model(m, lr=3e-4)
    arena = opt.arena
    assert arena is not None
    steps, total = 48, 60
    sched = torch.optim.lr_scheduler.OneCycleLR(opt, 3e-4, total, pct_start=0.25)
    opt._opt_called = True
    ref = torch.nn.Parameter(arena.params.detach().clone())
    ropt = torch.optim.Adam([ref], lr=3e-4)
    rsched = torch.optim.lr_scheduler.OneCycleLR(ropt, 3e-4, total, pct_start=0.25)
    gen = torch.Generator(device="cuda").manual_seed(11)
    grads = [torch.randn(arena.numel, device="cuda", generator=gen) * 0.01 for _ in range(steps)]
    # something long-running in front, so every one of the 48 hyper-parameter uploads is enqueued before the first runs
    big = torch.randn(8192, 8192, device="cuda")
    for _ in range(20):
        big = big @ big * 1e-4
    for t in range(steps):          # no .item(), no synchronize inside the loop
        arena.grads.copy_(grads[t])
        opt.update_hyper()
        opt.launch()
        sched.step()
    for t in range(steps):
        ref.grad = grads[t].clone()
        ropt.step()
        rsched.step()
    torch.cuda.synchronize()
    err = (arena.params - ref.data).abs().max().item()
    assert torch.allclose(arena.params, ref.data, atol=2e-6, rtol=1e-5), err     # a stale-hyper step is off by ~1e-4
    # (No bit-wise "graph replay with vs without host syncs" comparison of whole training steps here: the step itself
    # is not run-to-run deterministic -- attention backward accumulates dQ with fp32 TMA reduce-adds and the embedding
    # backward with atomics -- and 40 Adam steps amplify that noise to ~1e-3.  The end-to-end evidence is bench.py's
    # final_loss, which matches the reference arm's to 1e-4 since the ring went in.)


def test_checkpoint_roundtrip_from_arena(tmp_path):
    from distributed_pytorch_from_scratch_b200.utils import checkpoint as C
    m = _model(seed=7)
    m.compute_dtype = torch.bfloat16
    ids, pos, tgt = _batch()
    l0 = m.loss(ids, pos, tgt).item()        # builds the arena: parameters are now views
    p = C.save_checkpoint(m, str(tmp_path), 0, 5, l0, verbose=False)
    sd = torch.load(p, map_location="cpu", weights_only=True)
    assert len(sd) == 36 and all(v.dtype == torch.float32 for v in sd.values())
    assert os.path.getsize(p) < 1.2 * sum(v.numel() * 4 for v in sd.values()) + 100000     # no arena-sized storages
    m2 = _model(seed=8)
    m2.compute_dtype = torch.bfloat16
    _ = m2.loss(ids, pos, tgt)
    C.load_checkpoint(m2, p)                 # in-place load into the arena; shadow must be refreshed
    assert abs(m2.loss(ids, pos, tgt).item() - l0) < 1e-5


def test_standalone_layers_native_path():
    import process_manager as pm
    from models.layers import ColumnParallelLinear, RMSNorm, RowParallelLinear
    pm.init_pgm(1)
    torch.manual_seed(0)
    col = ColumnParallelLinear(256, 512).to(DEV)
    row = RowParallelLinear(512, 256).to(DEV)
    norm = RMSNorm(256).to(DEV)
    col.reset_parameters(); row.reset_parameters()
    x = torch.randn(4, 100, 256, device=DEV, requires_grad=True)
    xr = x.detach().clone().requires_grad_(True)
    ref = row(col(norm(xr)))
    ref.mean().backward()
    gref = col.weight.grad.clone()
    col.zero_grad(); row.zero_grad(); norm.zero_grad()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        y = row(col(norm(x)))
    y.float().mean().backward()
    assert (y.float() - ref).abs().max().item() < 0.05
    assert (col.weight.grad - gref).norm().item() / gref.norm().item() < 0.03
    assert (x.grad - xr.grad).norm().item() / xr.grad.norm().item() < 0.03
