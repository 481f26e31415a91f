"""Numerics of every hand-written sm_100a kernel against a plain PyTorch fp32 reference of the same op."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _bf(x):
    return x.to(torch.bfloat16)


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
# K <= 1024 runs the 8-epilogue-warp shape of the 2-SM kernel, K = 2048 the 4-warp / 6-stage one
@pytest.mark.parametrize("shape", [(128, 128, 64), (1000, 512, 512), (520, 264, 200), (4096, 1536, 512),
                                   (640, 512, 2048)])
def test_gemm_tcgen05(a_mn, b_mn, shape):
    from distributed_pytorch_from_scratch_b200.ops.gemm import gemm_nt
    M, N, K = shape
    torch.manual_seed(0)
    A, B = _bf(torch.randn(M, K, device=DEV)), _bf(torch.randn(N, K, device=DEV))
    bias, res = torch.randn(N, device=DEV), torch.randn(M, N, device=DEV)
    ref = A.float() @ B.float().t()
    a_in = A.t().contiguous() if a_mn else A
    b_in = B.t().contiguous() if b_mn else B
    scale = ref.abs().max().item()
    for bn in (128, 256):
        out = gemm_nt(a_in, b_in, a_mn_major=a_mn, b_mn_major=b_mn, out_dtype=torch.float32, bn=bn, splits=1,
                      two_sm=False)
        assert (out - ref).abs().max().item() < 2e-3 * scale
        out = gemm_nt(a_in, b_in, a_mn_major=a_mn, b_mn_major=b_mn, out_dtype=torch.bfloat16, bn=bn, bias=bias,
                      residual=res, splits=1)
        assert (out.float() - (ref + bias + res)).abs().max().item() < 2e-2 * scale
        out = gemm_nt(a_in, b_in, a_mn_major=a_mn, b_mn_major=b_mn, out_dtype=torch.float32, bn=bn, bias=bias,
                      splits=3)
        assert (out - (ref + bias)).abs().max().item() < 2e-3 * scale
    # 2-SM (cta_group::2, 256x256 tiles on CTA pairs)
    out = gemm_nt(a_in, b_in, a_mn_major=a_mn, b_mn_major=b_mn, out_dtype=torch.float32, two_sm=True, splits=1)
    assert (out - ref).abs().max().item() < 2e-3 * scale
    out = gemm_nt(a_in, b_in, a_mn_major=a_mn, b_mn_major=b_mn, out_dtype=torch.bfloat16, two_sm=True, bias=bias,
                  residual=res, splits=1)
    assert (out.float() - (ref + bias + res)).abs().max().item() < 2e-2 * scale
    out = gemm_nt(a_in, b_in, a_mn_major=a_mn, b_mn_major=b_mn, out_dtype=torch.float32, two_sm=True, bias=bias,
                  splits=2)
    assert (out - (ref + bias)).abs().max().item() < 2e-3 * scale


@pytest.mark.parametrize("shape", [(32000, 2048, 512), (1000, 256, 512), (520, 128, 192), (4096, 1024, 1024)])
def test_gemm_swiglu_epilogue(shape):
    """gate|up projection with SwiGLU in the 2-SM GEMM's epilogue vs fp32 PyTorch: gu = x W^T + b, act = silu(g) * u
    computed from the bf16-rounded g, u (what the backward reads)."""
    from distributed_pytorch_from_scratch_b200.ops.gemm import gemm_swiglu
    M, F, K = shape
    torch.manual_seed(0)
    x, w = _bf(torch.randn(M, K, device=DEV)), _bf(torch.randn(2 * F, K, device=DEV) * K ** -0.5)
    bias = torch.randn(2 * F, device=DEV)
    gu, act = gemm_swiglu(x, w, bias)
    ref = x.float() @ w.float().t() + bias
    sc = ref.abs().max().item()
    assert gu.shape == (M, 2 * F) and act.shape == (M, F)
    assert (gu.float() - ref).abs().max().item() < 1e-2 * sc
    g, u = gu.float()[:, :F], gu.float()[:, F:]
    act_ref = torch.nn.functional.silu(g) * u
    assert (act.float() - act_ref).abs().max().item() < 1e-2 * max(1.0, act_ref.abs().max().item())


@pytest.mark.parametrize("shape", [(32000, 2048, 512), (1000, 256, 512), (520, 512, 192), (4096, 1024, 1024)])
def test_gemm_swiglu_bwd_epilogue(shape):
    """down_proj dgrad with the SwiGLU backward in the epilogue vs fp32 PyTorch autograd of silu(g) * u."""
    from distributed_pytorch_from_scratch_b200.ops.gemm import gemm_swiglu_bwd
    M, F, K = shape
    torch.manual_seed(0)
    dy = _bf(torch.randn(M, K, device=DEV))
    w = _bf(torch.randn(K, F, device=DEV) * K ** -0.5)          # down_proj weight as stored: [d, F]
    gu = _bf(torch.randn(M, 2 * F, device=DEV))
    bg = torch.empty(2 * F, device=DEV)
    dgu = gemm_swiglu_bwd(dy, w, gu, bg)
    dact = _bf(dy.float() @ w.float()).float()                   # the stand-alone path rounds dact to bf16 too
    g = gu.float()[:, :F].clone().requires_grad_(True)
    u = gu.float()[:, F:].clone().requires_grad_(True)
    (torch.nn.functional.silu(g) * u).backward(dact)
    ref = torch.cat([g.grad, u.grad], dim=1)
    sc = max(1.0, ref.abs().max().item())
    assert (dgu.float() - ref).abs().max().item() < 2e-2 * sc
    bsum = dgu.float().sum(0)
    assert (bg - bsum).abs().max().item() < 1e-3 * max(1.0, bsum.abs().max().item()) + 1e-2


@pytest.mark.parametrize("B,T,H,HD", [(1, 128, 1, 64), (2, 200, 2, 64), (2, 1000, 2, 64), (2, 333, 2, 128)])
def test_flash_attention_fwd_bwd(B, T, H, HD):
    from distributed_pytorch_from_scratch_b200.ops import attention as A
    torch.manual_seed(0)
    qkv = _bf(torch.randn(B * T, 3 * H * HD, device=DEV))
    x = qkv.float().requires_grad_(True)
    ref = A.attention_reference(x, B, T, H, HD)
    g = torch.randn_like(ref)
    ref.backward(g)
    q = qkv.clone().requires_grad_(True)
    out = A.flash_attn_qkvpacked(q, B, T, H, HD)
    out.backward(_bf(g))
    assert (out.float() - ref).abs().max().item() < 0.03
    Dl = H * HD
    for i in range(3):
        e = (q.grad[:, i * Dl:(i + 1) * Dl].float() - x.grad[:, i * Dl:(i + 1) * Dl]).abs().max().item()
        assert e < 0.03 * max(1.0, x.grad.abs().max().item()), (i, e)


@pytest.mark.parametrize("D", [128, 512, 2048])
def test_rmsnorm_fwd_bwd(D):
    from distributed_pytorch_from_scratch_b200.ops import fused
    torch.manual_seed(0)
    M = 777
    x = torch.randn(M, D, device=DEV)
    scale = torch.rand(D, device=DEV) + 0.5
    dy = torch.randn(M, D, device=DEV)
    dres = torch.randn(M, D, device=DEV)
    xr, sr = x.clone().requires_grad_(True), scale.clone().requires_grad_(True)
    yr = sr * (xr * (xr.pow(2).mean(-1, keepdim=True) + 1e-5).rsqrt())
    yr.backward(_bf(dy).float())
    y, rstd = fused.rmsnorm_fwd(x, scale, 1e-5)
    assert (y.float() - yr).abs().max().item() < 0.03
    dx, dxb, dscale = fused.rmsnorm_bwd(_bf(dy), x, rstd, scale, dres=dres, want_fp32=True, want_bf16=True)
    assert torch.allclose(dx, xr.grad + dres, atol=2e-4, rtol=1e-4)
    assert (dxb.float() - dx).abs().max().item() < 0.05
    assert torch.allclose(dscale, sr.grad, atol=2e-2, rtol=2e-3)
    y32, _ = fused.rmsnorm_fwd(x, scale, 1e-5, out_dtype=torch.float32)
    assert torch.allclose(y32, yr, atol=1e-5)
    # same pass + column sums of dx (bias gradient of the linear in front of the norm); other outputs unchanged
    dbias = torch.full((D,), 7.0, device=DEV)
    dx2, dxb2, dscale2 = fused.rmsnorm_bwd(_bf(dy), x, rstd, scale, dres=dres, want_fp32=True, want_bf16=True,
                                           dbias_out=dbias)
    assert torch.equal(dx2, dx) and torch.equal(dxb2, dxb) and torch.equal(dscale2, dscale)
    assert torch.allclose(dbias, (xr.grad + dres).sum(0), atol=2e-3, rtol=1e-4)


def test_layernorm_fwd_bwd():
    from distributed_pytorch_from_scratch_b200.ops import fused
    torch.manual_seed(0)
    M, D = 300, 1024
    x = torch.randn(M, D, device=DEV)
    g, b = torch.rand(D, device=DEV) + 0.5, torch.randn(D, device=DEV)
    dy = torch.randn(M, D, device=DEV)
    xr, gr, br = (t.clone().requires_grad_(True) for t in (x, g, b))
    yr = torch.nn.functional.layer_norm(xr, (D,), gr, br, 1e-5)
    yr.backward(dy)
    y, mean, rstd = fused.layernorm_fwd(x, g, b, 1e-5, out_dtype=torch.float32)
    assert torch.allclose(y, yr, atol=1e-5)
    dx, dg, db = fused.layernorm_bwd(dy, x, mean, rstd, g)
    assert torch.allclose(dx, xr.grad, atol=1e-4)
    assert torch.allclose(dg, gr.grad, atol=1e-3, rtol=1e-4) and torch.allclose(db, br.grad, atol=1e-3, rtol=1e-4)


def test_swiglu_gelu():
    from distributed_pytorch_from_scratch_b200.ops import fused
    torch.manual_seed(0)
    M, F = 513, 1024
    gu = _bf(torch.randn(M, 2 * F, device=DEV))
    dact = _bf(torch.randn(M, F, device=DEV))
    gr = gu.float().requires_grad_(True)
    ar = torch.nn.functional.silu(gr[:, :F]) * gr[:, F:]
    ar.backward(dact.float())
    act = fused.swiglu_fwd(gu)
    assert (act.float() - ar).abs().max().item() < 0.05
    dgu = fused.swiglu_bwd(dact, gu)
    assert (dgu.float() - gr.grad).abs().max().item() < 0.06
    bg = torch.empty(2 * F, device=DEV)
    dgu2 = fused.swiglu_bwd_bias(dact, gu, bg)
    assert torch.equal(dgu2, dgu)
    assert torch.allclose(bg, dgu.float().sum(0), atol=2e-2, rtol=1e-3)
    xf = torch.randn(700, 520, device=DEV)
    cs = torch.empty(520, device=DEV)
    yb = fused.cast_colsum(xf, cs)
    assert torch.equal(yb, xf.to(torch.bfloat16)) and torch.allclose(cs, yb.float().sum(0), atol=1e-2, rtol=1e-4)
    x = _bf(torch.randn(M, F, device=DEV))
    xr = x.float().requires_grad_(True)
    yr = torch.nn.functional.gelu(xr, approximate="tanh")
    yr.backward(dact.float())
    assert (fused.gelu_fwd(x).float() - yr).abs().max().item() < 0.03
    assert (fused.gelu_bwd(dact, x).float() - xr.grad).abs().max().item() < 0.05


def test_rope_matches_reference_convention():
    from distributed_pytorch_from_scratch_b200.ops import fused
    from models.model import apply_rotary_pos_emb, get_cos_sin
    torch.manual_seed(0)
    B, T, H, hd = 2, 50, 3, 64
    Dl = H * hd
    qkv = _bf(torch.randn(B * T, 3 * Dl, device=DEV))
    cos, sin = get_cos_sin(64, hd, 10000., device=torch.device(DEV), dtype=torch.float32)
    pos = torch.randint(0, 64, (B, T), device=DEV)
    q = qkv[:, :Dl].float().view(B, T, H, hd).transpose(1, 2)
    k = qkv[:, Dl:2 * Dl].float().view(B, T, H, hd).transpose(1, 2)
    qr, kr = apply_rotary_pos_emb(q, k, cos[pos], sin[pos])
    buf = qkv.clone()
    half = hd // 2
    ch, sh = cos[:, :half].contiguous(), sin[:, :half].contiguous()
    fused.rope_inplace(buf, pos.reshape(-1).contiguous(), ch, sh, 2, Dl, H, hd)
    assert (buf[:, :Dl].float() - qr.transpose(1, 2).reshape(B * T, Dl)).abs().max().item() < 0.03
    assert (buf[:, Dl:2 * Dl].float() - kr.transpose(1, 2).reshape(B * T, Dl)).abs().max().item() < 0.03
    assert torch.equal(buf[:, 2 * Dl:], qkv[:, 2 * Dl:])
    fused.rope_inplace(buf, pos.reshape(-1).contiguous(), ch, sh, 2, Dl, H, hd, inverse=True)
    assert (buf.float() - qkv.float()).abs().max().item() < 0.06          # inverse rotation == transpose


def test_colsum_embedding_cast():
    from distributed_pytorch_from_scratch_b200.ops import fused
    torch.manual_seed(0)
    x = _bf(torch.randn(3001, 520, device=DEV))
    assert torch.allclose(fused.colsum(x), x.float().sum(0), atol=1e-2, rtol=1e-4)
    V, D, vs, ve = 256, 128, 64, 192
    w = torch.randn(ve - vs, D, device=DEV)
    ids = torch.randint(0, V, (4, 33), device=DEV)
    ids0 = ids.clone()
    out = fused.embedding_fwd(ids, w, vs, ve)
    m = (ids >= vs) & (ids < ve)
    ref = torch.where(m.unsqueeze(-1), w[(ids - vs).clamp(0, ve - vs - 1)], torch.zeros((), device=DEV)).view(-1, D)
    assert torch.equal(out, ref) and torch.equal(ids, ids0)
    dout = torch.randn(ids.numel(), D, device=DEV)
    dw = fused.embedding_bwd(ids, dout, ve - vs, vs, ve)
    dref = torch.zeros_like(w).index_add_(0, (ids.view(-1) - vs)[m.view(-1)], dout[m.view(-1)])
    assert torch.allclose(dw, dref, atol=1e-5)
    f = torch.randn(1000, 64, device=DEV)
    assert torch.equal(fused.cast_to_bf16(f), f.to(torch.bfloat16))


@pytest.mark.parametrize("nranks", [1, 4])
def test_vocab_parallel_cross_entropy(nranks):
    from distributed_pytorch_from_scratch_b200.ops import fused
    torch.manual_seed(0)
    M, V = 600, 1024
    logits = torch.randn(M, V, device=DEV) * 3
    tgt = torch.randint(0, V, (M,), device=DEV)
    tgt[::7] = -1
    lr = logits.clone().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(lr, tgt, ignore_index=-1)
    (ref * 0.5).backward()
    Vl = V // nranks
    shards = [logits[:, r * Vl:(r + 1) * Vl].contiguous() for r in range(nranks)]
    stats, accs = zip(*[fused.ce_stats(s, tgt, r * Vl, -1) for r, s in enumerate(shards)])
    all_stats = torch.stack(stats, 0).contiguous()
    up = torch.tensor([0.5], device=DEV)
    for r in range(nranks):
        fused.ce_loss(tgt, all_stats, nranks, accs[r], -1)
        loss = accs[r][0] / accs[r][1]
        assert abs(loss.item() - ref.item()) < 1e-4
        d = fused.ce_grad(shards[r], tgt, all_stats, nranks, accs[r], r * Vl, -1, upstream=up)
        assert (d.float() - lr.grad[:, r * Vl:(r + 1) * Vl]).abs().max().item() < 2e-5


def test_fused_adam_kernel_matches_torch():
    from distributed_pytorch_from_scratch_b200.ops import fused
    torch.manual_seed(0)
    n = 4096 * 3
    p = torch.randn(n, device=DEV)
    ref = torch.nn.Parameter(p.clone())
    opt = torch.optim.Adam([ref], lr=1e-3, betas=(0.9, 0.999), eps=1e-8)
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    shadow = torch.empty(n, dtype=torch.bfloat16, device=DEV)
    for t in range(1, 6):
        g = torch.randn(n, device=DEV)
        ref.grad = g.clone()
        opt.step()
        hyper = torch.tensor([1e-3, 0.9, 0.999, 1e-8, 1 - 0.9 ** t, 1 - 0.999 ** t, 0.0, 0.0], device=DEV)
        fused.adam_step(p, g, m, v, shadow, hyper)
    assert torch.allclose(p, ref.data, atol=1e-6, rtol=1e-5)
    assert torch.equal(shadow, p.to(torch.bfloat16))
