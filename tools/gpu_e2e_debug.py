"""GPU bring-up of the fused Transformer path: compares loss and grads against the fp32 module-path oracle on the
same weights (TP=1), then runs a few training steps and times them."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import process_manager as pm
from models.model import Transformer
from distributed_pytorch_from_scratch_b200.ops.optim import FusedAdam


def main():
    torch.manual_seed(0)
    pm.init_pgm(1)
    dev = "cuda"
    cfg = dict(attn_dim=256, ffn_dim=1024, num_heads=4, num_layers=2, vocab_size=1024, maxlen=256)
    model = Transformer(**cfg).to(dev)
    model.reset_parameters()
    B, T = 4, 200
    ids = torch.randint(3, 1024, (B, T), device=dev)
    tgt = torch.randint(3, 1024, (B, T), device=dev)
    tgt[:, -7:] = -1
    pos = torch.arange(T, device=dev).unsqueeze(0).repeat(B, 1)

    # oracle: fp32 module path
    model.compute_dtype = torch.float32
    loss_ref = model.loss(ids, pos, tgt)
    loss_ref.backward()
    ref_grads = {k: p.grad.clone() for k, p in model.named_parameters()}
    for p in model.parameters():
        p.grad = None
    # fused bf16 path
    model.compute_dtype = torch.bfloat16
    loss = model.loss(ids, pos, tgt)
    loss.backward()
    torch.cuda.synchronize()
    print(f"loss ref {loss_ref.item():.5f} fused {loss.item():.5f}", flush=True)
    worst = 0
    for k, p in model.named_parameters():
        g, r = p.grad, ref_grads[k]
        rel = (g - r).norm().item() / (r.norm().item() + 1e-12)
        worst = max(worst, rel)
        flag = "" if rel < 0.05 else "  <-- BAD"
        print(f"  {k:40s} |g|={g.norm().item():.4e} |ref|={r.norm().item():.4e} rel_err={rel:.4f}{flag}", flush=True)
    print(f"worst rel err {worst:.4f}", flush=True)

    # logits path
    with torch.no_grad():
        lg = model(ids, pos).float()
        model.compute_dtype = torch.float32
        lr_ = model(ids, pos).float()
        model.compute_dtype = torch.bfloat16
    print(f"logits max abs diff {(lg - lr_).abs().max().item():.4f} (scale {lr_.abs().max().item():.3f})", flush=True)

    # short training run, default 51M model
    from constants import ModelArgumments
    from dataclasses import asdict
    model = Transformer(**asdict(ModelArgumments())).to(dev)
    model.reset_parameters()
    model.compute_dtype = torch.bfloat16
    opt = FusedAdam.for_model(model, lr=3e-4)
    B, T = 32, 1000
    ids = torch.randint(3, 1024, (B, T), device=dev)
    tgt = torch.roll(ids, -1, dims=1)
    pos = torch.arange(T, device=dev).unsqueeze(0).repeat(B, 1)
    for step in range(8):
        torch.cuda.synchronize()
        t0 = time.time()
        loss = model.loss(ids, pos, tgt)
        loss.backward()
        opt.step()
        torch.cuda.synchronize()
        print(f"step {step} loss {loss.item():.4f} {1e3*(time.time()-t0):.1f} ms", flush=True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        loss = model.loss(ids, pos, tgt)
        loss.backward()
        opt.step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(f"eager: {ms:.2f} ms/step  {B*T/ms*1e3:.0f} tok/s  loss {loss.item():.4f}", flush=True)
    print(f"max mem {torch.cuda.max_memory_allocated()/2**30:.2f} GiB", flush=True)


if __name__ == "__main__":
    main()
