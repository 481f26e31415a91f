"""GPU bring-up for the tcgen05 flash attention (fwd + bwd) against an fp32 PyTorch oracle.

The kernel variant is picked by the library from the environment (read once per process), so the opt-in kernels are
checked and timed with the same script:
    DPFS_ATTN_PERSISTENT=0 python tools/gpu_attn_debug.py        # one-CTA-per-tile kernels
    DPFS_ATTN_FWD_PP=1     python tools/gpu_attn_debug.py fwd    # ping-pong forward (hd = 64)
    DPFS_ATTN_BWD_V2=1     python tools/gpu_attn_debug.py bwd    # backward with the lighter math path
"""
import os
import subprocess
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

CASES = [  # B, T, H, HD
    (1, 128, 1, 64), (1, 256, 1, 64), (2, 200, 2, 64), (2, 1000, 2, 64), (1, 128, 1, 128), (2, 333, 2, 128),
    (1, 2048, 2, 128),
]


def run(mode):
    import torch
    from distributed_pytorch_from_scratch_b200.ops import attention as A
    torch.manual_seed(0)
    for (B, T, H, HD) in CASES:
        qkv = (torch.randn(B * T, 3 * H * HD, device="cuda") * 1.0).to(torch.bfloat16)
        ref_in = qkv.float().requires_grad_(True)
        ref = A.attention_reference(ref_in, B, T, H, HD)
        t0 = time.time()
        out, lse = A.attn_fwd(qkv, B, T, H, HD, HD ** -0.5)
        torch.cuda.synchronize()
        err = (out.float() - ref).abs().max().item()
        print(f"[fwd] B={B} T={T} H={H} HD={HD} err={err:.4g} ref_max={ref.abs().max().item():.3g} "
              f"{'OK' if err < 0.03 else 'FAIL'} ({time.time()-t0:.2f}s)", flush=True)
        if err >= 0.03:
            d = (out.float() - ref).abs().reshape(B, T, H, HD)
            print("    worst per (b,h):", d.amax(dim=(1, 3)).tolist(), flush=True)
            rows = (d.amax(dim=(0, 2, 3)) > 0.03).nonzero().flatten()
            print("    bad rows:", rows[:20].tolist(), "count", rows.numel(), flush=True)
        if mode == "bwd":
            g = (torch.randn_like(ref) * 1.0)
            ref.backward(g)
            dref = ref_in.grad
            dqkv = A.attn_bwd(qkv, out, g.to(torch.bfloat16), lse, B, T, H, HD, HD ** -0.5)
            torch.cuda.synchronize()
            Dl = H * HD
            for name, sl in (("dq", slice(0, Dl)), ("dk", slice(Dl, 2 * Dl)), ("dv", slice(2 * Dl, 3 * Dl))):
                e = (dqkv[:, sl].float() - dref[:, sl]).abs().max().item()
                sc = dref[:, sl].abs().max().item()
                print(f"    [bwd] {name}: err={e:.4g} scale={sc:.3g} {'OK' if e < 0.03 * max(sc, 1) else 'FAIL'}",
                      flush=True)
    # timing, flagship shape
    for (B, T, H, HD) in [(32, 1000, 8, 64), (8, 2048, 16, 128)]:
        qkv = torch.randn(B * T, 3 * H * HD, device="cuda").to(torch.bfloat16)
        out, lse = A.attn_fwd(qkv, B, T, H, HD, HD ** -0.5)
        g = torch.randn_like(out)
        for fn, name in ((lambda: A.attn_fwd(qkv, B, T, H, HD, HD ** -0.5), "fwd"),
                         (lambda: A.attn_bwd(qkv, out, g, lse, B, T, H, HD, HD ** -0.5), "bwd")):
            if name == "bwd" and mode != "bwd":
                continue
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10):
                fn()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 10
            flops = 4 * B * H * T * T * HD / 2 * (1 if name == "fwd" else 2.5)
            print(f"[time] {name} B={B} T={T} H={H} HD={HD}: {ms*1e3:.1f} us  {flops/ms/1e9:.1f} TFLOP/s (causal)",
                  flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 1:
        run(sys.argv[1])
        sys.exit(0)
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/attn_debug.log", "w") as log:
        for mode in ("fwd", "bwd"):
            try:
                p = subprocess.run([sys.executable, __file__, mode], capture_output=True, text=True, timeout=200)
                txt = p.stdout + ("\nSTDERR:\n" + p.stderr[-3000:] if p.returncode != 0 else "")
                txt += f"\n[{mode}] exit code {p.returncode}\n"
            except subprocess.TimeoutExpired as e:
                txt = (e.stdout.decode() if e.stdout else "") + f"\n[{mode}] TIMEOUT\n"
            log.write(txt)
            log.flush()
            print(txt, flush=True)
