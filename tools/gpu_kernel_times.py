"""Device-timed micro-benchmarks of the memory-bound kernels at the default model's shapes, with achieved HBM bandwidth
as a fraction of the MEASURED copy bandwidth (MEASURED_PEAKS.json).  Inputs are far larger than L2 (126 MB) or a
256 MB buffer is rewritten between iterations."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_pytorch_from_scratch_b200.ops import fused
dev = "cuda"
try:
    HBM = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    HBM = 6650.0
M, D, F, V = 32000, 512, 2048, 1024
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def timeit(fn, iters=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / iters * 1e3
x = torch.randn(M, D, device=dev); scale = torch.ones(D, device=dev); dres = torch.randn(M, D, device=dev)
dy = torch.randn(M, D, device=dev).bfloat16()
y, rstd = fused.rmsnorm_fwd(x, scale, 1e-5)
gu = torch.randn(M, 2 * F, device=dev).bfloat16(); dact = torch.randn(M, F, device=dev).bfloat16()
bg = torch.empty(2 * F, device=dev); cs = torch.empty(D, device=dev)
qkv = torch.randn(M, 3 * D, device=dev).bfloat16()
pos = torch.arange(1000, device=dev).repeat(32)
cos = torch.randn(1000, 32, device=dev); sin = torch.randn(1000, 32, device=dev)
n = 51_500_032
p = torch.randn(n, device=dev); g = torch.randn(n, device=dev); m1 = torch.zeros(n, device=dev); v1 = torch.zeros(n, device=dev)
sh = torch.empty(n, dtype=torch.bfloat16, device=dev)
hyper = torch.tensor([1e-3, 0.9, 0.999, 1e-8, 0.1, 0.001, 0.0, 0.0], device=dev)
logits = torch.randn(M, V, device=dev); tgt = torch.randint(0, V, (M,), device=dev)
rows = [
 ("rmsnorm_fwd  [32000,512] fp32->bf16", lambda: fused.rmsnorm_fwd(x, scale, 1e-5), M*D*(4+2)),
 ("add_rmsnorm_fwd (x + y + b, norm)", lambda: fused.add_rmsnorm_fwd(x, dy, scale, scale, 1e-5), M*D*(4+2+4+2)),
 ("rmsnorm_bwd (+dres, fp32+bf16 out)", lambda: fused.rmsnorm_bwd(dy, x, rstd, scale, dres=dres, want_fp32=True, want_bf16=True), M*D*(2+4+4+4+2)),
 ("swiglu_fwd   [32000,2x2048]", lambda: fused.swiglu_fwd(gu), M*F*(4+2)),
 ("swiglu_bwd", lambda: fused.swiglu_bwd(dact, gu), M*F*(2+4+4)),
 ("swiglu_bwd + bias grads (fused)", lambda: fused.swiglu_bwd_bias(dact, gu, bg), M*F*(2+4+4)),
 ("colsum [32000,4096] bf16", lambda: fused.colsum(gu), M*2*F*2),
 ("cast fp32->bf16 [32000,512]", lambda: fused.cast_to_bf16(x), M*D*6),
 ("cast + colsum (fused)", lambda: fused.cast_colsum(x, cs), M*D*6),
 ("rope in place (q,k of [32000,1536])", lambda: fused.rope_inplace(qkv, pos, cos, sin, 2, D, 8, 64), M*2*D*2*2),
 ("adam 51.5M params (+bf16 shadow)", lambda: fused.adam_step(p, g, m1, v1, sh, hyper), n*(16+12+2)),
 ("ce_stats [32000,1024] fp32", lambda: fused.ce_stats(logits, tgt, 0, -1), M*V*4),
]
out = [f"| kernel | time (us) | bytes (MB) | GB/s | fraction of measured copy BW ({HBM:.0f} GB/s) |", "|---|---|---|---|---|"]
for name, fn, nbytes in rows:
    t = timeit(fn)
    line = f"| {name} | {t:.1f} | {nbytes/1e6:.0f} | {nbytes/t/1e3:.0f} | {100*nbytes/t/1e3/HBM:.0f}% |"
    out.append(line); print(line, flush=True)
os.makedirs("gpurun_out", exist_ok=True)
open("gpurun_out/kernel_times.md", "w").write("# Memory-bound kernels at the default model's shapes (B200, L2 flushed between iterations)\n\n" + "\n".join(out) + "\n")
