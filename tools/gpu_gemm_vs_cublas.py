"""Every GEMM of one decoder layer (default model, 32000 tokens, TP=1): this library's default dispatch against
torch.matmul / F.linear (cuBLAS, the reference's kernels) on the same box, L2 flushed between iterations.
The fused rows (SwiGLU epilogues) are compared with what they replace: cuBLAS GEMM + the ATen ops of the reference."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from distributed_pytorch_from_scratch_b200.ops import fused
from distributed_pytorch_from_scratch_b200.ops.gemm import gemm_nt, gemm_swiglu, gemm_swiglu_bwd
dev = "cuda"
T = 32000
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def timeit(fn, iters=8):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / iters * 1e3
def bf(*s): return torch.randn(*s, device=dev).bfloat16()
rows = []
def add(name, M, N, K, ours, ref):
    t, r = timeit(ours), timeit(ref)
    line = f"| {name} | {M},{N},{K} | {t:.1f} | {2*M*N*K/t/1e6:.0f} | {r:.1f} | {2*M*N*K/r/1e6:.0f} | {r/t:.2f}x |"
    rows.append(line); print(line, flush=True)
for name, N, K in [("fwd qkv (+bias)", 1536, 512), ("fwd wo", 512, 512), ("fwd down", 512, 2048), ("fwd lm_head (+bias, fp32 out)", 1024, 512)]:
    a, w, b = bf(T, K), bf(N, K), torch.randn(N, device=dev)
    if "fp32" in name:
        add(name, T, N, K, lambda: gemm_nt(a, w, bias=b, out_dtype=torch.float32), lambda: F.linear(a, w).float() + b)
    elif "bias" in name:
        add(name, T, N, K, lambda: gemm_nt(a, w, bias=b), lambda: F.linear(a, w, b.bfloat16()))
    else:
        add(name, T, N, K, lambda: gemm_nt(a, w), lambda: F.linear(a, w))
for name, Nout, Kred in [("dgrad qkv", 512, 1536), ("dgrad gate|up", 512, 4096), ("dgrad wo", 512, 512)]:
    dy, w = bf(T, Kred), bf(Kred, Nout)
    add(name, T, Nout, Kred, lambda: gemm_nt(dy, w, b_mn_major=True), lambda: torch.matmul(dy, w))
for name, Mo, No in [("wgrad qkv", 1536, 512), ("wgrad gate|up", 4096, 512), ("wgrad wo", 512, 512), ("wgrad down", 512, 2048), ("wgrad lm_head", 1024, 512)]:
    dy, x = bf(T, Mo), bf(T, No)
    out = torch.empty(Mo, No, device=dev)
    add(name + " (fp32 out)", Mo, No, T, lambda: gemm_nt(dy, x, a_mn_major=True, b_mn_major=True, out=out),
        lambda: torch.matmul(dy.t(), x).float())
# fused epilogues vs what the reference runs for the same maths
x, wgu, bgu = bf(T, 512), bf(4096, 512), torch.randn(4096, device=dev)
def ref_fwd():
    gu = F.linear(x, wgu, bgu.bfloat16())
    return gu, F.silu(gu[:, :2048]) * gu[:, 2048:]
add("fwd gate|up + bias + SwiGLU (one kernel) vs cuBLAS + ATen silu, mul", T, 4096, 512, lambda: gemm_swiglu(x, wgu, bgu), ref_fwd)
dy, wd, gu, bg = bf(T, 512), bf(512, 2048), bf(T, 4096), torch.empty(4096, device=dev)
def ref_bwd():
    dact = torch.matmul(dy, wd)
    g, u = gu[:, :2048].float(), gu[:, 2048:].float()
    s = torch.sigmoid(g)
    dg = dact * u * s * (1 + g * (1 - s))
    du = dact * g * s
    dgu = torch.cat([dg, du], 1).bfloat16()
    return dgu, dgu.float().sum(0)
add("dgrad down + SwiGLU backward + bias grads (one kernel) vs cuBLAS + ATen", T, 2048, 512, lambda: gemm_swiglu_bwd(dy, wd, gu, bg), ref_bwd)
hdr = ["# Every GEMM of one decoder layer against cuBLAS on the same B200 (default model, 32000 tokens, TP=1, bf16)", "",
       "`python tools/gpu_gemm_vs_cublas.py`; CUDA events, L2 flushed (256 MB write) before every timed launch, mean of 8.",
       "TFLOP/s count the GEMM's 2MNK only (the fused rows do more work in the same kernel).", "",
       "| GEMM | M,N,K | this library (us) | TFLOP/s | cuBLAS / reference ops (us) | TFLOP/s | speed-up |", "|---|---|---|---|---|---|---|"]
os.makedirs("gpurun_out", exist_ok=True)
open("gpurun_out/gemm_vs_cublas.md", "w").write("\n".join(hdr + rows) + "\n")
