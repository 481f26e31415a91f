"""Times every GEMM of one decoder layer (default model, TP=1) under the available kernel variants / split-K factors."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_pytorch_from_scratch_b200.ops.gemm import gemm_nt
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
cases = []
# forward: (name, A[M,K], B[N,K], kwargs)
for name, N, K, extra in [("fwd qkv", 1536, 512, "bias"), ("fwd gate|up", 4096, 512, "bias"), ("fwd wo (+res)", 512, 512, "res"),
                          ("fwd down (+res)", 512, 2048, "res"), ("fwd lm_head fp32", 1024, 512, "f32")]:
    a, w = bf(T, K), bf(N, K)
    kw = {}
    if extra == "bias": kw = dict(bias=torch.randn(N, device=dev))
    if extra == "res": kw = dict(bias=torch.randn(N, device=dev), residual=torch.randn(T, N, device=dev), out_dtype=torch.float32)
    if extra == "f32": kw = dict(bias=torch.randn(N, device=dev), out_dtype=torch.float32)
    cases.append((name, (T, N, K), lambda a=a, w=w, kw=kw, **o: gemm_nt(a, w, **kw, **o), False))
for name, Nout, Kred in [("dgrad qkv", 512, 1536), ("dgrad gate|up", 512, 4096), ("dgrad wo", 512, 512), ("dgrad down", 2048, 512)]:
    dy, w = bf(T, Kred), bf(Kred, Nout)
    cases.append((name, (T, Nout, Kred), lambda dy=dy, w=w, **o: gemm_nt(dy, w, b_mn_major=True, **o), False))
for name, Mo, No in [("wgrad qkv", 1536, 512), ("wgrad gate|up", 4096, 512), ("wgrad wo", 512, 512), ("wgrad down", 512, 2048), ("wgrad lm_head", 1024, 512)]:
    dy, x = bf(T, Mo), bf(T, No)
    out = torch.empty(Mo, No, device=dev)
    cases.append((name, (Mo, No, T), lambda dy=dy, x=x, out=out, **o: gemm_nt(dy, x, a_mn_major=True, b_mn_major=True, out=out, **o), True))
lines = ["| GEMM | M,N,K | default dispatch (us) | TFLOP/s | alternatives (us) |", "|---|---|---|---|---|"]
for name, (M, N, K), fn, is_wgrad in cases:
    t0 = timeit(lambda: fn())
    alts = []
    if is_wgrad:
        for label, o in [("1sm bn128 auto", dict(two_sm=False, bn=128)), ("1sm bn256 auto", dict(two_sm=False, bn=256)),
                         ("2sm s4", dict(two_sm=True, splits=4)), ("2sm s8", dict(two_sm=True, splits=8)),
                         ("2sm s16", dict(two_sm=True, splits=16)), ("2sm s32", dict(two_sm=True, splits=32)),
                         ("1sm128 s8", dict(two_sm=False, bn=128, splits=8)), ("1sm128 s16", dict(two_sm=False, bn=128, splits=16))]:
            try: alts.append(f"{label} {timeit(lambda: fn(**o)):.0f}")
            except Exception as e: alts.append(f"{label} ERR")
    else:
        for label, o in [("1sm bn128", dict(two_sm=False, bn=128, splits=1)), ("1sm bn256", dict(two_sm=False, bn=256, splits=1)), ("2sm", dict(two_sm=True, splits=1))]:
            alts.append(f"{label} {timeit(lambda: fn(**o)):.0f}")
    line = f"| {name} | {M},{N},{K} | {t0:.1f} | {2*M*N*K/t0/1e6:.0f} | {'; '.join(alts)} |"
    lines.append(line); print(line, flush=True)
os.makedirs("gpurun_out", exist_ok=True)
open("gpurun_out/gemm_shapes.md", "w").write("# GEMMs of one decoder layer (default model, 32000 tokens, TP=1), L2 flushed between iterations\n\n" + "\n".join(lines) + "\n")
