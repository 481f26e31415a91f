"""GPU bring-up script for the tcgen05 GEMM: runs every operand-major variant over a ladder of shapes, each variant
in its own subprocess with a timeout (a hung kernel must not hide the results of the other variants).

    python tools/gpu_gemm_debug.py            # writes gpurun_out/gemm_debug.log
"""
import os
import subprocess
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

VARIANTS = {"kk": (False, False), "kmn": (False, True), "mnmn": (True, True), "mnk": (True, False)}
SHAPES = [
    # M, N, K
    (128, 128, 64), (128, 128, 128), (128, 128, 512), (256, 256, 256), (128, 256, 64),
    (1000, 512, 512), (4096, 1536, 512), (32000, 512, 2048), (520, 264, 200),
]


def run_variant(name):
    import torch
    from distributed_pytorch_from_scratch_b200.ops.gemm import gemm_nt
    a_mn, b_mn = VARIANTS[name]
    torch.manual_seed(0)
    dev = "cuda"
    for (M, N, K) in SHAPES:
        if (a_mn and M % 8) or (b_mn and N % 8) or ((not a_mn or not b_mn) and K % 8):
            continue
        for bn in ((128, 256, 512) if not os.environ.get("ONLY_2SM") else (512,)):
            for out_dtype in (torch.bfloat16, torch.float32):
                A = torch.randn(M, K, device=dev).to(torch.bfloat16)
                B = torch.randn(N, K, device=dev).to(torch.bfloat16)
                a_in = A.t().contiguous() if a_mn else A
                b_in = B.t().contiguous() if b_mn else B
                bias = torch.randn(N, device=dev)
                res = torch.randn(M, N, device=dev)
                ref = A.float() @ B.float().t()
                t0 = time.time()
                kw = dict(two_sm=True) if bn == 512 else dict(bn=bn, two_sm=False)
                out = gemm_nt(a_in, b_in, a_mn_major=a_mn, b_mn_major=b_mn, out_dtype=out_dtype, splits=1, **kw)
                torch.cuda.synchronize()
                err = (out.float() - ref).abs().max().item()
                scale = ref.abs().max().item()
                out2 = gemm_nt(a_in, b_in, a_mn_major=a_mn, b_mn_major=b_mn, out_dtype=out_dtype,
                               bias=bias, residual=res, splits=1, **kw)
                torch.cuda.synchronize()
                err2 = (out2.float() - (ref + bias + res)).abs().max().item()
                out3 = gemm_nt(a_in, b_in, a_mn_major=a_mn, b_mn_major=b_mn, out_dtype=out_dtype,
                               bias=bias, splits=3, **kw)
                torch.cuda.synchronize()
                err3 = (out3.float() - (ref + bias)).abs().max().item()
                tol = (0.02 if out_dtype == torch.bfloat16 else 2e-3) * max(scale, 1.0)
                ok = err < tol and err2 < tol and err3 < tol
                print(f"[{name}] M={M} N={N} K={K} bn={bn} out={str(out_dtype)[6:]} err={err:.4g} err_epi={err2:.4g} "
                      f"err_splitk={err3:.4g} scale={scale:.3g} {'OK' if ok else 'FAIL'} ({time.time()-t0:.2f}s)",
                      flush=True)
                if not ok:
                    d = (out.float() - ref).abs()
                    bad_rows = (d.max(dim=1).values > tol).nonzero().flatten()[:16].tolist()
                    bad_cols = (d.max(dim=0).values > tol).nonzero().flatten()[:16].tolist()
                    print(f"    bad rows (first 16): {bad_rows}  bad cols: {bad_cols} "
                          f"frac_bad={(d > tol).float().mean().item():.4f}", flush=True)
                    print("    out[0,:8]=", out[0, :8].float().tolist(), "\n    ref[0,:8]=", ref[0, :8].tolist(),
                          flush=True)
                    # one-hot K probe on the first shape: which k does each k0 land on?
                    if (M, N, K) == SHAPES[0]:
                        Bp = torch.arange(N * K, device=dev).reshape(N, K).remainder(251).float().to(torch.bfloat16)
                        for k0 in (0, 1, 8, 16, 17, 63):
                            Ap = torch.zeros(M, K, device=dev, dtype=torch.bfloat16)
                            Ap[:, k0] = 1
                            o = gemm_nt(Ap.t().contiguous() if a_mn else Ap, Bp.t().contiguous() if b_mn else Bp,
                                        a_mn_major=a_mn, b_mn_major=b_mn, out_dtype=torch.float32, bn=bn, splits=1)
                            torch.cuda.synchronize()
                            match = [(o[0] - Bp[:, kk].float()).abs().max().item() < 0.5 for kk in range(K)]
                            print(f"    one-hot k0={k0}: row0 matches B[:,k] for k in "
                                  f"{[i for i, m in enumerate(match) if m][:8]}; "
                                  f"rows equal to row0: {(o == o[0:1]).all(dim=1).sum().item()}/{M}", flush=True)
    # quick timing of the flagship shapes
    for (M, N, K) in [(32000, 1536, 512), (32000, 512, 2048), (32000, 4096, 512), (8192, 8192, 8192)]:
        for bn in (128, 256, 512):
            kw = dict(two_sm=True) if bn == 512 else dict(bn=bn, two_sm=False)
            A = torch.randn(M, K, device=dev).to(torch.bfloat16)
            B = torch.randn(N, K, device=dev).to(torch.bfloat16)
            a_in = A.t().contiguous() if a_mn else A
            b_in = B.t().contiguous() if b_mn else B
            for _ in range(3):
                gemm_nt(a_in, b_in, a_mn_major=a_mn, b_mn_major=b_mn, splits=1, **kw)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10):
                gemm_nt(a_in, b_in, a_mn_major=a_mn, b_mn_major=b_mn, splits=1, **kw)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 10
            print(f"[{name}] time M={M} N={N} K={K} bn={bn}: {ms*1e3:.1f} us  {2*M*N*K/ms/1e9:.1f} TFLOP/s",
                  flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--variant":
        run_variant(sys.argv[2])
        sys.exit(0)
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/gemm_debug.log", "w") as log:
        for name in VARIANTS:
            try:
                p = subprocess.run([sys.executable, __file__, "--variant", name], capture_output=True, text=True,
                                   timeout=120)
                txt = p.stdout + ("\nSTDERR:\n" + p.stderr[-3000:] if p.returncode != 0 else "")
                txt += f"\n[{name}] exit code {p.returncode}\n"
            except subprocess.TimeoutExpired as e:
                txt = (e.stdout.decode() if e.stdout else "") + f"\n[{name}] TIMEOUT\n"
            log.write(txt)
            log.flush()
            print(txt, flush=True)
