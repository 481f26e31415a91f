"""Launch each flagship kernel a few times (for `ncu -k ... -c 1` captures of the default-model shapes)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_pytorch_from_scratch_b200.ops import attention as A
from distributed_pytorch_from_scratch_b200.ops.gemm import gemm_nt
which = sys.argv[1] if len(sys.argv) > 1 else "all"
dev = "cuda"
torch.manual_seed(0)
if which in ("all", "attn"):
    B, T, H, HD = 32, 1000, 8, 64
    qkv = torch.randn(B * T, 3 * H * HD, device=dev).to(torch.bfloat16)
    for _ in range(3):
        out, lse = A.attn_fwd(qkv, B, T, H, HD, HD ** -0.5)
        g = torch.randn_like(out)
        A.attn_bwd(qkv, out, g, lse, B, T, H, HD, HD ** -0.5)
if which in ("all", "gemm"):
    M, N, K = 32000, 4096, 512     # gate|up forward of the default model
    a = torch.randn(M, K, device=dev).to(torch.bfloat16); w = torch.randn(N, K, device=dev).to(torch.bfloat16)
    bias = torch.randn(N, device=dev)
    for _ in range(3):
        gemm_nt(a, w, bias=bias, two_sm=True, splits=1)
    dy = torch.randn(M, 1536, device=dev).to(torch.bfloat16)   # wgrad of q|k|v: [1536, 512] over 32000 tokens
    x = torch.randn(M, 512, device=dev).to(torch.bfloat16)
    for _ in range(3):
        gemm_nt(dy, x, a_mn_major=True, b_mn_major=True, out_dtype=torch.float32, two_sm=True)
torch.cuda.synchronize()
print("done")
