"""Launch the memory-bound kernels of one decoder layer at the default-model shapes (for `ncu -k regex:... -c N`)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_pytorch_from_scratch_b200.ops import fused
M, F, D = 32000, 2048, 512
torch.manual_seed(0)
gu = torch.randn(M, 2 * F, device="cuda").to(torch.bfloat16)
dact = torch.randn(M, F, device="cuda").to(torch.bfloat16)
bg = torch.empty(2 * F, device="cuda")
x = torch.randn(M, D, device="cuda")
dy = torch.randn(M, D, device="cuda").to(torch.bfloat16)
scale = torch.rand(D, device="cuda") + 0.5
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for _ in range(3):
    flush.zero_()
    act = fused.swiglu_fwd(gu)
    flush.zero_()
    dgu = fused.swiglu_bwd_bias(dact, gu, bg)
    flush.zero_()
    h, rstd = fused.rmsnorm_fwd(x, scale, 1e-5)
    flush.zero_()
    fused.rmsnorm_bwd(dy, x, rstd, scale, dres=x, want_fp32=True, want_bf16=True)
torch.cuda.synchronize()
print("done")
