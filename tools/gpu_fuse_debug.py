import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
import process_manager as pm
pm.init_pgm(world)
from distributed_pytorch_from_scratch_b200.parallel.symm import try_enable_symm
symm = try_enable_symm(required=True)
def timeit(fn, iters=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item() * 1e3
M, N, K = 32000, 512, 256
a = torch.randn(M, K, device="cuda").to(torch.bfloat16); w = torch.randn(N, K, device="cuda").to(torch.bfloat16)
from distributed_pytorch_from_scratch_b200.ops.gemm import gemm_nt
o1 = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
o2 = symm.buffer((M, N), torch.bfloat16)
t1 = timeit(lambda: gemm_nt(a, w, out=o1))
t2 = timeit(lambda: gemm_nt(a, w, out=o2))
t3 = timeit(lambda: o1.copy_(o2))
t4 = timeit(lambda: o2.copy_(o1))
if rank == 0: print(f"DEBUG mc={hex(symm.mc_base)} gemm->normal {t1:.1f} us, gemm->symm {t2:.1f} us, copy symm->normal {t3:.1f}, normal->symm {t4:.1f}", flush=True)
t = timeit(lambda: symm.gemm_allreduce(a, w))
if rank == 0: print(f"DEBUG={os.environ.get('DPFS_FUSE_DEBUG')} MODE={os.environ.get('DPFS_AR_MODE')}: fused {t:.1f} us", flush=True)
dist.barrier(); dist.destroy_process_group()
