import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
import process_manager as pm
pm.init_pgm(world)
from distributed_pytorch_from_scratch_b200.parallel.symm import try_enable_symm
from distributed_pytorch_from_scratch_b200.ops import _lib
symm = try_enable_symm(required=True)
def timeit(fn, iters=30):
    for _ in range(5): fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item() * 1e3
NAMES = {1: "ld_reduce only", 2: "multimem.st only", 3: "p2p ld + multimem.st", 4: "p2p ld + p2p st"}
for shape in [(32000, 512), (32000, 2048)]:
    S = shape[0] * shape[1] * 2
    x = torch.randn(shape, device="cuda").to(torch.bfloat16)
    ref = x.clone(); dist.all_reduce(ref)
    buf = symm.buffer(shape, torch.bfloat16)
    for mode in (1, 2, 3, 4):
        for blocks in (32, 128):
            def run():
                rc = _lib.lib().dpfs_allreduce_exp(symm.mc_base, symm.data_ptrs, 0, S, mode, blocks, symm.flag_ptrs,
                                                   symm.rank, symm.n, _lib.stream())
                assert rc == 0, rc
            buf.copy_(x); run(); torch.cuda.synchronize()
            err = (buf.float() - ref.float()).abs().max().item() if mode >= 3 else float("nan")
            t = timeit(run)
            if rank == 0:
                print(f"{shape} mode {mode} ({NAMES[mode]}) blocks {blocks}: {t:.1f} us ({S/t/1e3:.0f} GB/s algbw) err {err:.3g}", flush=True)
dist.barrier(); dist.destroy_process_group()
