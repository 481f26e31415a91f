"""All-reduce data-path comparison at the current world size: P2P two-shot vs NVLS (multimem) two-shot vs NCCL."""
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
    for _ in range(4): fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item() * 1e3
lines = []
for shape in [(4000, 512), (32000, 512), (32000, 2048)]:
    S = shape[0] * shape[1] * 2
    buf = symm.buffer(shape, torch.bfloat16); buf.normal_()
    res = {}
    for mode in ("p2p", "nvls"):
        symm.mode = mode
        res[mode] = timeit(lambda: symm.all_reduce(buf, out=buf))
    y = torch.randn(shape, device="cuda").to(torch.bfloat16)
    res["nccl"] = timeit(lambda: dist.all_reduce(y))
    # sequence-parallel halves
    part = symm.buffer(shape, torch.bfloat16)
    loc = torch.randn(shape[0] // world, shape[1], device="cuda").to(torch.bfloat16)
    for mode in ("p2p", "nvls"):
        symm.mode = mode
        res["rs_" + mode] = timeit(lambda: symm.reduce_scatter_rows(part))
        res["ag_" + mode] = timeit(lambda: symm.all_gather_into_buffer(loc))
    line = (f"TP={world} bf16 {shape} ({S/1e6:.1f} MB): all-reduce P2P {res['p2p']:.1f} us, NVLS {res['nvls']:.1f} us, "
            f"NCCL {res['nccl']:.1f} us | reduce-scatter P2P {res['rs_p2p']:.1f} / NVLS {res['rs_nvls']:.1f} us, "
            f"all-gather P2P {res['ag_p2p']:.1f} / NVLS {res['ag_nvls']:.1f} us")
    lines.append(line)
    if rank == 0: print(line, flush=True)
if rank == 0:
    os.makedirs("gpurun_out", exist_ok=True)
    open(f"gpurun_out/ar_modes_tp{world}.txt", "w").write("\n".join(lines) + "\n")
dist.barrier(); dist.destroy_process_group()
