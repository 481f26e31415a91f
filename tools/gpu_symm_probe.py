"""Probe torch symmetric memory on the box: peer pointers, multicast pointer, signal pads (run under torchrun)."""
import os
import torch
import torch.distributed as dist

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
import torch.distributed._symmetric_memory as symm_mem
try:
    t = symm_mem.empty(1 << 20, dtype=torch.float32, device=f"cuda:{local}")
    hdl = symm_mem.rendezvous(t, dist.group.WORLD.group_name)
    if rank == 0:
        print("buffer_ptrs", [hex(p) for p in hdl.buffer_ptrs])
        print("signal_pad_ptrs", [hex(p) for p in hdl.signal_pad_ptrs])
        print("multicast_ptr", hex(hdl.multicast_ptr) if hdl.multicast_ptr else hdl.multicast_ptr)
        print("buffer_size", hdl.buffer_size, "signal_pad_size", hdl.signal_pad_size, "world", hdl.world_size)
        print("attrs", [a for a in dir(hdl) if not a.startswith("_")])
    t.fill_(rank + 1)
    hdl.barrier()
    peer = hdl.get_buffer((rank + 1) % world, (4,), torch.float32)
    print(rank, "peer read", peer.tolist())
    hdl.barrier()
except Exception as e:
    import traceback; traceback.print_exc()
p2p = [torch.cuda.can_device_access_peer(local, j) for j in range(torch.cuda.device_count()) if j != local]
if rank == 0:
    print("can_device_access_peer", p2p, "device_count", torch.cuda.device_count())
dist.barrier()
dist.destroy_process_group()
