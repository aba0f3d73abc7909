import os, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem
local = int(os.environ["LOCAL_RANK"]); torch.cuda.set_device(local)
dist.init_process_group("cpu:gloo,cuda:nccl", device_id=torch.device("cuda", local))
rank, world = dist.get_rank(), dist.get_world_size()
try:
    print(rank, "backend", symm_mem.get_backend(torch.device("cuda", local)) if hasattr(symm_mem, "get_backend") else "?", flush=True)
except Exception as e:
    print(rank, "get_backend err", e, flush=True)
t = symm_mem.empty(1 << 20, dtype=torch.uint8, device=torch.device("cuda", local))
hdl = symm_mem.rendezvous(t, dist.group.WORLD)
print(rank, "ptrs", [hex(p) for p in hdl.buffer_ptrs], "mc", hex(hdl.multicast_ptr) if hdl.multicast_ptr else None,
      "sig", [hex(p) for p in hdl.signal_pad_ptrs][:2], "size", hdl.buffer_size, flush=True)
dist.barrier()
# write through peer pointer sanity
t.fill_(rank + 1)
torch.cuda.synchronize(); dist.barrier()
peer = hdl.get_buffer((rank + 1) % world, (16,), torch.uint8)
print(rank, "peer sees", peer[:4].tolist(), flush=True)
dist.barrier()
dist.destroy_process_group()
