"""NVLink P2P write bandwidth for the patterns the TP epilogue produces.
   torchrun --nproc-per-node G tools/p2p_bench.py"""
import ctypes, os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch, torch.distributed as dist
from gru_mpi_cuda_b200 import _native
from gru_mpi_cuda_b200.parallel.symm import SymmBuffer

local = int(os.environ["LOCAL_RANK"]); torch.cuda.set_device(local)
dist.init_process_group("cpu:gloo,cuda:nccl", device_id=torch.device("cuda", local))
rank, world = dist.get_rank(), dist.get_world_size()
lib = _native.lib()
NB = 256 << 20
USE_TORCH = os.environ.get("P2P_TORCH_SYMM") == "1"
if USE_TORCH:
    import torch.distributed._symmetric_memory as symm_mem
    t_sym = symm_mem.empty(NB, dtype=torch.uint8, device=torch.device("cuda", local))
    hdl = symm_mem.rendezvous(t_sym, dist.group.WORLD.group_name)
    class _B: pass
    buf = _B(); buf.ptrs = list(hdl.buffer_ptrs); buf.close = lambda: None
else:
    buf = SymmBuffer(NB, local)
src = torch.empty(NB, dtype=torch.uint8, device="cuda")
stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

def run(peers, row_bytes, pitch, blocks, who="all"):
    rows = (NB // pitch) if pitch else 0
    total = rows * row_bytes * len(peers)
    arr = (ctypes.c_void_p * len(peers))(*[ctypes.c_void_p(buf.ptrs[p]) for p in peers])
    active = who == "all" or rank == 0
    torch.cuda.synchronize(); dist.barrier()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for it in range(3):
        if it == 1: s.record()
        if active:
            _native.check(lib.b200gru_p2p_write_bench(arr, len(peers), src.data_ptr(), rows, row_bytes, pitch, blocks, stream))
    e.record(); torch.cuda.synchronize()
    ms = s.elapsed_time(e) / 2
    gbs = total / ms / 1e6
    t = torch.tensor([gbs], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MIN if who == "all" else dist.ReduceOp.MAX)
    return t.item()

others = [p for p in range(world) if p != rank]
for who in ("all",):
    for row_bytes, pitch in ((4096, 4096), (128, 4096), (128, 16384), (1024, 16384), (2048, 16384)):
        for blocks in (296,):
            g1 = run([(rank + 1) % world], row_bytes, pitch, blocks, who)
            gall = run(others, row_bytes, pitch, blocks, who)
            if rank == 0:
                print(f"writers={who:3s} row={row_bytes:5d}B pitch={pitch:6d} blocks={blocks:4d}: 1 peer {g1:7.1f} GB/s | {len(others)} peers total out {gall:7.1f} GB/s", flush=True)
dist.barrier(); buf.close(); dist.destroy_process_group()
