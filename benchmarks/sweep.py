#!/usr/bin/env python
"""names-per-call sweep (BASELINE.json config 5): 1K..1M names at G GPUs, DP.

    python benchmarks/sweep.py                              # 1 GPU
    torchrun --nproc-per-node 8 benchmarks/sweep.py         # 8 GPUs

For each total N prints device-timed names/s (max over ranks) and the end-to-end
number through NameGenerator.generate (pinned host floats in, host bytes out).
"""
import argparse
import json
import os
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

import torch
import torch.distributed as dist

from gru_mpi_cuda_b200 import GRUConfig
from gru_mpi_cuda_b200.models import NameGenerator
from gru_mpi_cuda_b200.utils import weights as W


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="1024,4096,16384,65536,262144,1048576")
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--dp", default="p2p")
    ap.add_argument("--path", default="auto", choices=["auto", "multi", "persistent"],
                    help="tensor-core kernel path (auto: persistent kernel for small calls)")
    ap.add_argument("--persist-max", type=int, default=-1, help="auto: largest call (names per GPU) for the persistent kernel")
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("cpu:gloo,cuda:nccl", device_id=torch.device("cuda", local))
        from gru_mpi_cuda_b200.parallel.dp import DPContext
    cfg = GRUConfig()
    T = cfg.max_len
    sizes = [int(x) for x in a.sizes.split(",")]
    nmax = max(sizes)
    gen = NameGenerator(cfg, W.random_init(cfg, 1), device=local, max_batch=(nmax + world - 1) // world, mode="tc")
    gen.set_path(a.path, a.persist_max)
    comm = None
    if world > 1:
        comm = DPContext(cfg, local, nmax, method=a.dp)
        gen.comm = comm
    rf = W.random_floats(nmax * T, 2, pin=True) if rank == 0 else None
    rows = []
    for n in sizes:
        def dev_step():
            if comm is None:
                gen.generate_device(gen._rf_dev, n)
            else:
                comm.generate(gen, n, None, to_host=False)
        gen.generate(n, rf)  # stage floats, warm
        for _ in range(2):
            dev_step()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(a.steps):
            dev_step()
        e.record()
        torch.cuda.synchronize()
        ms = torch.tensor([s.elapsed_time(e) / a.steps], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(a.steps):
            gen.generate(n, rf)
        torch.cuda.synchronize()
        e2e = torch.tensor([(time.perf_counter() - t0) / a.steps], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(e2e, op=dist.ReduceOp.MAX)
        if rank == 0:
            row = {"n_names": n, "gpus": world, "path": a.path, "launches": gen.last_launches, "ms_device": ms.item(), "names_per_sec_device": n / (ms.item() / 1e3),
                   "ms_e2e": e2e.item() * 1e3, "names_per_sec_e2e": n / e2e.item()}
            rows.append(row)
            print(json.dumps(row), flush=True)
    if comm is not None:
        torch.cuda.synchronize(); dist.barrier(); comm.close()
    gen.finalize()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
