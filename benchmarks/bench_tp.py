#!/usr/bin/env python
"""Tensor-parallel benchmark (BASELINE.json config 4: hidden 8192 over 8 GPUs).

    torchrun --nproc-per-node 8 benchmarks/bench_tp.py --hidden 8192 --names 8192

Reports names/s (device-timed, max over ranks) and the fused all-gather->GEMM
roofline: the slower of (MMA FLOPs per rank / measured bf16 peak) and (bytes
each rank must push over NVLink / measured 770 GB/s per direction).
"""
import argparse
import json
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

import torch
import torch.distributed as dist

from gru_mpi_cuda_b200 import GRUConfig
from gru_mpi_cuda_b200.parallel.tp import TPContext
from gru_mpi_cuda_b200.utils import weights as W
from gru_mpi_cuda_b200.utils.clocks import ClockSampler


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--hidden", type=int, default=8192)
    ap.add_argument("--names", type=int, default=8192)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    a = ap.parse_args()
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("cpu:gloo,cuda:nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    cfg = GRUConfig(hidden=a.hidden)
    T = cfg.max_len
    params = W.random_init(cfg, 1)
    tp = TPContext(cfg, params, local, a.names)
    del params
    rf = W.random_floats(a.names * T, 2, pin=True) if rank == 0 else None
    tp.generate(a.names, rf)                      # stages the floats on the root device
    for _ in range(a.warmup):
        tp.generate(a.names, None, to_host=False)
    torch.cuda.synchronize(); dist.barrier()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.steps)]
    with ClockSampler(local, period_ms=20) as cs:
        for s, e in evs:
            s.record(); tp.generate(a.names, None, to_host=False); e.record()
        torch.cuda.synchronize(); dist.barrier()
    ms = torch.tensor([sum(s.elapsed_time(e) for s, e in evs) / a.steps], dtype=torch.float64, device="cuda")
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    tp.gen.set_trace(True)
    tp.generate(a.names, None, to_host=False)
    tr = tp.gen.trace()
    tp.gen.set_trace(False)
    out = tp.generate(a.names, rf)
    if rank == 0:
        agg = {}
        for name, t in tr:
            agg.setdefault(name, []).append(t)
        print("trace(ms): " + "; ".join(f"{k}: n={len(v)} mean={sum(v)/len(v):.3f} first={v[0]:.3f} max={max(v):.3f}"
                                         for k, v in agg.items()), file=sys.stderr)
        H, C, B = cfg.hidden, cfg.num_char, a.names
        Hs = H // world
        kx = (C + 63) // 64 * 64
        full = 2.0 * B * 3 * Hs * ((kx + H) + 2 * H) + 2.0 * B * C * H          # per rank, steps >= 1
        first = 2.0 * B * 3 * Hs * (kx + H) + 2.0 * B * C * H                   # step 0: h == 0
        flops = full * (T - 1) + first
        nvl_bytes = 2.0 * B * Hs * 2 * (world - 1) * T                           # 2 layers, bf16, to 7 peers
        peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").exists() else {}
        peak = peaks.get("bf16_tflops_sustained", 1388.2) * 1e12
        t_compute = flops / peak * 1e3
        t_link = nvl_bytes / 770e9 * 1e3
        print(json.dumps({
            "bench": "tensor_parallel", "hidden": H, "tp": world, "names": B, "ms_per_generate": ms.item(),
            "names_per_sec": B / (ms.item() / 1e3), "executed_tflops_per_gpu": flops / (ms.item() / 1e3) / 1e12,
            "nvlink_out_gb_per_gpu_per_generate": nvl_bytes / 1e9,
            "roofline_ms": {"compute_at_measured_sustained_bf16": t_compute, "nvlink_at_measured_770GBs": t_link,
                            "bound": max(t_compute, t_link)},
            "achieved_fraction_of_roofline": max(t_compute, t_link) / ms.item(),
            "clocks": cs.summary(), "launches_per_generate": tp.gen.last_launches,
            "checksum": int(out.to(torch.int64).sum().item())}))
    tp.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
