#!/usr/bin/env python
"""Headline benchmark: generated names/sec, whole job, device-timed, max over ranks.

    python bench.py --gpus N --steps K --warmup W            # this framework
    python bench.py --impl reference --gpus N ...            # unmodified reference

Config (BASELINE.json): the reference GRU name generator at its default shape
(NUM_CHAR=256, EMB=512, HIDDEN=1024, 2 layers, MAX_LEN=10), data-parallel over
N GPUs, random-init weights, synthetic uniform random floats.  One "step" is
one namegen() call: generate every name of the batch, all MAX_LEN characters,
including sampling and the output write.  Weak scaling: names per GPU fixed.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "generated names/sec (whole box, device-timed, max over ranks)"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--names-per-gpu", type=int, default=65536)
    ap.add_argument("--ref-names-per-gpu", type=int, default=32,
                    help="names per rank for the reference arm (its cost is strictly per-name serial)")
    ap.add_argument("--mode", default="tc", choices=["tc", "exact"])
    ap.add_argument("--dp", default="p2p", choices=["p2p", "nccl"])
    ap.add_argument("--seed", type=int, default=1)
    return ap.parse_args()


def env_ranks():
    return (int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)),
            int(os.environ.get("LOCAL_RANK", 0)))


# ------------------------------------------------------------------ reference
def run_reference(args):
    rank, world, local = env_ranks()
    exe = ROOT / "baseline" / "_ref" / "namegen_ref"
    if not exe.exists():
        try:
            sys.path.insert(0, str(ROOT / "baseline"))
            import build_ref
            build_ref.build(verbose=False)
        except Exception as e:  # noqa: BLE001
            if rank == 0:
                print(json.dumps({"impl": "reference", "unavailable": f"reference build failed: {e}"[:200]}))
            return 0
    if not exe.exists():
        if rank == 0:
            print(json.dumps({"impl": "reference",
                              "unavailable": "baseline/_ref/namegen_ref missing and /root/reference not present to build it"}))
        return 0
    from gru_mpi_cuda_b200 import GRUConfig
    from gru_mpi_cuda_b200.utils import weights as W
    cfg = GRUConfig()
    wfile = Path(f"/tmp/b200gru_ref_weights_{os.environ.get('MASTER_PORT', 'solo')}.bin")
    if rank == 0:
        W.save(wfile, W.random_init(cfg, args.seed))
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("gloo")
        dist.barrier()
    n_total = args.ref_names_per_gpu * world
    env = dict(os.environ, RANK=str(rank), WORLD_SIZE=str(world), CUDA_VISIBLE_DEVICES=str(local),
               B200GRU_SHM_TAG=os.environ.get("MASTER_PORT", "solo"))
    from gru_mpi_cuda_b200.utils.clocks import ClockSampler
    t0 = time.time()
    with ClockSampler(local) as cs:
        res = subprocess.run([str(exe), str(wfile), str(n_total), str(args.seed + 1), str(args.steps),
                              str(args.warmup)], env=env, capture_output=True, text=True)
    wall = time.time() - t0
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return 0
    if res.returncode != 0:
        print(json.dumps({"impl": "reference", "unavailable": f"reference run failed rc={res.returncode}: "
                          + (res.stderr.strip().splitlines() or ['?'])[-1][:160]}))
        return 0
    line = [l for l in res.stdout.splitlines() if l.startswith("{")][-1]
    r = json.loads(line)
    out = {
        "impl": "reference", "metric": METRIC, "value": r["names_per_sec"], "unit": "names/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp32",
        "data": "synthetic (random-init weights, uniform random floats)",
        "config": {"model": "gru-namegen C256 E512 H1024 L2 T10", "global_batch": r["generated"], "seq_len": cfg.max_len,
                   "parallelism": f"dp{world}", "names_per_gpu": args.ref_names_per_gpu,
                   "note": "unmodified namegensf.cu built against baseline/shim (no MPI in image: shared-memory "
                           "shim); batch is small because the reference generates one name at a time, so its "
                           "names/s does not depend on N; timing = wall clock between cudaDeviceSynchronize fences"},
        "clocks": cs.summary(), "wall_s": round(wall, 2),
        "e2e": {"value": r["names_per_sec"], "unit": "names/s", "h2d_bytes_per_step": None,
                "d2h_bytes_per_step": None, "note": "reference copies 4 B H2D + 1 KiB D2H per character itself"},
        "gpu_launches": None,
    }
    print(json.dumps(out))
    return 0


# ----------------------------------------------------------------------- ours
def run_ours(args):
    import torch
    import torch.distributed as dist
    from gru_mpi_cuda_b200 import GRUConfig
    from gru_mpi_cuda_b200.models import NameGenerator
    from gru_mpi_cuda_b200.utils import weights as W
    from gru_mpi_cuda_b200.utils.clocks import ClockSampler

    rank, world, local = env_ranks()
    if world != args.gpus and rank == 0:
        print(f"[bench] note: --gpus {args.gpus} but WORLD_SIZE={world}; using WORLD_SIZE", file=sys.stderr)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    comm = None
    if world > 1:
        dist.init_process_group("cpu:gloo,cuda:nccl", device_id=dev)
        from gru_mpi_cuda_b200.parallel.dp import DPContext
    cfg = GRUConfig()
    per_gpu = args.names_per_gpu
    n_total = per_gpu * world
    T = cfg.max_len
    params = W.random_init(cfg, args.seed)
    gen = NameGenerator(cfg, params, device=local, max_batch=per_gpu, mode=args.mode)
    if world > 1:
        comm = DPContext(cfg, local, n_total, method=args.dp)
        gen.comm = comm
    rf_host = W.random_floats(n_total * T, args.seed + 1, pin=True) if rank == 0 else None
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def device_step():
        """Device-resident inputs: kernels (+ in-kernel P2P traffic and flags) only."""
        if comm is None:
            gen.generate_device(gen._rf_dev, per_gpu)
        else:
            comm.generate(gen, n_total, None, to_host=False)

    # stage the floats on the device once for the kernel-only measurement
    if comm is None:
        gen._rf_dev[: n_total * T].copy_(rf_host)
    else:
        comm.generate(gen, n_total, rf_host, to_host=True)
    barrier()

    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    # the sampler starts before warm-up (nvidia-smi takes ~100 ms to come up); its
    # summary keeps the samples taken under load
    with ClockSampler(local, period_ms=20) as cs:
        time.sleep(0.3)
        for _ in range(args.warmup):
            flush.fill_(1)
            device_step()
        barrier()
        launches_per_step = gen.last_launches
        t_wall0 = time.perf_counter()
        t_epoch0 = time.time()
        for a, b in evs:
            flush.fill_(2)          # evict L2 between timed iterations (outside the event pair)
            a.record()
            device_step()
            b.record()
        barrier()
        t_wall = time.perf_counter() - t_wall0
        t_epoch1 = time.time()
        time.sleep(0.05)  # let the sampler flush its last rows
    ms_local = sum(a.elapsed_time(b) for a, b in evs)
    ms_t = torch.tensor([ms_local], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms_t, op=dist.ReduceOp.MAX)
    ms_total = ms_t.item()
    clocks = cs.summary(t_epoch0, t_epoch1)

    # ---- end to end through the public API: pinned host floats in, host bytes out
    # like the reference's caller-allocated `output` (namegensf.cu:624-627): a pinned host buffer
    out_host = torch.empty(n_total, T + 1, dtype=torch.uint8).pin_memory() if (rank == 0 and comm is None) else None
    e2e_kw = {"out": out_host} if comm is None else {}
    for _ in range(2):
        gen.generate(n_total, rf_host, **e2e_kw)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out = gen.generate(n_total, rf_host, **e2e_kw)
    barrier()
    e2e_s = time.perf_counter() - t0
    e2e_t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_s = e2e_t.item()

    if rank == 0:
        ms_per_step = ms_total / args.steps
        value = n_total / (ms_per_step / 1e3)
        flops_per_name_step = 2.0 * (3 * cfg.hidden * cfg.hidden * (2 * cfg.num_layers - 1)
                                     + cfg.hidden * cfg.num_char)
        # step 0 has h == 0: the recurrent GEMMs are skipped there
        flops_per_name = flops_per_name_step * T - 2.0 * 3 * cfg.hidden * cfg.hidden * cfg.num_layers
        peaks = {}
        try:
            peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
        except Exception:  # noqa: BLE001
            pass
        tflops = per_gpu * flops_per_name / (ms_per_step / 1e3) / 1e12
        res = {
            "metric": METRIC, "value": value, "unit": "names/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16" if args.mode == "tc" else "fp32",
            "data": "synthetic (random-init weights of the named shape, uniform random floats)",
            "config": {"model": "gru-namegen C256 E512 H1024 L2 T10", "global_batch": n_total,
                       "seq_len": T, "parallelism": f"dp{world}" + (f"-{args.dp}" if world > 1 else ""),
                       "names_per_gpu": per_gpu, "kernel_family": args.mode,
                       "l2": "256 MiB flush write between timed iterations; per-step working set "
                             f"(hidden state {4 * per_gpu * cfg.hidden * 2 >> 20} MiB + weights) also exceeds L2",
                       "step": "one namegen() call: all names, all MAX_LEN characters, sampling and output write"},
            "clocks": clocks,
            "e2e": {"value": n_total / (e2e_s / args.steps), "unit": "names/s",
                    "h2d_bytes_per_step": n_total * T * 4, "d2h_bytes_per_step": n_total * (T + 1),
                    "ms_per_step": e2e_s / args.steps * 1e3},
            "gpu_launches": launches_per_step * args.steps,
            "launches_per_step": launches_per_step,
            "achieved_tflops_per_gpu": tflops,
            "frac_of_measured_bf16_sustained": (tflops / peaks["bf16_tflops_sustained"]) if peaks.get("bf16_tflops_sustained") else None,
            "wall_s_timed_region": t_wall,
            "sample_names_checksum": int(out.to(torch.int64).sum().item()),
        }
        print(json.dumps(res))
    if comm is not None:
        barrier()
        comm.close()
    gen.finalize()
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
