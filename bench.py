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
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--names-per-gpu", type=int, default=65536)
    ap.add_argument("--ref-names-per-gpu", type=int, default=32,
                    help="names per rank for the reference arm (its cost is strictly per-name serial)")
    ap.add_argument("--mode", default="tc", choices=["tc", "exact"])
    ap.add_argument("--dp", default="p2p", choices=["p2p", "nccl"])
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--exact-names", type=int, default=8192,
                    help="names per GPU for the same-precision (fp32 exact path) side measurement; 0 = skip")
    ap.add_argument("--parallel", default="dp", choices=["dp", "tp"],
                    help="dp: BASELINE config 2/3 (names split over GPUs); tp: config 4 (hidden split over GPUs)")
    ap.add_argument("--hidden", type=int, default=None, help="hidden size (tp arm default 8192)")
    ap.add_argument("--tp-names", type=int, default=8192, help="names per call on the tp arm")
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
    # Nothing of this repo's package or native library is imported on this arm:
    # the weight file is written with numpy and the clock sampler is loaded by
    # file path (stdlib only), so the only GPU code in these processes is the
    # reference binary itself.
    import importlib.util
    import numpy as np
    C, E, H, L, T = 256, 512, 1024, 2, 10     # the reference's default shape
    n_params = C * E + 3 * H * E + 3 * H * H * (2 * L - 1) + 6 * H * L + C * H + C
    wfile = Path(f"/tmp/b200gru_ref_weights_{os.environ.get('MASTER_PORT', 'solo')}.bin")
    if rank == 0:
        k = 1.0 / np.sqrt(H)
        np.random.default_rng(args.seed).uniform(-k, k, n_params).astype("<f4").tofile(wfile)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("gloo")
        dist.barrier()
    n_total = args.ref_names_per_gpu * world
    env = dict(os.environ, RANK=str(rank), WORLD_SIZE=str(world), CUDA_VISIBLE_DEVICES=str(local),
               B200GRU_SHM_TAG=os.environ.get("MASTER_PORT", "solo"))
    spec = importlib.util.spec_from_file_location("_b200gru_clocks", ROOT / "gru_mpi_cuda_b200" / "utils" / "clocks.py")
    clocks_mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(clocks_mod)
    ClockSampler = clocks_mod.ClockSampler
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
        "config": {"model": "gru-namegen C256 E512 H1024 L2 T10", "global_batch": r["generated"], "seq_len": T,
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
    # L2 rule: the per-call working set (two hidden buffers per layer, bf16, every
    # layer kernel streams them once per character) is compared with 2x the 126 MB
    # L2; above that the inputs themselves exceed L2 and nothing is flushed, so the
    # timed region is kernels only (clock samples inside it are all under load).
    # Below it, a 256 MiB write evicts L2 between timed iterations.
    work_bytes = 2 * cfg.num_layers * per_gpu * cfg.hidden * 2
    use_flush = work_bytes < (2 * 126 << 20)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev) if use_flush else None

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
    with ClockSampler(local, period_ms=10) as cs:
        time.sleep(0.3)
        for _ in range(args.warmup):
            if use_flush:
                flush.fill_(1)
            device_step()
        barrier()
        launches_per_step = gen.last_launches
        t_wall0 = time.perf_counter()
        t_epoch0 = time.time()
        for a, b in evs:
            if use_flush:
                flush.fill_(2)      # evict L2 between timed iterations (outside the event pair)
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
    out_host = torch.empty(n_total, T + 1, dtype=torch.uint8).pin_memory() if rank == 0 else None
    e2e_kw = {"out": out_host} if rank == 0 else {}
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

    # ---- same precision as the reference (fp32, bit-identical to the CPU oracle):
    # the exact path on each GPU's own names, device-timed, kernels only
    exact = None
    if args.exact_names > 0 and args.mode == "tc":
        ne = min(args.exact_names, per_gpu)
        if comm is not None:     # ranks other than the root never staged floats locally
            gen._rf_dev[: ne * T].uniform_()
        for _ in range(2):
            gen.generate_device(gen._rf_dev, ne, mode="exact")
        barrier()
        ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 3
        ea.record()
        for _ in range(reps):
            gen.generate_device(gen._rf_dev, ne, mode="exact")
        eb.record()
        barrier()
        ex_t = torch.tensor([ea.elapsed_time(eb) / reps], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ex_t, op=dist.ReduceOp.MAX)
        exact = {"value": ne * world / (ex_t.item() / 1e3), "unit": "names/s", "dtype": "fp32",
                 "names_per_gpu": ne, "ms_per_step": ex_t.item(), "n_gpus": world,
                 "note": "mode='exact': fp32 SIMT kernels in the reference's association order, names "
                         "bit-identical to the CPU oracle; each GPU runs its own names (no host traffic)"}

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
                       "l2": ("256 MiB flush write between timed iterations" if use_flush else
                              f"inputs larger than L2: hidden state {work_bytes >> 20} MiB per GPU is streamed "
                              "once per character (no flush between iterations)"),
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
            "precision_note": "headline is bf16 tensor-core generation (BASELINE config 3); the reference arm is "
                              "fp32 at a small batch; same_precision_fp32 is this framework's fp32 exact path",
            "same_precision_fp32": exact,
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


# ------------------------------------------------------------ tensor parallel
def run_tp(args):
    """BASELINE.json config 4: hidden 8192 sharded over the GPUs of the node, every
    rank runs the same names; the per-step hidden all-gather is fused into the
    next GEMM's K loop (in-kernel NVLink stores + per-tile counters)."""
    import torch
    import torch.distributed as dist
    from gru_mpi_cuda_b200 import GRUConfig
    from gru_mpi_cuda_b200.models import NameGenerator
    from gru_mpi_cuda_b200.utils import weights as W
    from gru_mpi_cuda_b200.utils.clocks import ClockSampler

    rank, world, local = env_ranks()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    H = args.hidden or 8192
    cfg = GRUConfig(hidden=H)
    T, C, B = cfg.max_len, cfg.num_char, args.tp_names
    params = W.random_init(cfg, args.seed)
    rf_host = W.random_floats(B * T, args.seed + 1, pin=True) if rank == 0 else None
    if world > 1:
        dist.init_process_group("cpu:gloo,cuda:nccl", device_id=dev)
        from gru_mpi_cuda_b200.parallel.tp import TPContext
        tp = TPContext(cfg, params, local, B)
        gen = tp.gen
        step = lambda: tp.generate(B, None, to_host=False)          # noqa: E731
        e2e_step = lambda: tp.generate(B, rf_host)                   # noqa: E731
    else:
        tp = None
        gen = NameGenerator(cfg, params, device=local, max_batch=B, mode="tc")
        out_host = torch.empty(B, T + 1, dtype=torch.uint8).pin_memory()
        step = lambda: gen.generate_device(gen._rf_dev, B)           # noqa: E731
        e2e_step = lambda: gen.generate(B, rf_host, out=out_host)    # noqa: E731
    del params

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    e2e_step()                      # stages the floats on the (root) device
    barrier()
    # working set per call: 2 layers x 2 buffers x B x H bf16 + this rank's weight slice
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    with ClockSampler(local, period_ms=10) as cs:
        time.sleep(0.3)
        for _ in range(max(args.warmup, 3)):
            step()
        barrier()
        t_epoch0 = time.time()
        for a, b in evs:
            a.record(); step(); b.record()
        barrier()
        t_epoch1 = time.time()
        time.sleep(0.05)
    launches = gen.last_launches
    ms_t = torch.tensor([sum(a.elapsed_time(b) for a, b in evs) / args.steps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms_t, op=dist.ReduceOp.MAX)
    ms = ms_t.item()
    clocks = cs.summary(t_epoch0, t_epoch1)
    for _ in range(2):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out = e2e_step()
    barrier()
    e2e_t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_ms = e2e_t.item() / args.steps * 1e3
    if rank == 0:
        Hs = H // world
        full = 2.0 * B * 3 * Hs * (H + 2 * H) + 2.0 * B * C * H     # per rank, steps >= 1 (useful MACs only)
        first = 2.0 * B * 3 * Hs * H + 2.0 * B * C * H               # step 0: h == 0
        flops = full * (T - 1) + first
        nvl_bytes = 2.0 * B * Hs * 2 * T * (1 if world > 1 else 0)   # unique bytes a rank must publish (2 layers)
        peaks = {}
        try:
            peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
        except Exception:  # noqa: BLE001
            pass
        peak = peaks.get("bf16_tflops_sustained", 1388.2) * 1e12
        t_compute = flops / peak * 1e3
        t_link = nvl_bytes * max(world - 1, 0) / 770e9 * 1e3          # unicast egress: one copy per peer
        t_link_mc = nvl_bytes / 770e9 * 1e3                           # NVSwitch multicast: one copy out
        bound = max(t_compute, t_link_mc)
        print(json.dumps({
            "metric": METRIC, "value": B / (ms / 1e3), "unit": "names/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic (random-init weights of the named shape, uniform random floats)",
            "config": {"model": f"gru-namegen C256 E512 H{H} L2 T10", "global_batch": B, "seq_len": T,
                       "parallelism": f"tp{world}", "kernel_family": "tc",
                       "allgather": ("nvswitch multicast (multimem.st / multimem.red)" if (tp is not None and tp.multicast)
                                     else "unicast p2p stores" if tp is not None else "none"),
                       "l2": f"inputs larger than L2: {2 * 2 * B * H * 2 >> 20} MiB of hidden state + "
                             f"{3 * 3 * Hs * H * 2 >> 20} MiB of weights per rank streamed per character (no flush)",
                       "step": "one namegen() call on every rank: all names, all MAX_LEN characters"},
            "clocks": clocks,
            "e2e": {"value": B / (e2e_ms / 1e3), "unit": "names/s", "h2d_bytes_per_step": B * T * 4,
                    "d2h_bytes_per_step": B * (T + 1), "ms_per_step": e2e_ms},
            "gpu_launches": launches * args.steps, "launches_per_step": launches,
            "achieved_tflops_per_gpu": flops / (ms / 1e3) / 1e12,
            "roofline_ms": {"compute_at_measured_sustained_bf16": t_compute,
                            "nvlink_unicast_at_770GBs": t_link, "nvlink_multicast_at_770GBs": t_link_mc,
                            "bound": bound},
            "achieved_fraction_of_roofline": bound / ms,
            "sample_names_checksum": int(out.to(torch.int64).sum().item()) if out is not None else None,
        }))
    if tp is not None:
        tp.close()
        dist.destroy_process_group()
    else:
        gen.finalize()
    return 0


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)
    if args.parallel == "tp":
        return run_tp(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
