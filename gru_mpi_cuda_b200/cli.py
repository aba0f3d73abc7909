"""Command-line driver — the role of the reference's unshipped main.cpp
(`/root/reference/namegensf.cu:21`, `:625`): parse arguments, make the random
floats, call initialize / generate / finalize, time the call, print or save
the names.

    python -m gru_mpi_cuda_b200.cli --params model.bin --n 1000 --seed 7 --out names.bin
    torchrun --nproc-per-node 8 -m gru_mpi_cuda_b200.cli --random-init --n 1000000
    python -m gru_mpi_cuda_b200.cli --make-weights model.bin          # write random-init weights
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import torch

from . import GRUConfig
from .utils import weights as W


def _cfg(a) -> GRUConfig:
    return GRUConfig(num_char=a.num_char, emb_dim=a.emb_dim, hidden=a.hidden, num_layers=a.layers,
                     max_len=a.max_len, sos=a.sos, eos=a.eos)


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="gru_mpi_cuda_b200.cli", description=__doc__,
                                 formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--params", help="weight file (headerless fp32, reference layout)")
    ap.add_argument("--random-init", action="store_true", help="use deterministic random weights")
    ap.add_argument("--make-weights", metavar="PATH", help="write random-init weights to PATH and exit")
    ap.add_argument("--import-torch", nargs=2, metavar=("STATE.pt", "OUT.bin"),
                    help="convert a torch state dict (embedding/gru/fc keys) to the reference weight file and exit")
    ap.add_argument("--export-torch", nargs=2, metavar=("IN.bin", "STATE.pt"),
                    help="convert a reference weight file to a torch state dict and exit")
    ap.add_argument("--n", type=int, default=16, help="number of names (N)")
    ap.add_argument("--seed", type=int, default=0, help="seed of the random floats (rng_seed)")
    ap.add_argument("--out", help="write the N x (MAX_LEN+1) byte matrix here")
    ap.add_argument("--mode", default="tc", choices=["tc", "bf16", "exact", "fp32", "cpu"])
    ap.add_argument("--parallel", default="dp", choices=["dp", "tp"], help="multi-GPU strategy under torchrun")
    ap.add_argument("--dp-method", default="p2p", choices=["p2p", "nccl"])
    ap.add_argument("--print", type=int, default=8, dest="show", help="names to print")
    ap.add_argument("--json", action="store_true", help="print a one-line JSON result record")
    for name, default in (("num-char", 256), ("emb-dim", 512), ("hidden", 1024), ("layers", 2), ("max-len", 10),
                          ("sos", 0), ("eos", 1)):
        ap.add_argument(f"--{name}", type=int, default=default)
    a = ap.parse_args(argv)
    cfg = _cfg(a)

    if a.import_torch:
        src, dst = a.import_torch
        sd = torch.load(src, map_location="cpu", weights_only=True)
        W.save(dst, W.from_torch_state(cfg, sd))
        print(f"wrote {W.param_count(cfg) * 4} bytes to {dst}")
        return 0
    if a.export_torch:
        src, dst = a.export_torch
        torch.save(W.to_torch_state(cfg, W.load(src, cfg)), dst)
        print(f"wrote torch state dict to {dst}")
        return 0
    if a.make_weights:
        W.save(a.make_weights, W.random_init(cfg, a.seed))
        print(f"wrote {W.param_count(cfg) * 4} bytes to {a.make_weights}")
        return 0
    if not a.params and not a.random_init:
        ap.error("give --params FILE or --random-init")

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    params = W.load(a.params, cfg) if a.params else W.random_init(cfg, 1)
    # the reference prints one hello line per rank (:365-366)
    print(f"({os.uname().nodename}) Hello world, rank {rank} out of {world}", file=sys.stderr)
    rf = W.random_floats(a.n * cfg.max_len, a.seed, pin=torch.cuda.is_available()) if rank == 0 else None

    if a.mode == "cpu":
        from .models import oracle
        t0 = time.perf_counter()
        rows = oracle.generate(cfg, params, rf, n_names=a.n)
        dt = time.perf_counter() - t0
    else:
        from .models import NameGenerator
        torch.cuda.set_device(local)
        comm = gen = tp = None
        if world > 1:
            import torch.distributed as dist
            dist.init_process_group("cpu:gloo,cuda:nccl", device_id=torch.device("cuda", local))
        if world > 1 and a.parallel == "tp":
            from .parallel.tp import TPContext
            tp = TPContext(cfg, params, local, a.n)
            run = lambda: tp.generate(a.n, rf)
        else:
            per_rank = max(1, -(-a.n // world))
            gen = NameGenerator(cfg, params, device=local, max_batch=per_rank, mode=a.mode)
            if world > 1:
                from .parallel.dp import DPContext
                comm = DPContext(cfg, local, a.n, method=a.dp_method)
                gen.comm = comm
            run = lambda: gen.generate(a.n, rf)
        run()                                   # warm-up (module load, graph capture)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        rows = run()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if tp is not None:
            tp.close()
        if comm is not None:
            torch.cuda.synchronize(); dist.barrier(); comm.close()
        if gen is not None:
            gen.finalize()
        if world > 1:
            dist.destroy_process_group()
    if rank != 0:
        return 0
    if a.out:
        with open(a.out, "wb") as f:
            f.write(rows.numpy().tobytes())
    from .models.namegen import decode
    for i, name in enumerate(decode(rows[: a.show])):
        print(f"{i:6d}  {name!r}")
    rec = {"n_names": a.n, "ranks": world, "mode": a.mode, "parallel": a.parallel if world > 1 else "none",
           "seconds": dt, "names_per_sec": a.n / dt, "config": cfg.as_dict()}
    print(json.dumps(rec) if a.json else f"{a.n} names in {dt * 1e3:.3f} ms ({a.n / dt:,.0f} names/s)")
    return 0


if __name__ == "__main__":
    sys.exit(main())
