#!/usr/bin/env python
"""EOS-heavy generation: what compaction of finished names buys.

Random-init weights almost never sample EOS (1/256 per step); a trained model
does.  b_fc[eos] is raised so that P(EOS) is about 0.25 per step (mean length
about 4 of MAX_LEN 10) and one 65,536-name call is timed with compaction on and
off (device time, CUDA events, multi-launch tensor-core path)."""
import argparse
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from gru_mpi_cuda_b200 import GRUConfig
from gru_mpi_cuda_b200.models import NameGenerator
from gru_mpi_cuda_b200.utils import weights as W


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--names", type=int, default=65536)
    ap.add_argument("--bias", type=float, default=4.44, help="b_fc[eos]; exp(b)/(exp(b)+255) ~ P(EOS)")
    ap.add_argument("--steps", type=int, default=20)
    a = ap.parse_args()
    cfg = GRUConfig()
    p = W.random_init(cfg, 1)
    W.split(cfg, p)["b_fc"][cfg.eos] = a.bias
    n, T = a.names, cfg.max_len
    rf = W.random_floats(n * T, 2)
    res = {}
    with NameGenerator(cfg, p, device=0, max_batch=n) as g:
        g.set_path("multi")
        rows = {}
        for on in (True, False):
            g.set_compaction(on)
            rows[on] = g.generate(n, rf)
            for _ in range(3):
                g.generate_device(g._rf_dev, n)
            torch.cuda.synchronize()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            for _ in range(a.steps):
                g.generate_device(g._rf_dev, n)
            e.record()
            torch.cuda.synchronize()
            res["compaction_on_ms" if on else "compaction_off_ms"] = s.elapsed_time(e) / a.steps
            res["launches_on" if on else "launches_off"] = g.last_launches
    out = rows[True]
    lens = torch.where((out[:, :T] == cfg.eos).any(dim=1), (out[:, :T] == cfg.eos).float().argmax(dim=1) + 1,
                       torch.full((n,), T))
    res.update(names=n, mean_length=lens.float().mean().item(), identical_rows=bool(torch.equal(rows[True], rows[False])),
               speedup=res["compaction_off_ms"] / res["compaction_on_ms"])
    print(json.dumps(res))


if __name__ == "__main__":
    main()
