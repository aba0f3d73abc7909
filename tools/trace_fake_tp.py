"""Per-kernel times of ONE tensor-parallel rank's share, alone on one GPU (B200GRU_TP_FAKE=1)."""
import argparse, os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
os.environ["B200GRU_TP_FAKE"] = "1"
import torch
from gru_mpi_cuda_b200 import GRUConfig
from gru_mpi_cuda_b200.models import NameGenerator
from gru_mpi_cuda_b200.utils import weights as W
ap = argparse.ArgumentParser()
ap.add_argument("--hidden", type=int, default=2048)
ap.add_argument("--names", type=int, default=8192)
ap.add_argument("--tp", type=int, default=2)
a = ap.parse_args()
cfg = GRUConfig(hidden=a.hidden)
g = NameGenerator(cfg, W.random_init(cfg, 1), device=0, max_batch=a.names, tp_rank=0, tp_world=a.tp)
rf = W.random_floats(a.names * cfg.max_len, 2)
for _ in range(20):
    g.generate(a.names, rf)
g.set_trace(True); g.generate(a.names, rf); tr = g.trace()
agg = {}
for k, t in tr:
    agg.setdefault(k, []).append(t)
print(f"fake TP{a.tp} H={a.hidden} B={a.names} (us, median):", {k: round(sorted(v)[len(v)//2] * 1e3) for k, v in agg.items()})
g.finalize()
