"""One generate() call at a given batch — the command line profiled under ncu."""
import argparse
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from gru_mpi_cuda_b200 import GRUConfig
from gru_mpi_cuda_b200.models import NameGenerator
from gru_mpi_cuda_b200.utils import weights as W

ap = argparse.ArgumentParser()
ap.add_argument("--names", type=int, default=65536)
ap.add_argument("--hidden", type=int, default=1024)
ap.add_argument("--max-len", type=int, default=3)
ap.add_argument("--mode", default="tc")
ap.add_argument("--reps", type=int, default=1)
ap.add_argument("--fake-tp", type=int, default=1, help="profile one rank's share of a TP model (B200GRU_TP_FAKE=1)")
a = ap.parse_args()
cfg = GRUConfig(hidden=a.hidden, max_len=a.max_len)
if a.fake_tp > 1:
    import os
    os.environ["B200GRU_TP_FAKE"] = "1"
gen = NameGenerator(cfg, W.random_init(cfg, 1), device=0, max_batch=a.names, mode=a.mode, tp_rank=0,
                    tp_world=a.fake_tp)
rf = W.random_floats(a.names * cfg.max_len, 2)
for _ in range(a.reps):
    out = gen.generate(a.names, rf)
torch.cuda.synchronize()
print("ok", out.shape, int(out.sum()))
gen.finalize()
