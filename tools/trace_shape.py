"""Per-kernel device times (engine trace) for a given shape on one GPU."""
import argparse, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from gru_mpi_cuda_b200 import GRUConfig
from gru_mpi_cuda_b200.models import NameGenerator
from gru_mpi_cuda_b200.utils import weights as W

ap = argparse.ArgumentParser()
ap.add_argument("--hidden", type=int, default=1024)
ap.add_argument("--names", type=int, default=65536)
a = ap.parse_args()
cfg = GRUConfig(hidden=a.hidden)
gen = NameGenerator(cfg, W.random_init(cfg, 1), device=0, max_batch=a.names, mode="tc")
rf = W.random_floats(a.names * cfg.max_len, 2)
for _ in range(3):
    gen.generate(a.names, rf)
gen.set_trace(True)
gen.generate(a.names, rf)
tr = gen.trace()
agg = {}
for k, t in tr:
    agg.setdefault(k, []).append(t)
H, B, C = cfg.hidden, a.names, cfg.num_char
fl = {"layer0": 2.0 * B * 3 * H * H, "layerN": 2.0 * B * 3 * H * 2 * H, "fc_sample": 2.0 * B * C * H}
print(f"H={H} B={B}: " + "; ".join(
    f"{k}: steady {sorted(v)[len(v)//2]*1e3:.0f} us" + (f" = {fl[k]/(sorted(v)[len(v)//2]*1e-3)/1e12:.0f} TFLOP/s" if k in fl else "")
    for k, v in agg.items()))
gen.finalize()
