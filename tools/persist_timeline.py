#!/usr/bin/env python
"""Print the in-kernel timeline of one persistent generate (default shape).

    python tools/persist_timeline.py [n_names]

Two CTAs are instrumented (the first weight-holding CTA and the first FC CTA);
each role stamps %globaltimer at its hand-off points.  Output: one line per
event, microseconds since the first stamp."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from gru_mpi_cuda_b200 import GRUConfig
from gru_mpi_cuda_b200.models import NameGenerator
from gru_mpi_cuda_b200.utils import weights as W

EV = {1: "operand counter reached", 2: "loads issued", 3: "first operand landed", 4: "MMAs issued",
      5: "accumulator ready", 6: "published", 7: "chars visible", 8: "role start", 9: "weights resident",
      10: "fc: exponentials done", 11: "fc: segment sums exchanged", 12: "fc: sampled", 13: "fc: bookkeeping done"}


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    cfg = GRUConfig()
    g = NameGenerator(cfg, W.random_init(cfg, 1), device=0, max_batch=max(n, 256))
    g.set_path("persistent")
    r = W.random_floats(n * cfg.max_len, 2)
    for _ in range(3):
        g.generate(n, r)
    g.set_trace(True)
    g.generate(n, r)
    spans = g.trace()
    tl = g.persist_timeline()
    g.set_trace(False)
    print(f"n={n} kernel span (CUDA events): {spans}")
    rows = []
    for (cta, role), evs in tl.items():
        for t, step, layer, ev in evs:
            rows.append((t, cta, role, step, layer, EV.get(ev, ev)))
    rows.sort()
    for t, cta, role, step, layer, ev in rows:
        print(f"{t:9.2f} us  {cta:8s} {role:8s} step {step} layer {layer}  {ev}")
    g.finalize()


if __name__ == "__main__":
    main()
