"""CPU restatement of the reference's distributed flow, for tests without GPUs.

`MPI_Scatter` of floats (`/root/reference/namegensf.cu:636`), per-rank
generation with Oracle A, `MPI_Gather` of rows (`:889`), root-zeroed output
(`:638-641`) — over a gloo process group.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from ..models import oracle
from .partition import partition


def generate_cpu(cfg, params: torch.Tensor, n_names: int, random_floats: torch.Tensor | None, group=None):
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    T = cfg.max_len
    sl = partition(n_names, world, rank)
    mine = torch.empty(sl.jpp * T, dtype=torch.float32)
    chunks = None
    if rank == 0:
        chunks = list(random_floats.reshape(-1)[: sl.jpp * world * T].view(world, sl.jpp * T).contiguous().unbind(0))
    dist.scatter(mine, chunks, src=0, group=group)
    rows = oracle.generate(cfg, params, mine, n_names=sl.jpp)
    gathered = [torch.empty_like(rows) for _ in range(world)] if rank == 0 else None
    dist.gather(rows, gathered, dst=0, group=group)
    if rank != 0:
        return None
    out = torch.zeros(n_names, T + 1, dtype=torch.uint8)
    out[: sl.jpp * world] = torch.cat(gathered, 0)
    return out
