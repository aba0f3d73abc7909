"""Data-parallel work split of the reference (`namegensf.cu:628-630`).

`JPP = N / mpi_size`; rank r owns names `[JPP*r, JPP*r + JPP)`; the
`N % mpi_size` trailing names are never generated and stay zero in the root's
output (`:640`).  That quirk is kept by default (`drop_remainder=True`) and can
be switched off, in which case the tail goes to the last rank.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass

from .. import _native


@dataclass(frozen=True)
class Slice:
    jpp: int     # names per process
    start: int   # first name owned by this rank
    end: int     # one past the last

    @property
    def count(self) -> int:
        return self.end - self.start


def partition(n_names: int, world_size: int, rank: int, drop_remainder: bool = True) -> Slice:
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError(f"bad rank/size {rank}/{world_size}")
    jpp, s, e = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    _native.lib().b200gru_partition(n_names, world_size, rank, ctypes.byref(jpp), ctypes.byref(s), ctypes.byref(e))
    end = e.value
    if not drop_remainder and rank == world_size - 1:
        end = n_names
    return Slice(jpp.value, s.value, end)


def rank_info() -> tuple[int, int, int]:
    """(rank, world_size, local_rank) from the launcher env — stands in for
    MPI_Comm_rank/size (`namegensf.cu:363-364`)."""
    import os
    return (int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)),
            int(os.environ.get("LOCAL_RANK", 0)))
