"""gru_mpi_cuda_b200 — Blackwell-native GRU name generation.

Capabilities of junyongeom/gru-mpi-cuda (`/root/reference/namegensf.cu`),
rebuilt for B200: same weight file, same initialize / generate-N-names /
finalize entry points, same rank/size split and output format; batched
sm_100a kernels instead of 51 launches per character.
"""
from .config import GRUConfig  # noqa: F401

__all__ = ["GRUConfig"]
__version__ = "0.1.0"
