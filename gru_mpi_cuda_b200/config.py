"""Runtime model configuration.

The reference hard-codes its shape as macros in `namegen.h`, which it does not
ship (`/root/reference/namegensf.cu:1`, used at `:375-407`).  `HIDDEN_DIM=1024`
and `NUM_CHAR=256` are confirmed by comments (`:694`, `:862`); `EMBEDDING_DIM`,
`MAX_LEN`, `SOS`, `EOS` are the course-skeleton values and are plain runtime
fields here.
"""
from __future__ import annotations

import ctypes
from dataclasses import asdict, dataclass


class CConfig(ctypes.Structure):
    """Mirror of `b200gru_config` in include/b200gru.h."""

    _fields_ = [(n, ctypes.c_int32) for n in
                ("num_char", "emb_dim", "hidden", "num_layers", "max_len", "sos", "eos")]


@dataclass(frozen=True)
class GRUConfig:
    num_char: int = 256
    emb_dim: int = 512
    hidden: int = 1024
    num_layers: int = 2
    max_len: int = 10
    sos: int = 0
    eos: int = 1

    def to_c(self) -> CConfig:
        return CConfig(**asdict(self))

    @property
    def row_bytes(self) -> int:
        """Bytes per output row: MAX_LEN chars + NUL (`namegensf.cu:624`)."""
        return self.max_len + 1

    def as_dict(self) -> dict:
        return asdict(self)

    # named shapes from BASELINE.json
    @staticmethod
    def default() -> "GRUConfig":
        return GRUConfig()

    @staticmethod
    def tiny() -> "GRUConfig":
        """Small shape for unit tests (still multiples of the 64-wide tiles)."""
        return GRUConfig(num_char=64, emb_dim=64, hidden=128, max_len=6)

    @staticmethod
    def tp8192() -> "GRUConfig":
        return GRUConfig(hidden=8192)
