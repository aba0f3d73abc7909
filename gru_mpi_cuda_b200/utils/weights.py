"""Weight-file I/O: the reference's headerless fp32 parameter file.

Layout and order: `/root/reference/namegensf.cu:375-407` (see
csrc/weights_io.cpp).  The native library owns the layout table; this module
exposes it as torch tensors and converts to/from a `torch.nn.GRU` state dict.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from pathlib import Path

import numpy as np
import torch

from .. import _native
from ..config import GRUConfig


@dataclass(frozen=True)
class TensorSpec:
    name: str
    offset: int
    size: int
    rows: int
    cols: int


def layout(cfg: GRUConfig) -> list[TensorSpec]:
    L = _native.lib()
    c = cfg.to_c()
    n = _native.check(L.b200gru_num_tensors(ctypes.byref(c)), "num_tensors")
    off = (ctypes.c_int64 * n)()
    sz = (ctypes.c_int64 * n)()
    rows = (ctypes.c_int32 * n)()
    cols = (ctypes.c_int32 * n)()
    _native.check(L.b200gru_param_layout(ctypes.byref(c), off, sz, rows, cols), "param_layout")
    return [TensorSpec(L.b200gru_tensor_name(ctypes.byref(c), i).decode(), off[i], sz[i], rows[i], cols[i])
            for i in range(n)]


def param_count(cfg: GRUConfig) -> int:
    c = cfg.to_c()
    return _native.check(_native.lib().b200gru_param_count(ctypes.byref(c)), "param_count")


def random_init(cfg: GRUConfig, seed: int = 0) -> torch.Tensor:
    """Deterministic random-init weights of the named shape (flat fp32, CPU)."""
    n = param_count(cfg)
    out = torch.empty(n, dtype=torch.float32)
    c = cfg.to_c()
    _native.check(_native.lib().b200gru_random_init(ctypes.byref(c), seed, out.data_ptr(), n), "random_init")
    return out


def random_floats(n: int, seed: int = 0, pin: bool = False) -> torch.Tensor:
    """`n` uniform floats in [0,1): the `random_floats` argument of namegen()."""
    out = torch.empty(n, dtype=torch.float32, pin_memory=pin)
    _native.check(_native.lib().b200gru_random_floats(seed, out.data_ptr(), n), "random_floats")
    return out


def save(path: str | Path, flat: torch.Tensor) -> None:
    flat = flat.detach().to(torch.float32).contiguous().cpu()
    _native.check(_native.lib().b200gru_write_binary(str(path).encode(), flat.data_ptr(), flat.numel() * 4),
                  f"write {path}")


def load(path: str | Path, cfg: GRUConfig) -> torch.Tensor:
    """read_binary() + size check; returns the flat fp32 parameter vector."""
    L = _native.lib()
    buf = ctypes.c_void_p()
    nbytes = ctypes.c_size_t()
    _native.check(L.b200gru_read_binary(str(path).encode(), ctypes.byref(buf), ctypes.byref(nbytes)),
                  f"read {path}")
    try:
        need = param_count(cfg) * 4
        if nbytes.value != need:
            raise ValueError(f"{path}: {nbytes.value} bytes, but config {cfg} needs {need}")
        arr = np.ctypeslib.as_array(ctypes.cast(buf, ctypes.POINTER(ctypes.c_float)), shape=(need // 4,))
        return torch.from_numpy(arr.copy())
    finally:
        L.b200gru_free_host(buf)


def split(cfg: GRUConfig, flat: torch.Tensor) -> dict[str, torch.Tensor]:
    """Views of the flat vector by reference tensor name (W_ir0, b_hn1, ...)."""
    out = {}
    for t in layout(cfg):
        v = flat[t.offset:t.offset + t.size]
        out[t.name] = v.view(t.rows, t.cols) if t.cols > 1 else v
    return out


def to_torch_state(cfg: GRUConfig, flat: torch.Tensor) -> dict[str, torch.Tensor]:
    """State dict for `models.reference.TorchNameGen` (nn.Embedding/GRU/Linear)."""
    p = split(cfg, flat)
    sd = {"embedding.weight": p["character_embedding"].clone(),
          "fc.weight": p["W_fc"].clone(), "fc.bias": p["b_fc"].clone()}
    for l in range(cfg.num_layers):
        sd[f"gru.weight_ih_l{l}"] = torch.cat([p[f"W_i{g}{l}"] for g in "rzn"], 0)
        sd[f"gru.weight_hh_l{l}"] = torch.cat([p[f"W_h{g}{l}"] for g in "rzn"], 0)
        sd[f"gru.bias_ih_l{l}"] = torch.cat([p[f"b_i{g}{l}"] for g in "rzn"], 0)
        sd[f"gru.bias_hh_l{l}"] = torch.cat([p[f"b_h{g}{l}"] for g in "rzn"], 0)
    return sd


def from_torch_state(cfg: GRUConfig, sd: dict[str, torch.Tensor]) -> torch.Tensor:
    """Inverse of `to_torch_state`: pack a torch state dict into file order."""
    flat = torch.empty(param_count(cfg), dtype=torch.float32)
    H = cfg.hidden
    for t in layout(cfg):
        if t.name == "character_embedding":
            src = sd["embedding.weight"]
        elif t.name == "W_fc":
            src = sd["fc.weight"]
        elif t.name == "b_fc":
            src = sd["fc.bias"]
        else:
            kind, gate, layer = t.name[:3], "rzn".index(t.name[3]), int(t.name[4:])
            key = {"W_i": "weight_ih", "W_h": "weight_hh", "b_i": "bias_ih", "b_h": "bias_hh"}[kind]
            src = sd[f"gru.{key}_l{layer}"][gate * H:(gate + 1) * H]
        flat[t.offset:t.offset + t.size] = src.reshape(-1).to(torch.float32)
    return flat
