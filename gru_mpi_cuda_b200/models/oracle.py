"""Oracle A — the reference's CPU semantics (csrc/cpu_oracle.cpp)."""
from __future__ import annotations

import ctypes

import torch

from .. import _native
from ..config import GRUConfig

LIBM = 1


def generate(cfg: GRUConfig, params: torch.Tensor, rfloats: torch.Tensor, n_names: int | None = None,
             return_logits: bool = False, libm: bool = False):
    """Run `namegen()`'s math for `n_names` names on the CPU.

    Returns `uint8[n, max_len+1]` (and fp32 logits `[n, max_len, num_char]`).
    Mirrors `/root/reference/namegensf.cu:649-885`.
    """
    params = params.detach().to(torch.float32).contiguous().cpu()
    rfloats = rfloats.detach().to(torch.float32).contiguous().cpu()
    if n_names is None:
        n_names = rfloats.numel() // cfg.max_len
    assert rfloats.numel() >= n_names * cfg.max_len
    out = torch.zeros(n_names, cfg.max_len + 1, dtype=torch.uint8)
    logits = torch.zeros(n_names, cfg.max_len, cfg.num_char) if return_logits else None
    c = cfg.to_c()
    _native.check(_native.lib().b200gru_oracle_generate(
        ctypes.byref(c), params.data_ptr(), n_names, rfloats.data_ptr(), out.data_ptr(),
        logits.data_ptr() if return_logits else None, LIBM if libm else 0), "oracle_generate")
    return (out, logits) if return_logits else out


def random_select(probs: torch.Tensor, r: float) -> int:
    """Inverse-CDF sampler, `namegensf.cu:322-333`."""
    probs = probs.detach().to(torch.float32).contiguous().cpu()
    return _native.lib().b200gru_random_select(probs.data_ptr(), probs.numel(), float(r))


def softmax(x: torch.Tensor, libm: bool = False) -> torch.Tensor:
    x = x.detach().to(torch.float32).contiguous().cpu()
    y = torch.empty_like(x)
    _native.check(_native.lib().b200gru_oracle_softmax(x.data_ptr(), y.data_ptr(), x.numel(), LIBM if libm else 0))
    return y


def matvec(w: torch.Tensor, x: torch.Tensor) -> torch.Tensor:
    w = w.detach().to(torch.float32).contiguous().cpu()
    x = x.detach().to(torch.float32).contiguous().cpu()
    y = torch.empty(w.shape[0])
    _native.check(_native.lib().b200gru_oracle_matvec(w.data_ptr(), x.data_ptr(), y.data_ptr(), w.shape[0], w.shape[1]))
    return y


def gru_cell(w_ih, w_hh, b_ih, b_hh, x, h, libm: bool = False) -> torch.Tensor:
    """One layer step; returns the new hidden vector."""
    t = [a.detach().to(torch.float32).contiguous().cpu() for a in (w_ih, w_hh, b_ih, b_hh, x)]
    h = h.detach().to(torch.float32).contiguous().cpu().clone()
    _native.check(_native.lib().b200gru_oracle_gru_cell(*[a.data_ptr() for a in t], h.data_ptr(),
                                                        t[0].shape[1], h.numel(), LIBM if libm else 0))
    return h
