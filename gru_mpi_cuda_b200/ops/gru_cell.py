"""Batched fp32 GRU layer step (the reference's 23-launch layer block,
`/root/reference/namegensf.cu:671-763`, as one op)."""
from __future__ import annotations

import ctypes

import torch

from .. import _native


def gru_cell_fp32(x, h, w_ih, w_hh, b_ih, b_hh) -> torch.Tensor:
    """`h' = GRUCell(x, h)` for `x [B,in]`, `h [B,H]`; weights stacked r,z,n.

    Bit-identical per row to `models.oracle.gru_cell`.
    """
    ts = [t.to(torch.float32).contiguous() for t in (x, h, w_ih, w_hh, b_ih, b_hh)]
    assert all(t.is_cuda for t in ts)
    x, h, w_ih, w_hh, b_ih, b_hh = ts
    B, in_dim = x.shape
    H = h.shape[1]
    assert w_ih.shape == (3 * H, in_dim) and w_hh.shape == (3 * H, H)
    out = torch.empty_like(h)
    stream = torch.cuda.current_stream(x.device).cuda_stream
    with torch.cuda.device(x.device):
        _native.check(_native.lib().b200gru_gru_cell_fp32(
            x.data_ptr(), h.data_ptr(), w_ih.data_ptr(), w_hh.data_ptr(), b_ih.data_ptr(), b_hh.data_ptr(),
            out.data_ptr(), B, in_dim, H, ctypes.c_void_p(stream)), "gru_cell_fp32")
    return out
