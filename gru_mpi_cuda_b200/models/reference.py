"""Oracle B — the same network written with stock PyTorch modules.

The reference's launch sequence (`/root/reference/namegensf.cu:664-873`) is
exactly `nn.Embedding -> nn.GRU(E, H, num_layers) -> nn.Linear(H, C)` with
PyTorch's gate order (r, z, n) and `b_hn` inside the `r *` product, so a plain
torch model loaded from the same weight file is an independent check on
Oracle A (different summation order, so agreement is to tolerance).
"""
from __future__ import annotations

import torch
from torch import nn

from ..config import GRUConfig
from ..utils import weights as W


class TorchNameGen(nn.Module):
    def __init__(self, cfg: GRUConfig):
        super().__init__()
        self.cfg = cfg
        self.embedding = nn.Embedding(cfg.num_char, cfg.emb_dim)
        self.gru = nn.GRU(cfg.emb_dim, cfg.hidden, num_layers=cfg.num_layers, batch_first=True)
        self.fc = nn.Linear(cfg.hidden, cfg.num_char)

    @classmethod
    def from_flat(cls, cfg: GRUConfig, flat: torch.Tensor) -> "TorchNameGen":
        m = cls(cfg)
        m.load_state_dict(W.to_torch_state(cfg, flat))
        return m.eval()

    @torch.no_grad()
    def step(self, chars: torch.Tensor, h: torch.Tensor):
        """chars [B] int64, h [L,B,H] -> logits [B,C], new h."""
        x = self.embedding(chars).unsqueeze(1)
        y, h = self.gru(x, h)
        return self.fc(y[:, 0]), h

    @torch.no_grad()
    def generate(self, rfloats: torch.Tensor, n_names: int | None = None, return_logits: bool = False):
        """Batched generation with the reference's sampler and output format."""
        cfg = self.cfg
        dev = self.fc.weight.device
        if n_names is None:
            n_names = rfloats.numel() // cfg.max_len
        r = rfloats.reshape(-1)[: n_names * cfg.max_len].view(n_names, cfg.max_len).to(dev)
        out = torch.zeros(n_names, cfg.max_len + 1, dtype=torch.uint8, device=dev)
        logits_all = torch.zeros(n_names, cfg.max_len, cfg.num_char, device=dev) if return_logits else None
        chars = torch.full((n_names,), cfg.sos, dtype=torch.long, device=dev)
        h = torch.zeros(cfg.num_layers, n_names, cfg.hidden, device=dev)
        alive = torch.ones(n_names, dtype=torch.bool, device=dev)
        for t in range(cfg.max_len):
            logits, h = self.step(chars, h)
            # softmax without max-subtraction (namegensf.cu:302-313)
            e = torch.exp(logits)
            p = e / e.sum(dim=1, keepdim=True)
            # sequential fp32 prefix sum, first index with psum > r, else C-1 (:322-333)
            cdf = torch.cumsum(p, dim=1)
            hit = cdf > r[:, t:t + 1]
            sel = torch.where(hit.any(dim=1), hit.float().argmax(dim=1),
                              torch.full_like(chars, cfg.num_char - 1))
            out[:, t] = torch.where(alive, sel, torch.zeros_like(sel)).to(torch.uint8)
            if return_logits:
                logits_all[:, t] = torch.where(alive[:, None], logits, torch.zeros_like(logits))
            chars = sel
            alive = alive & (sel != cfg.eos)
        return (out, logits_all) if return_logits else out
