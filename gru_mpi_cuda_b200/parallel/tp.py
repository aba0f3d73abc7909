"""Tensor-parallel generation: hidden dimension sharded over the GPUs of a node.

Not in the reference (every rank there holds all weights,
`/root/reference/namegensf.cu:370-407`); this is BASELINE.json's
"hidden 8192 tensor-parallel across 8 GPUs" configuration.

Every rank runs the SAME names.  Rank r owns hidden units
`[r*H/G, (r+1)*H/G)` of every layer and only those weight rows.  After the gate
epilogue of a layer kernel, each thread stores its slice of `h'` into the
full-width hidden buffer of *every* rank (plain stores to NVLink-mapped peer
addresses) and bumps a per-(source rank, 128-name tile) counter there with a
system-scope release.  The TMA producer of the next kernel starts its K loop
with the locally produced chunk and acquires each remote rank's counter right
before loading that rank's chunk — the all-gather is overlapped with the GEMM
tile by tile and there is no collective call on the hot path.  FC + sampling is
computed redundantly on every rank (2.8% of the FLOPs) so no sampled character
ever has to be communicated.
"""
from __future__ import annotations

import ctypes

import torch
import torch.distributed as dist

from .. import _native
from ..models.namegen import NameGenerator
import os

from .symm import McBuffer, SymmBuffer


class TPContext:
    def __init__(self, cfg, params, device: int, max_batch: int, group=None, multicast: bool | None = None):
        """`multicast`: publish the hidden slices through the NVSwitch multicast
        mapping (one store per 16 bytes instead of one per peer).  Default: on
        when the fabric offers it (`B200GRU_TP_MC=0` turns it off for A/B runs)."""
        self.cfg = cfg
        self.device = device
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.lib = _native.lib()
        self.gen = NameGenerator(cfg, params, device=device, max_batch=max_batch, mode="tc",
                                 tp_rank=self.rank, tp_world=self.world)
        T = cfg.max_len
        self.max_batch = max_batch
        nbytes = self.lib.b200gru_tp_shared_bytes(self.gen._h)
        if multicast is None:
            multicast = os.environ.get("B200GRU_TP_MC", "1") != "0"
        self.shared = None
        if multicast:
            try:
                self.shared = McBuffer(nbytes, device, group)
            except Exception as e:  # noqa: BLE001  (no symmetric-memory support here: unicast P2P instead)
                self.mc_error = repr(e)
        # every rank must take the same route
        ok = torch.tensor([1 if (self.shared is not None and self.shared.mc_ptr) else 0], device=torch.device("cuda", device))
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
        if multicast and not int(ok.item()):
            if self.shared is not None:
                self.shared.close()
            self.shared = None
        if self.shared is None:
            self.shared = SymmBuffer(nbytes, device, group)
        self.multicast = isinstance(self.shared, McBuffer)
        peers = (ctypes.c_void_p * self.world)(*[ctypes.c_void_p(p) for p in self.shared.ptrs])
        _native.check(self.lib.b200gru_tp_bind_mc(self.gen._h, ctypes.c_void_p(self.shared.local_ptr), peers,
                                                   ctypes.c_void_p(self.shared.mc_ptr) if self.multicast else None),
                      "tp_bind")
        # root-owned I/O, read / written in place by every rank's kernels (as in DP)
        # two halves, used alternately: a peer may still be reading call e's floats
        # (every rank's sampling kernels load them over NVLink) while the root
        # already uploads call e+1's; half (e & 1) is only rewritten for call e+2,
        # whose upload is ordered after the epoch fence of call e+1 on the root's stream
        self.rf_half = (max_batch * T * 4 + 255) // 256 * 256
        self.rf = SymmBuffer(2 * self.rf_half, device, group)
        self.flags = SymmBuffer(256, device, group)
        self.out = torch.zeros(max_batch, T + 1, dtype=torch.uint8, device=torch.device("cuda", device))
        self._host_out = torch.empty(max_batch, T + 1, dtype=torch.uint8).pin_memory()
        self.epoch = 0
        dist.barrier(group=group)

    def generate(self, n_names: int, random_floats: torch.Tensor | None, to_host: bool = True):
        """Collective.  Rank 0's `random_floats` drive every rank; rank 0 returns
        `uint8[n_names, max_len+1]`, the others `None`."""
        cfg, T = self.cfg, self.cfg.max_len
        if n_names > self.max_batch:
            raise ValueError("n_names exceeds max_batch")
        self.epoch += 1
        e = self.epoch
        stream = ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        flag_root = self.flags.ptrs[0]
        launches = 0
        with torch.cuda.device(self.device):
            if self.rank == 0:
                if random_floats is not None:
                    rf = random_floats.reshape(-1)[: n_names * T]
                    _native.check(self.lib.b200gru_memcpy_async(
                        ctypes.c_void_p(self.rf.local_ptr + (e & 1) * self.rf_half), rf.data_ptr(),
                        rf.numel() * 4, 1, stream))
                elif e > 1:  # reuse the previous call's floats: bring them over to this call's half
                    _native.check(self.lib.b200gru_memcpy_async(
                        ctypes.c_void_p(self.rf.local_ptr + (e & 1) * self.rf_half),
                        ctypes.c_void_p(self.rf.local_ptr + ((e - 1) & 1) * self.rf_half),
                        n_names * T * 4, 3, stream))
                _native.check(self.lib.b200gru_signal_flag(ctypes.c_void_p(flag_root), e, stream))
            else:
                _native.check(self.lib.b200gru_wait_flags(ctypes.c_void_p(flag_root), 1, e, stream))
            launches += 1
            io = _native.CIo(self.rf.ptrs[0] + (e & 1) * self.rf_half, self.out.data_ptr(), None, 0)
            _native.check(self.lib.b200gru_generate(self.gen._h, 1, n_names, ctypes.byref(io), stream), "generate")
            launches += self.lib.b200gru_last_launch_count(self.gen._h)
            self.gen.last_launches = launches
            result = None
            if to_host:
                if self.rank == 0:
                    self._host_out[:n_names].copy_(self.out[:n_names], non_blocking=True)
                torch.cuda.current_stream(self.device).synchronize()
                if self.rank == 0:
                    result = self._host_out[:n_names].clone()
        return result

    def local_rows(self, n_names: int) -> torch.Tensor:
        """This rank's own copy of the sampled rows (every rank samples all names)."""
        torch.cuda.current_stream(self.device).synchronize()
        return self.out[:n_names].cpu()

    def close(self):
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)
        self.gen.finalize()
        for b in (self.rf, self.flags, self.shared):
            b.close()
