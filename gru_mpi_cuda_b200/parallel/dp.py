"""Data-parallel name generation across GPUs.

Reference behaviour (`/root/reference/namegensf.cu`): rank 0 owns
`random_floats[N*MAX_LEN]` and `output[N*(MAX_LEN+1)]` in host memory;
`MPI_Scatter` (:636) hands each rank `JPP*MAX_LEN` floats, every rank generates
its `JPP = N // size` names (:628-630, :649), `MPI_Gather` (:889) collects
`JPP*(MAX_LEN+1)` bytes per rank into the root's `output` in rank order.

B200 design ("p2p", the product): rank 0's device buffers are mapped into every
rank (CUDA IPC over NVLink).  There is no scatter and no gather call: each
rank's sampling kernel *loads* its own floats from rank 0's buffer and *stores*
its characters into rank 0's output rows while it runs, then release-stores a
completion flag in rank 0's memory; rank 0 runs a tiny kernel that
acquire-polls the flags.  Epoch numbers make the flags reusable.

"nccl" (the baseline kept for comparison): `dist.scatter` -> local generate ->
`dist.gather`, i.e. the reference's structure with NCCL in place of MPI.
"""
from __future__ import annotations

import ctypes

import torch
import torch.distributed as dist

from .. import _native
from .partition import partition
from .symm import SymmBuffer


class DPContext:
    def __init__(self, cfg, device: int, max_names_total: int, group=None, method: str = "p2p",
                 drop_remainder: bool = True):
        assert method in ("p2p", "nccl")
        self.cfg = cfg
        self.device = device
        self.group = group
        self.method = method
        self.drop_remainder = drop_remainder
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.max_total = int(max_names_total)
        self.epoch = 0
        self.lib = _native.lib()
        T = cfg.max_len
        self.rf_bytes = self.max_total * T * 4
        self.out_bytes = self.max_total * (T + 1)
        # [0] ready flag, [1..world] done flags (uint32), padded
        self.flag_bytes = 4 * (1 + self.world + 15) // 16 * 16 + 64
        if method == "p2p":
            self.rf = SymmBuffer(self.rf_bytes, device, group)
            self.out = SymmBuffer(self.out_bytes, device, group)
            self.flags = SymmBuffer(self.flag_bytes, device, group)
        self._host_out = None
        self._copy_stream = None
        self._epoch_host = torch.zeros(64, dtype=torch.int32).pin_memory() if torch.cuda.is_available() else None
        self.extra_launches = 0

    # root-side views
    def _root(self, buf: SymmBuffer) -> int:
        return buf.ptrs[0]

    def generate(self, gen, n_names: int, random_floats: torch.Tensor | None, mode: str | None = None,
                 to_host: bool = True, out: torch.Tensor | None = None):
        """Collective: every rank calls it; only rank 0's `random_floats` is
        read and only rank 0 gets the `[n_names, max_len+1]` result.

        `out` (root only) is the reference's caller-allocated `char *output`
        (`namegensf.cu:624-627`): a pinned host `uint8[n_names, max_len+1]`
        tensor that receives the rows directly (one device-to-host copy, no
        staging copy) and is returned."""
        if n_names > self.max_total:
            raise ValueError(f"n_names={n_names} > max_names_total={self.max_total}")
        if self.method == "nccl":
            res = self._generate_nccl(gen, n_names, random_floats, mode, to_host)
            if out is not None and res is not None:
                out.copy_(res)
                return out
            return res
        if out is not None and self.rank == 0:
            if (out.dtype != torch.uint8 or out.device.type != "cpu" or not out.is_contiguous()
                    or out.numel() != n_names * (self.cfg.max_len + 1)):
                raise ValueError("out must be a contiguous host uint8 tensor of n_names*(max_len+1) bytes")
        cfg = self.cfg
        T = cfg.max_len
        sl = partition(n_names, self.world, self.rank, self.drop_remainder)
        self.epoch += 1
        e = self.epoch
        stream = ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        flags_root = self._root(self.flags)
        launches = 0
        with torch.cuda.device(self.device):
            main = torch.cuda.current_stream(self.device)
            if self.rank == 0:
                # The floats go up on a side stream and are announced with a flag;
                # every rank (the root included) waits for that flag only right
                # before its first sampling kernel, so the first character's GRU
                # kernels overlap the host-to-device copy.
                if self._copy_stream is None:
                    self._copy_stream = torch.cuda.Stream(device=self.device)
                cs = self._copy_stream
                cs.wait_stream(main)  # the previous call (all ranks done reading) is behind us
                cstream = ctypes.c_void_p(cs.cuda_stream)
                if random_floats is not None:  # None: reuse the floats already on the device
                    rf = random_floats.reshape(-1)[: n_names * T]
                    if rf.dtype != torch.float32 or rf.device.type != "cpu":
                        raise ValueError("random_floats must be a host float32 tensor on the root rank")
                    _native.check(self.lib.b200gru_memcpy_async(ctypes.c_void_p(self.rf.local_ptr), rf.data_ptr(),
                                                                rf.numel() * 4, 1, cstream), "H2D floats")
                # announce with a 4-byte copy (no kernel: a spinning wait kernel and a
                # signalling kernel on one GPU are not guaranteed to be co-scheduled)
                slot = self._epoch_host[e % self._epoch_host.numel():][:1]
                slot.fill_(e)
                _native.check(self.lib.b200gru_memcpy_async(ctypes.c_void_p(flags_root), slot.data_ptr(), 4, 1,
                                                            cstream), "ready flag")
                covered = sl.jpp * self.world if self.drop_remainder else n_names
                if covered < n_names:  # rows nobody generates stay zero (ref:640)
                    _native.check(self.lib.b200gru_memset_async(
                        ctypes.c_void_p(self.out.local_ptr + covered * (T + 1)), 0,
                        (n_names - covered) * (T + 1), stream))
            io = _native.CIo(self._root(self.rf) + sl.start * T * 4,
                             self._root(self.out) + sl.start * (T + 1),
                             flags_root + 4 * (1 + self.rank), e,
                             flags_root, e)  # ready flag: acquire-polled (over NVLink off the root)
            m = gen.mode if mode is None else {"exact": 0, "fp32": 0, "tc": 1, "bf16": 1}[mode]
            _native.check(self.lib.b200gru_generate(gen._h, m, sl.count, ctypes.byref(io), stream), "generate")
            launches += self.lib.b200gru_last_launch_count(gen._h)
            result = None
            if self.rank == 0:
                _native.check(self.lib.b200gru_wait_flags(ctypes.c_void_p(flags_root + 4), self.world, e, stream))
                launches += 1
                if to_host:
                    nb = n_names * (T + 1)
                    if out is not None:
                        dst = out
                    else:
                        if self._host_out is None or self._host_out.numel() < nb:
                            self._host_out = torch.empty(self.out_bytes, dtype=torch.uint8).pin_memory()
                        dst = self._host_out
                    _native.check(self.lib.b200gru_memcpy_async(dst.data_ptr(), ctypes.c_void_p(self.out.local_ptr),
                                                                nb, 2, stream))
                    torch.cuda.current_stream(self.device).synchronize()
                    # without a caller buffer the rows are copied out of the reused staging buffer
                    result = out.view(n_names, T + 1) if out is not None else \
                        self._host_out[:nb].view(n_names, T + 1).clone()
            elif to_host:
                torch.cuda.current_stream(self.device).synchronize()
        gen.last_launches = launches
        return result

    def _generate_nccl(self, gen, n_names, random_floats, mode, to_host=True):
        """Baseline: the reference's structure with NCCL in place of MPI —
        scatter (:636), local generate, gather (:889)."""
        cfg = self.cfg
        T = cfg.max_len
        sl = partition(n_names, self.world, self.rank, True)
        dev = torch.device("cuda", self.device)
        if getattr(self, "_mine", None) is None or self._mine.numel() != sl.jpp * T:
            self._mine = torch.empty(sl.jpp * T, dtype=torch.float32, device=dev)
        if self.rank == 0:
            if random_floats is not None:  # None: reuse the floats already on the device
                self._full = random_floats.reshape(-1)[: sl.jpp * self.world * T].to(dev, non_blocking=True)
            chunks = list(self._full.view(self.world, sl.jpp * T).unbind(0))
        else:
            chunks = None
        dist.scatter(self._mine, chunks, src=0, group=self.group)
        out = gen.generate_device(self._mine, sl.jpp, mode=mode).contiguous()
        gathered = [torch.empty_like(out) for _ in range(self.world)] if self.rank == 0 else None
        dist.gather(out, gathered, dst=0, group=self.group)
        gen.last_launches += 2
        if self.rank != 0 or not to_host:
            if to_host:
                torch.cuda.current_stream(self.device).synchronize()
            return None
        res = torch.zeros(n_names, T + 1, dtype=torch.uint8)
        res[: sl.jpp * self.world] = torch.cat(gathered, 0).cpu()
        return res

    def close(self):
        if self.method == "p2p":
            for b in (self.rf, self.out, self.flags):
                b.close()
