"""Peer-visible device buffers (CUDA IPC) and raw-pointer tensor views.

Bootstrap only: the process group is used once to all-gather 64-byte IPC
handles.  After that kernels address peer memory directly over NVLink; see
csrc/symm.cu.  The reference's equivalent plumbing is MPI itself
(`/root/reference/namegensf.cu:362-364, :615, :636, :889`).
"""
from __future__ import annotations

import ctypes

import torch
import torch.distributed as dist

from .. import _native


class _RawCudaArray:
    """Minimal `__cuda_array_interface__` carrier so torch can view a raw
    device pointer owned by the native library."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {
            "shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3, "strides": None}


def view_u8(ptr: int, nbytes: int, device: int) -> torch.Tensor:
    return torch.as_tensor(_RawCudaArray(ptr, nbytes), device=torch.device("cuda", device))


class SymmBuffer:
    """One buffer per rank, same size everywhere, each rank holding mapped
    pointers to every peer's copy."""

    def __init__(self, nbytes: int, device: int, group=None):
        self.lib = _native.lib()
        self.device = device
        self.nbytes = int(nbytes)
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        ptr = ctypes.c_void_p()
        handle = (ctypes.c_ubyte * 64)()
        _native.check(self.lib.b200gru_symm_alloc(device, self.nbytes, ctypes.byref(ptr), handle), "symm_alloc")
        self.local_ptr = ptr.value
        self.ptrs = [0] * self.world
        self.ptrs[self.rank] = self.local_ptr
        self._opened = []
        if self.world > 1:
            mine = (bytes(handle), device)
            gathered = [None] * self.world
            dist.all_gather_object(gathered, mine, group=group)
            for r, (h, _peer_dev) in enumerate(gathered):
                if r == self.rank:
                    continue
                p = ctypes.c_void_p()
                buf = (ctypes.c_ubyte * 64).from_buffer_copy(h)
                _native.check(self.lib.b200gru_symm_open(device, buf, ctypes.byref(p)), f"symm_open(rank {r})")
                self.ptrs[r] = p.value
                self._opened.append(p.value)
            dist.barrier(group=group)
        self.local = view_u8(self.local_ptr, self.nbytes, device)

    def close(self):
        if self.local_ptr is None:
            return
        if self.world > 1 and dist.is_initialized():
            torch.cuda.synchronize(self.device)
            dist.barrier(group=self.group)
        for p in self._opened:
            self.lib.b200gru_symm_close(self.device, ctypes.c_void_p(p))
        self._opened = []
        self.local = None
        if self.world > 1 and dist.is_initialized():
            dist.barrier(group=self.group)
        self.lib.b200gru_symm_free(self.device, ctypes.c_void_p(self.local_ptr))
        self.local_ptr = None


class McBuffer:
    """Peer-visible buffer that also has an NVSwitch MULTICAST address.

    Same role as `SymmBuffer` (one buffer per rank, every rank holds pointers
    to all of them) plus `mc_ptr`: an address in a multicast mapping of all the
    copies, so that a single `multimem.st` / `multimem.red` issued by a kernel
    lands in every rank's buffer (NVLink SHARP).  CUDA IPC handles cannot carry
    that, the memory has to come from `cuMemCreate` and be bound to a
    `cuMulticast` object on every rank; torch's symmetric memory does exactly
    this bootstrap (handle exchange over the process group), so it is used for
    the plumbing only.  `mc_ptr` is 0 where the fabric has no multicast.
    """

    def __init__(self, nbytes: int, device: int, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        self.device = device
        self.nbytes = int(nbytes)
        self.group = group if group is not None else dist.group.WORLD
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self._t = symm_mem.empty(self.nbytes, dtype=torch.uint8, device=torch.device("cuda", device))
        self._t.zero_()
        self._hdl = symm_mem.rendezvous(self._t, self.group)
        self.ptrs = [int(p) for p in self._hdl.buffer_ptrs]
        self.local_ptr = self.ptrs[self.rank]
        self.mc_ptr = int(self._hdl.multicast_ptr or 0)
        self.local = self._t
        torch.cuda.synchronize(device)
        dist.barrier(group=group)

    def close(self):
        if self._t is None:
            return
        torch.cuda.synchronize(self.device)
        if dist.is_initialized():
            dist.barrier(group=self.group)
        self._hdl = None
        self._t = None
        self.local = None
