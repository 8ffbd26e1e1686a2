"""Symmetric memory (N1): the same buffer on every rank, mapped into every rank's address space.

This is what replaces the per-task gRPC server of the reference (``tf.train.Server``,
example.py:38-40, lr2.py:331): a worker does not *connect* to the parameter server, it *maps*
every shard, and the kernels in ``csrc/kernels/ps_comm.cu`` move the bytes over NVLink.

* multi-process (one process per GPU): ``cudaMalloc`` + CUDA IPC handles, exchanged once over
  the control-plane process group (gloo) — ``b200ps_symm_alloc/open`` in ``csrc/symm/symm.cu``;
* single process: ``world`` *virtual ranks* on one device (all "peer" pointers are local) —
  the configuration every multi-rank kernel is unit-tested in, and the degenerate world=1
  case of a single-GPU run.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import List, Optional

import torch

from ..ops import native


class _RawCudaBuffer:
    """Exposes a raw device address to torch through ``__cuda_array_interface__``."""

    def __init__(self, ptr: int, nbytes: int, owner=None):
        self.ptr, self.nbytes, self._owner = int(ptr), int(nbytes), owner
        self.__cuda_array_interface__ = {
            "shape": (self.nbytes,), "typestr": "|u1", "data": (self.ptr, False), "version": 2,
            "strides": None,
        }


def tensor_from_ptr(ptr: int, nbytes: int, device, owner=None) -> torch.Tensor:
    """uint8 tensor aliasing ``nbytes`` of device memory at ``ptr`` (no ownership transfer)."""
    holder = _RawCudaBuffer(ptr, nbytes, owner)
    t = torch.as_tensor(holder, device=device)
    t._b200ps_holder = holder  # keep the holder (and its owner) alive with the tensor
    return t


@dataclass
class SymmBuffer:
    """One symmetric allocation: ``nbytes`` on every rank."""
    name: str
    nbytes: int
    ptrs: List[int]                 # address of every rank's copy, valid in THIS process
    local: List[torch.Tensor]       # uint8 views of the copies this process owns (1 per local rank)
    local_ranks: List[int]
    mc_ptr: int = 0                 # NVSwitch multicast address of the buffer (0: not bound to a multicast object)

    def view(self, rank: int, dtype: torch.dtype, offset_bytes: int = 0, numel: Optional[int] = None):
        """Typed tensor over the copy of local rank ``rank``."""
        i = self.local_ranks.index(rank)
        esz = torch.empty((), dtype=dtype).element_size()
        if numel is None:
            numel = (self.nbytes - offset_bytes) // esz
        return self.local[i][offset_bytes: offset_bytes + numel * esz].view(dtype)


class SymmPool:
    """Allocator of symmetric buffers for a group of ``world`` ranks.

    ``rank`` is this process's rank; with ``virtual=True`` this single process plays all
    ``world`` ranks on one device.
    """

    def __init__(self, rank: int, world: int, device, group=None, virtual: bool = False):
        self.rank, self.world, self.device = rank, world, torch.device(device)
        self.group = group
        self.virtual = virtual or world == 1
        self.local_ranks = list(range(world)) if self.virtual else [rank]
        self._buffers = []
        import os
        # NVLS: off with B200PS_NVLS=0; switched off for the rest of the run by the first allocation that fails
        self.multicast_ok = os.environ.get("B200PS_NVLS", "1") != "0"
        self._opened: List[int] = []
        self._owned: List[int] = []
        if not self.virtual and group is None:
            raise ValueError("multi-process symmetric memory needs a control-plane process group")

    def _alloc_multicast(self, name: str, nbytes: int) -> Optional[SymmBuffer]:
        """A symmetric buffer that is ALSO bound to an NVSwitch multicast object, so kernels can use
        ``multimem.ld_reduce`` / ``multimem.st`` on it (csrc/kernels/ps_nvls.cu).  cudaMalloc + CUDA-IPC memory cannot be
        bound, so these buffers come from torch's symmetric-memory allocator (cuMemCreate + cuMulticastBindMem under
        the hood): one rendezvous over the control-plane group yields every rank's unicast address and the multicast
        address.  Returns None when the platform has no multicast support (the caller falls back to the IPC pool).
        All ranks take the same decision: the outcome is agreed on over the control plane."""
        import torch.distributed as dist
        ok, buf = 1, None
        try:
            import torch.distributed._symmetric_memory as symm_mem
            with torch.cuda.device(self.device):
                t = symm_mem.empty(nbytes, dtype=torch.uint8, device=self.device)
                hdl = symm_mem.rendezvous(t, self.group.group_name)
            mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
            if not mc:
                ok = 0
            else:
                t.zero_()
                torch.cuda.synchronize(self.device)
                buf = SymmBuffer(name, nbytes, [int(p) for p in hdl.buffer_ptrs], [t], [self.rank], mc_ptr=mc)
                buf._handle = hdl                     # keeps the mapping alive with the buffer
        except Exception as e:                        # no symmetric-memory support in this torch / on this box
            ok = 0
            self._mc_error = f"{type(e).__name__}: {e}"
        flag = torch.tensor([ok], dtype=torch.int32)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.group)
        if int(flag.item()) == 0:
            self.multicast_ok = False
            return None
        self._buffers.append(buf)
        dist.barrier(group=self.group)
        return buf

    def alloc(self, name: str, nbytes: int, multicast: bool = False) -> SymmBuffer:
        nbytes = (int(nbytes) + 255) // 256 * 256
        if multicast and not self.virtual and self.multicast_ok:
            buf = self._alloc_multicast(name, nbytes)
            if buf is not None:
                return buf
        if self.virtual:
            local = [torch.zeros(nbytes, dtype=torch.uint8, device=self.device) for _ in self.local_ranks]
            buf = SymmBuffer(name, nbytes, [t.data_ptr() for t in local], local, list(self.local_ranks))
            self._buffers.append(buf)
            return buf
        import torch.distributed as dist
        lib = native.load()
        with torch.cuda.device(self.device):
            ptr = C.c_void_p()
            handle = (C.c_ubyte * 64)()
            native.check(lib.b200ps_symm_alloc(nbytes, C.byref(ptr), handle), "symm_alloc")
        self._owned.append(ptr.value)
        handles: List[Optional[bytes]] = [None] * self.world
        dist.all_gather_object(handles, bytes(handle), group=self.group)
        ptrs = []
        with torch.cuda.device(self.device):
            for r in range(self.world):
                if r == self.rank:
                    ptrs.append(ptr.value)
                    continue
                p = C.c_void_p()
                hb = (C.c_ubyte * 64).from_buffer_copy(handles[r])
                native.check(lib.b200ps_symm_open(hb, C.byref(p)), f"symm_open(rank {r})")
                self._opened.append(p.value)
                ptrs.append(p.value)
        local = [tensor_from_ptr(ptr.value, nbytes, self.device, owner=self)]
        buf = SymmBuffer(name, nbytes, ptrs, local, [self.rank])
        self._buffers.append(buf)
        dist.barrier(group=self.group)   # every rank has mapped every copy before first use
        return buf

    def close(self):
        if self.virtual:
            self._buffers.clear()
            return
        lib = native.load()
        torch.cuda.synchronize(self.device)
        with torch.cuda.device(self.device):
            for p in self._opened:
                lib.b200ps_symm_close(p)
            self._opened.clear()
            if self.group is not None:
                import torch.distributed as dist
                try:
                    dist.barrier(group=self.group)  # nobody still maps our memory
                except Exception:
                    pass
            for p in self._owned:
                lib.b200ps_symm_free(p)
            self._owned.clear()
        self._buffers.clear()
