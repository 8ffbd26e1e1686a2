"""NVLS (NVSwitch multicast) buffers and the multimem push (K1-NVLS, csrc/kernels/ps_nvls.cu).

The b200ps pool (``parallel/symm.py``) maps peers' ``cudaMalloc`` memory through CUDA IPC; such memory cannot be
bound to a multicast object.  ``NvlsGroup`` allocates a gradient buffer and a working copy through torch's
symmetric-memory allocator instead — one rendezvous over the NCCL process group yields the peer pointers and the
*multicast* pointer of each buffer — and ``push_sync`` then runs the synchronous-replica update
(example.py:110-123: mean of the replicas' gradients, SGD on the owner's fp32 master, fresh values to every
replica) with ``multimem.ld_reduce`` / ``multimem.st``: the switch does the reduction and the fan-out, each owner
touches every element of its slice once.

    g = NvlsGroup(numel, lr)              # under torchrun, NCCL group initialised
    g.grads.copy_(my_gradients)           # each rank writes its own replica's gradients
    g.push_sync()                         # barrier, multimem kernel on the owned slice, barrier
    g.work                                # every rank now holds the updated parameters

``tools/nvls_probe.py`` checks that the box supports this at all.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from ..ops import native


class NvlsGroup:
    def __init__(self, numel: int, lr: float, group=None, init: torch.Tensor = None):
        import torch.distributed._symmetric_memory as symm_mem
        self.group = group or dist.group.WORLD
        self.rank, self.world = dist.get_rank(self.group), dist.get_world_size(self.group)
        dev = torch.device("cuda", torch.cuda.current_device())
        align = 4 * self.world
        self.numel = (int(numel) + align - 1) // align * align          # every owner slice a multiple of 4 elements
        self.lr = float(lr)
        self.grads = symm_mem.empty(self.numel, dtype=torch.float32, device=dev)
        self.work = symm_mem.empty(self.numel, dtype=torch.float32, device=dev)
        self._hg = symm_mem.rendezvous(self.grads, self.group.group_name)
        self._hw = symm_mem.rendezvous(self.work, self.group.group_name)
        if not int(self._hg.multicast_ptr) or not int(self._hw.multicast_ptr):
            raise RuntimeError("NVLS multicast is not available for symmetric memory on this system")
        shard = self.numel // self.world
        self.lo, self.hi = self.rank * shard, (self.rank + 1) * shard
        # fp32 master of the owned slice, indexable by the global element index (like the b200ps stores)
        self.master = torch.zeros(self.numel, dtype=torch.float32, device=dev)
        self.gstep = torch.zeros(1, dtype=torch.int64, device=dev)
        self.grads.zero_()
        self.work.zero_()
        if init is not None:
            self.work[:init.numel()].copy_(init.reshape(-1))
            self.master.copy_(self.work)
        torch.cuda.synchronize()
        self._hw.barrier()
        self._lib = native.load()

    def push_sync(self) -> None:
        """Mean-gradient SGD step on this rank's slice through the switch; all ranks call it together.
        The two symmetric-memory barriers and the kernel are all issued on torch's CURRENT stream (the barriers
        have no stream argument, so a launch on any other stream would be unordered against them)."""
        self._hg.barrier()                                   # every replica's gradients are written
        native.check(self._lib.b200ps_push_sync_nvls(int(self._hg.multicast_ptr), int(self._hw.multicast_ptr),
                                                     self.master.data_ptr(), self.lo, self.hi, self.lr / self.world,
                                                     self.gstep.data_ptr(), 0, native.stream_ptr(None)),
                     "push_sync_nvls")
        self._hw.barrier()                                   # every replica's working copy has the fresh values

    @property
    def global_step(self) -> int:
        return int(self.gstep.item())
