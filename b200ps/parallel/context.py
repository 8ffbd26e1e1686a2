"""Process context: who am I, which GPU do I drive, how do I reach my peers.

Control plane = ``torch.distributed`` (gloo over 127.0.0.1/TCP) — used only to bootstrap
(exchange CUDA IPC handles, ready/done barriers, checkpoint broadcast).  Data plane =
symmetric memory + the kernels of libb200ps.so; NCCL is never called on b200ps's own paths
(the PyTorch+NCCL baseline lives in ``baseline/ps_nccl.py``).

Launch conventions understood by :func:`init_from_env`:
* ``torchrun`` / ``python -m torch.distributed.run`` (RANK, LOCAL_RANK, WORLD_SIZE, MASTER_*),
* a single process (world = 1),
* ``virtual_world=N``: one process emulating N ranks on one device (tests).
"""
from __future__ import annotations

import datetime
import os
from dataclasses import dataclass
from typing import Optional

import torch

from .symm import SymmPool


class PeerLost(RuntimeError):
    """A rank of the job stopped answering on the control plane (see ``Context.check_peers``)."""


@dataclass
class Context:
    rank: int
    world: int
    device: torch.device
    pool: SymmPool
    group: Optional[object] = None       # gloo control-plane group (None when single process)
    virtual: bool = False

    @property
    def is_chief(self) -> bool:
        return self.rank == 0

    def control_barrier(self, timeout_s: float = 120.0) -> None:
        if self.group is not None:
            import torch.distributed as dist
            dist.monitored_barrier(group=self.group, timeout=datetime.timedelta(seconds=timeout_s))

    def all_reduce_max(self, value: float) -> float:
        """Max of a host scalar over the ranks (used for max-over-ranks device timings)."""
        if self.group is None:
            return float(value)
        import torch.distributed as dist
        t = torch.tensor([float(value)], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        return float(t.item())

    def all_reduce_sum(self, value: float) -> float:
        if self.group is None:
            return float(value)
        import torch.distributed as dist
        t = torch.tensor([float(value)], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return float(t.item())

    def check_peers(self, timeout_s: float = 5.0) -> None:
        """Liveness check over the control plane: returns once every rank has answered, raises ``PeerLost`` when
        one has not within ``timeout_s``.  On the NVLink backend every worker is also a parameter-server shard, so
        a dead worker takes its shard with it (the dead ps task of the reference, lr2.py:223-229): a synchronous
        step launched after that would spin on the dead rank's packets until its bounded in-kernel wait traps.
        Calling this between steps (apps: ``--heartbeat_steps``) turns the loss of a peer into an exception on
        the host BEFORE the next launch; recovery is the reference's — restart, the chief restores the newest
        checkpoint."""
        if self.group is None:
            return
        import torch.distributed as dist
        try:
            dist.monitored_barrier(group=self.group, timeout=datetime.timedelta(seconds=timeout_s),
                                   wait_all_ranks=False)
        except Exception as e:           # gloo reports the silent rank(s) in the message
            raise PeerLost(f"rank {self.rank}: a peer stopped answering within {timeout_s:.1f} s: "
                           f"{str(e).splitlines()[0][:200]}") from None

    def abandon(self) -> None:
        """Leave a job that lost a rank: no barriers (they cannot complete any more) — unmap what we mapped, free
        what we own, drop the process group.  The interpreter then exits normally."""
        if self.device.type == "cuda":
            try:
                torch.cuda.synchronize(self.device)
            except Exception:
                pass
        group, self.pool.group, self.group = self.group, None, None
        try:
            self.pool.close()
        except Exception:
            pass
        if group is not None:
            import torch.distributed as dist
            try:
                dist.destroy_process_group()
            except Exception:
                pass

    def close(self) -> None:
        """Release the symmetric memory: unmap the peers' buffers, control barrier, free our own."""
        self.pool.close()

    def shutdown(self) -> None:
        """Orderly end of a multi-process run: ``close()``, a last control barrier, then the control-plane
        process group is destroyed so that the interpreter exits normally (no ``os._exit``)."""
        import gc
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        gc.collect()
        self.close()
        if self.group is not None:
            import torch.distributed as dist
            try:
                self.control_barrier()
            finally:
                self.group = None
                if dist.is_initialized():
                    dist.destroy_process_group()


def init_from_env(virtual_world: int = 0, device: Optional[str] = None, timeout_s: float = 300.0) -> Context:
    if virtual_world and virtual_world > 0:
        dev = torch.device(device or ("cuda:0" if torch.cuda.is_available() else "cpu"))
        if dev.type == "cuda":
            torch.cuda.set_device(dev)
        return Context(0, virtual_world, dev, SymmPool(0, virtual_world, dev, virtual=True), None, True)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", str(rank)))
    if torch.cuda.is_available():
        dev = torch.device(device or f"cuda:{local_rank % torch.cuda.device_count()}")
        torch.cuda.set_device(dev)
    else:
        dev = torch.device("cpu")
    if world == 1:
        return Context(0, 1, dev, SymmPool(0, 1, dev, virtual=True), None, False)
    import torch.distributed as dist
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29512")
    if not dist.is_initialized():
        dist.init_process_group("gloo", rank=rank, world_size=world,
                                timeout=datetime.timedelta(seconds=timeout_s))
        group = dist.group.WORLD
    else:
        # someone (e.g. the NCCL baseline) already owns the default group: add a gloo one
        group = dist.new_group(backend="gloo", timeout=datetime.timedelta(seconds=timeout_s))
    return Context(rank, world, dev, SymmPool(rank, world, dev, group=group), group, False)
