"""Server (R2, P3, P4): the per-task runtime endpoint — ``tf.train.Server(cluster, job_name, task_index)``.

Reference behaviour (example.py:38-40,53-55; lr2.py:331-335): every task starts an in-process
server bound to its ``host:port``; a ``ps`` task then calls ``server.join()`` and serves
variable reads / gradient applies forever; workers open sessions on ``server.target``.

Two transports implement that contract here:

* ``nvlink`` (GPU mode): workers are ranks of one NVLink process group and EACH WORKER IS ALSO
  A PARAMETER-SERVER SHARD (symmetric memory, kernels in libb200ps.so).  A ``ps`` task is
  still a valid role — so the reference's 4-command launch recipe keeps working — but owns no
  tensors: ``join()`` just waits for the workers' done-tokens.
* ``gloo`` (CPU mode, BASELINE.json config 1): a real ps process holds its variables and runs
  the service loop below over gloo/TCP: INIT, PULL, PUSH_ASYNC (unlocked apply per message =
  Hogwild at message granularity), PUSH_SYNC (SyncReplicas accumulate / mean / apply / token
  release / stale-gradient drop), sparse row PULL/PUSH (the gather / ScatterSub of lr2.py:390,396),
  GET_STEP, DONE.

``join()`` returns once every worker has posted a done-token — the graceful-shutdown protocol
the reference only sketches in comments (lr2.py:329,337-346).
"""
from __future__ import annotations

import datetime
import struct
from typing import Dict, List

import torch

from ..io import log
from .spec import ClusterSpec

# message opcodes (header = int64[6]: op, worker_task, a0, a1, a2, a3)
OP_INIT, OP_IS_READY, OP_PULL, OP_PUSH_ASYNC, OP_PUSH_SYNC, OP_GET_STEP, OP_DONE = 1, 2, 3, 4, 5, 6, 7
OP_SPARSE_PULL, OP_SPARSE_PUSH, OP_REGISTER, OP_SET_STEP, OP_SPARSE_INIT = 8, 9, 10, 11, 12
HDR = 6
TAG_HDR, TAG_DATA = 11, 12


def f2i(x: float) -> int:
    return struct.unpack("<q", struct.pack("<d", float(x)))[0]


def i2f(x: int) -> float:
    return struct.unpack("<d", struct.pack("<q", int(x)))[0]


def _cap_cpu_threads(cluster) -> None:
    """Tasks started by hand on one host (the reference's recipe) would each spin one intra-op thread per core on
    matrices of a few thousand elements and fight for the cores (gloo backend, 1 ps + 2 workers on 8 cores:
    54 ms/step; 9 ms with the cap).  When the user has not chosen (OMP_NUM_THREADS) and every task of the cluster
    lives on this host, keep cores / (2 * tasks) threads."""
    import os
    if "OMP_NUM_THREADS" in os.environ:
        return
    addrs = [a for tasks in cluster.as_dict().values() for a in tasks]
    if {a.rsplit(":", 1)[0] for a in addrs} <= {"localhost", "127.0.0.1"}:
        tasks = len(addrs)
        torch.set_num_threads(max(1, min(torch.get_num_threads(), (os.cpu_count() or 1) // max(1, 2 * tasks))))


class Server:
    def __init__(self, cluster, job_name: str, task_index: int, backend: str = "auto", start: bool = True,
                 timeout_s: float = 300.0):
        self.cluster = cluster if isinstance(cluster, ClusterSpec) else ClusterSpec(cluster)
        if job_name not in ("ps", "worker"):
            raise ValueError(f"job_name must be 'ps' or 'worker', got {job_name!r}")
        self.job_name, self.task_index = job_name, int(task_index)
        self.address = self.cluster.task_address(job_name, task_index)
        if backend == "auto":
            backend = "nvlink" if torch.cuda.is_available() else "gloo"
        self.backend = backend
        self.timeout_s = timeout_s
        self.rank = self.cluster.global_rank(job_name, task_index)
        self.world = self.cluster.world_size()
        self.group = None
        log.set_role(job_name, task_index)
        if start:
            self.start()

    @property
    def target(self) -> str:
        """``server.target`` (example.py:145): what a session attaches to."""
        return f"b200ps+{self.backend}://{self.address}"

    def start(self) -> None:
        """Join the control-plane process group (every ps and worker task is one rank)."""
        import torch.distributed as dist
        if self.group is not None:
            return
        host, port = self.cluster.rendezvous()
        _cap_cpu_threads(self.cluster)
        if not dist.is_initialized():
            dist.init_process_group("gloo", init_method=f"tcp://{host}:{port}", rank=self.rank,
                                    world_size=self.world,
                                    timeout=datetime.timedelta(seconds=self.timeout_s))
        self.group = dist.group.WORLD
        # workers-only subgroup (collective: every task creates it) — the NVLink process group
        nw = self.cluster.num_workers
        self.worker_group = dist.new_group(ranks=list(range(nw)), backend="gloo") if self.cluster.num_ps else self.group

    def post_done(self) -> None:
        """Worker: send this task's done-token to every ps (lr2.py:337-346's intent)."""
        import torch.distributed as dist
        for k in range(self.cluster.num_ps):
            dist.send(torch.tensor([OP_DONE, self.task_index, 0, 0, 0, 0], dtype=torch.int64),
                      dst=self.cluster.global_rank("ps", k), tag=TAG_HDR)

    def nvlink_context(self, device=None):
        """GPU mode: this worker's Context (rank = task_index, symmetric-memory pool over the workers)."""
        from ..parallel.context import Context
        from ..parallel.symm import SymmPool
        nw = self.cluster.num_workers
        dev = torch.device(device or f"cuda:{self.task_index % max(torch.cuda.device_count(), 1)}")
        torch.cuda.set_device(dev)
        if nw == 1:
            return Context(0, 1, dev, SymmPool(0, 1, dev, virtual=True), None, False)
        return Context(self.task_index, nw, dev, SymmPool(self.task_index, nw, dev, group=self.worker_group),
                       self.worker_group, False)

    # ------------------------------------------------------------------ ps side
    def join(self) -> None:
        """Block serving requests until all workers have sent their done-token."""
        if self.job_name != "ps":
            raise RuntimeError("join() is the parameter-server loop; workers do not call it")
        _PsService(self).serve()

    def shutdown(self) -> None:
        import torch.distributed as dist
        if dist.is_initialized():
            try:
                dist.destroy_process_group()
            except Exception:
                pass
        self.group = None


class _PsService:
    """State and request loop of one ps task (gloo transport)."""

    def __init__(self, server: Server):
        self.s = server
        self.num_workers = server.cluster.num_workers
        self.slices: Dict[int, torch.Tensor] = {}       # group id -> this ps's flat fp32 slice
        self.sparse: Dict[int, torch.Tensor] = {}       # var id -> rows held here
        self.initialized = False
        self.global_step = 0
        self.applied_steps = 0                          # aggregated sync updates applied here
        self.done = 0
        # sync-replica state per group
        self.acc: Dict[int, torch.Tensor] = {}
        self.acc_count: Dict[int, int] = {}
        self.tokens = 0
        self.waiting: List[int] = []                    # worker ranks blocked on a token
        self.deferred_pulls: List[tuple] = []           # (worker_rank, gid, min_step)

    def _send(self, t: torch.Tensor, dst: int) -> None:
        import torch.distributed as dist
        if t.numel() > 0:
            dist.send(t.contiguous(), dst=dst, tag=TAG_DATA)

    def _recv(self, n: int, src: int, dtype=torch.float32) -> torch.Tensor:
        import torch.distributed as dist
        t = torch.empty(n, dtype=dtype)
        if n > 0:                                   # empty slices are never put on the wire
            dist.recv(t, src=src, tag=TAG_DATA)
        return t

    def _reply_pull(self, wrank: int, gid: int) -> None:
        self._send(self.slices[gid], wrank)
        self._send(torch.tensor([self.global_step], dtype=torch.int64), wrank)

    def _release_tokens(self) -> None:
        while self.tokens > 0 and self.waiting:
            w = self.waiting.pop(0)
            self.tokens -= 1
            self._send(torch.tensor([self.global_step], dtype=torch.int64), w)

    def serve(self) -> None:
        import torch.distributed as dist
        log.info("start ...")
        hdr = torch.zeros(HDR, dtype=torch.int64)
        while self.done < self.num_workers:
            src = dist.recv(hdr, src=None, tag=TAG_HDR)
            op, _wtask, a0, a1, a2, a3 = (int(v) for v in hdr.tolist())
            if op == OP_REGISTER:          # a0 = group id, a1 = slice length held by this ps
                if a0 not in self.slices:
                    self.slices[a0] = torch.zeros(a1, dtype=torch.float32)
            elif op == OP_INIT:            # chief pushes initial values of this ps's slice
                self.slices[a0] = self._recv(a1, src)
                self.initialized = bool(a2) or self.initialized
            elif op == OP_SPARSE_INIT:     # a0 = var id, a1 = rows held here, a2 = row width (0 = 1)
                dim = max(a2, 1)
                self.sparse[a0] = self._recv(a1 * dim, src).view(a1, dim)
            elif op == OP_SET_STEP:
                self.global_step = a0
                self.initialized = True
            elif op == OP_IS_READY:
                self._send(torch.tensor([1 if self.initialized else 0, self.global_step], dtype=torch.int64), src)
            elif op == OP_PULL:            # a0 = gid, a1 = minimum applied sync step the reply must reflect
                if a1 > self.applied_steps:
                    self.deferred_pulls.append((src, a0, a1))
                else:
                    self._reply_pull(src, a0)
            elif op == OP_PUSH_ASYNC:      # a0 = gid, a1 = n, a2 = lr bits, a3 = count global_step
                g = self._recv(a1, src)
                self.slices[a0].add_(g, alpha=-i2f(a2))          # ApplyGradientDescent, unlocked
                if a3:
                    self.global_step += 1
            elif op == OP_PUSH_SYNC:       # a0 = gid, a1 = n, a2 = lr bits, a3 = packed(local_step, R, total, count)
                g = self._recv(a1, src)
                local_step, R, total, count = a3 >> 24, (a3 >> 16) & 0xFF, (a3 >> 8) & 0xFF, a3 & 1
                if local_step >= self.applied_steps:             # fresh gradient
                    if a0 not in self.acc:
                        self.acc[a0] = torch.zeros_like(self.slices[a0])
                        self.acc_count[a0] = 0
                    self.acc[a0].add_(g)
                    self.acc_count[a0] += 1
                    if self.acc_count[a0] == R:
                        self.slices[a0].add_(self.acc[a0], alpha=-i2f(a2) / R)
                        self.acc[a0].zero_()
                        self.acc_count[a0] = 0
                        if count:                                  # last group of the step on this ps
                            self.applied_steps += 1
                            self.global_step += 1
                            self.tokens += total
                            still = []
                            for (w, gid, ms) in self.deferred_pulls:
                                if ms <= self.applied_steps:
                                    self._reply_pull(w, gid)
                                else:
                                    still.append((w, gid, ms))
                            self.deferred_pulls = still
                # else: stale gradient (computed before the last aggregate) is dropped
                if count:
                    self.waiting.append(src)
                    self._release_tokens()
            elif op == OP_SPARSE_PULL:     # a0 = var id, a1 = number of ids (local row indices)
                ids = self._recv(a1, src, torch.int64)
                self._send(self.sparse[a0][ids].reshape(-1), src)
            elif op == OP_SPARSE_PUSH:     # a0 = var id, a1 = n ids, a2 = lr bits
                ids = self._recv(a1, src, torch.int64)
                dim = self.sparse[a0].shape[1]
                vals = self._recv(a1 * dim, src).view(a1, dim)
                self.sparse[a0].index_add_(0, ids, vals, alpha=-i2f(a2))   # ScatterSub (duplicates summed)
            elif op == OP_GET_STEP:
                self._send(torch.tensor([self.global_step], dtype=torch.int64), src)
            elif op == OP_DONE:
                self.done += 1
                log.info(f"done-token {self.done}/{self.num_workers}")
            else:
                raise RuntimeError(f"ps: unknown opcode {op} from rank {src}")
        log.info("Terminating parameter server: {}".format(self.s.task_index))
