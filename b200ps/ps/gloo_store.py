"""Worker-side client of the CPU/gloo parameter server (BASELINE.json config 1).

Same surface as :class:`b200ps.ps.store.ParameterStore` (register / build / initialize / var /
grad / pull / push_sync / push_async / global_step / state_dict), but the variables live in
the ``ps`` processes (``cluster/server.py``) and every pull/push is a message over gloo/TCP —
the reference's architecture (worker <-> ps star, example.py:64-118) kept as the plumbing
test bed.  Variables are placed on ps tasks with the same planner as the GPU store
(``round_robin_variable`` reproduces ``replica_device_setter``).
"""
from __future__ import annotations

import zlib
from typing import Dict, List, Optional

import torch

from ..cluster import server as S
from .placement import VarSpec, plan_layout


class GlooParameterStore:
    def __init__(self, server: "S.Server", placement: str = "round_robin_variable", seed: int = 1):
        if server.job_name != "worker":
            raise ValueError("GlooParameterStore is the worker-side client")
        self.server = server
        self.cluster = server.cluster
        self.num_ps = self.cluster.num_ps
        self.world = self.cluster.num_workers
        self.rank = server.task_index
        self.device = torch.device("cpu")
        self.placement, self.seed = placement, int(seed)
        self._specs: List[VarSpec] = []
        self._order: Dict[str, int] = {}
        self.local_step = 0            # last sync token seen (SyncReplicas local_step)
        self._built = False
        self.sparse: Dict[str, dict] = {}

    # ------------------------------------------------------------------ plumbing
    def _ps_rank(self, k: int) -> int:
        return self.cluster.global_rank("ps", k)

    def _hdr(self, k: int, op: int, a0=0, a1=0, a2=0, a3=0) -> None:
        import torch.distributed as dist
        dist.send(torch.tensor([op, self.rank, a0, a1, a2, a3], dtype=torch.int64), dst=self._ps_rank(k),
                  tag=S.TAG_HDR)

    def _send(self, k: int, t: torch.Tensor) -> None:
        import torch.distributed as dist
        if t.numel() > 0:                           # empty slices are never put on the wire
            dist.send(t.contiguous(), dst=self._ps_rank(k), tag=S.TAG_DATA)

    def _recv(self, k: int, n: int, dtype=torch.float32) -> torch.Tensor:
        import torch.distributed as dist
        t = torch.empty(n, dtype=dtype)
        if n > 0:
            dist.recv(t, src=self._ps_rank(k), tag=S.TAG_DATA)
        return t

    # ------------------------------------------------------------------ registration / build
    def note_untracked_variable(self, name: str) -> None:
        if name not in self._order:
            self._order[name] = len(self._order)

    def register(self, name: str, shape, *, group: str = "fp32", init: str = "zeros", stddev: float = 1.0,
                 mean: float = 0.0, value: float = 0.0, trainable: bool = True) -> None:
        if self._built:
            raise RuntimeError("register() after build()")
        if group == "bf16":
            group = "fp32"                                # the CPU path keeps everything in fp32
        self._order[name] = len(self._order)
        self._specs.append(VarSpec(name, tuple(int(s) for s in shape), group, init, stddev, mean, value, trainable))

    def build(self) -> "GlooParameterStore":
        dense = [s for s in self._specs if s.group == "fp32" and s.trainable]
        self.layout = plan_layout(dense, self.num_ps, self.placement, self._order)
        self.work = torch.zeros(self.layout.total)
        self.grads = torch.zeros(self.layout.total)
        for k in range(self.num_ps):
            lo, hi = self.layout.bounds[k], self.layout.bounds[k + 1]
            self._hdr(k, S.OP_REGISTER, 0, hi - lo)
        for vid, s in enumerate(x for x in self._specs if x.group == "sparse"):
            rows, dim = int(s.shape[0]), int(s.shape[1])
            if self.placement == "round_robin_variable":
                owner = self._order[s.name] % self.num_ps
                self.sparse[s.name] = dict(spec=s, vid=vid, rows=rows, rows_per_shard=rows, owner=owner, dim=dim)
            else:
                rps = (rows + self.num_ps - 1) // self.num_ps
                self.sparse[s.name] = dict(spec=s, vid=vid, rows=rows, rows_per_shard=rps, owner=None, dim=dim)
        self._built = True
        return self

    def variables(self) -> List[VarSpec]:
        return [s for s in self._specs if s.trainable and s.group == "fp32"]

    def owner_of(self, name: str) -> int:
        if name in self.sparse:
            o = self.sparse[name]["owner"]
            return 0 if o is None else o
        return self.layout.owners[name]

    def _spec(self, name: str) -> VarSpec:
        for s in self._specs:
            if s.name == name:
                return s
        raise KeyError(name)

    def var(self, name: str, rank=None) -> torch.Tensor:
        s = self._spec(name)
        off = self.layout.offsets[name]
        return self.work[off: off + s.numel].view(s.shape)

    def grad(self, name: str, rank=None) -> torch.Tensor:
        s = self._spec(name)
        off = self.layout.offsets[name]
        return self.grads[off: off + s.numel].view(s.shape)

    # ------------------------------------------------------------------ init / readiness (Supervisor)
    def _initial_value(self, s: VarSpec) -> torch.Tensor:
        if s.init == "normal":
            gen = torch.Generator().manual_seed(self.seed * 1000003 + (zlib.crc32(s.name.encode()) & 0x7FFFFFFF))
            return torch.randn(s.numel, generator=gen) * s.stddev + s.mean
        if s.init == "constant":
            return torch.full((s.numel,), float(s.value))
        return torch.zeros(s.numel)

    def initialize(self, state: Optional[Dict[str, torch.Tensor]] = None) -> None:
        """Chief only: run the initialisers (or load ``state``) and ship every slice to its ps."""
        for s in self.variables():
            v = self._initial_value(s) if state is None else torch.as_tensor(state[s.name], dtype=torch.float32).reshape(-1)
            off = self.layout.offsets[s.name]
            self.work[off: off + s.numel] = v
        for k in range(self.num_ps):
            lo, hi = self.layout.bounds[k], self.layout.bounds[k + 1]
            self._hdr(k, S.OP_INIT, 0, hi - lo, 0)
            self._send(k, self.work[lo:hi])
        for name, sv in self.sparse.items():
            s = sv["spec"]
            full = self._initial_value(s) if state is None else torch.as_tensor(state[name], dtype=torch.float32).reshape(-1)
            for k in range(self.num_ps):
                lo, hi = self._sparse_range(sv, k)
                if hi > lo:
                    self._hdr(k, S.OP_SPARSE_INIT, sv["vid"], hi - lo, sv["dim"])
                    self._send(k, full.reshape(-1)[lo * sv["dim"]:hi * sv["dim"]])
        step = 0 if state is None or "global_step" not in state else int(round(float(state["global_step"])))
        for k in range(self.num_ps):
            self._hdr(k, S.OP_SET_STEP, step)           # also flips the "initialized" flag
        self.local_step = 0

    def is_ready(self) -> bool:
        """Have the chief's initial values reached every ps?  (non-chief workers poll this)"""
        ok = True
        for k in range(self.num_ps):
            self._hdr(k, S.OP_IS_READY)
            r = self._recv(k, 2, torch.int64)
            ok = ok and bool(r[0].item())
        return ok

    def _sparse_range(self, sv: dict, k: int):
        if sv["owner"] is not None:
            return (0, sv["rows"]) if k == sv["owner"] else (0, 0)
        lo = min(sv["rows"], k * sv["rows_per_shard"])
        return lo, min(sv["rows"], lo + sv["rows_per_shard"])

    # ------------------------------------------------------------------ traffic
    def pull(self, groups=None, include_self: bool = False, min_step: int = 0) -> None:
        for k in range(self.num_ps):
            self._hdr(k, S.OP_PULL, 0, min_step)
        for k in range(self.num_ps):
            lo, hi = self.layout.bounds[k], self.layout.bounds[k + 1]
            self.work[lo:hi] = self._recv(k, hi - lo)
            gs = self._recv(k, 1, torch.int64)
            if k == 0:
                self._last_step = int(gs.item())

    def push_async(self, lr: float, rank=None, groups=None, count_step: bool = True) -> None:
        for k in range(self.num_ps):
            lo, hi = self.layout.bounds[k], self.layout.bounds[k + 1]
            self._hdr(k, S.OP_PUSH_ASYNC, 0, hi - lo, S.f2i(lr), int(count_step and k == 0))
            self._send(k, self.grads[lo:hi])

    def push_sync(self, lr: float, replicas_to_aggregate: Optional[int] = None, groups=None,
                  count_step: bool = True) -> None:
        R = replicas_to_aggregate or self.world
        packed = (self.local_step << 24) | (R << 16) | (self.world << 8) | 1
        active = list(range(self.num_ps))
        for k in active:
            lo, hi = self.layout.bounds[k], self.layout.bounds[k + 1]
            self._hdr(k, S.OP_PUSH_SYNC, 0, hi - lo, S.f2i(lr), packed)
            self._send(k, self.grads[lo:hi])
        step = self.local_step
        for k in active:                                   # block on a token from every ps (the barrier)
            tok = int(self._recv(k, 1, torch.int64).item())
            step = max(step, tok) if k != 0 else tok
        self.local_step = step
        self.pull(min_step=0)                              # fresh values for the next step

    def barrier(self) -> None:
        pass

    def zero_grads(self) -> None:
        self.grads.zero_()

    # sparse rows (E3/E4 of SURVEY §2.5.3): ids up, rows down; ids+values up
    def sparse_pull(self, name: str, ids: torch.Tensor) -> torch.Tensor:
        sv = self.sparse[name]
        dim = sv["dim"]
        out = torch.empty(ids.numel(), dim, dtype=torch.float32)
        owner = (ids // sv["rows_per_shard"]) if sv["owner"] is None else torch.full_like(ids, sv["owner"])
        local = ids - owner * sv["rows_per_shard"] if sv["owner"] is None else ids
        masks = []
        for k in range(self.num_ps):
            m = owner == k
            n = int(m.sum().item())
            masks.append((m, n))
            if n:
                self._hdr(k, S.OP_SPARSE_PULL, sv["vid"], n)
                self._send(k, local[m])
        for k, (m, n) in enumerate(masks):
            if n:
                out[m] = self._recv(k, n * dim).view(n, dim)
        return out.view(-1) if dim == 1 else out

    def sparse_push(self, name: str, ids: torch.Tensor, vals: torch.Tensor, lr: float) -> None:
        sv = self.sparse[name]
        owner = (ids // sv["rows_per_shard"]) if sv["owner"] is None else torch.full_like(ids, sv["owner"])
        local = ids - owner * sv["rows_per_shard"] if sv["owner"] is None else ids
        for k in range(self.num_ps):
            m = owner == k
            n = int(m.sum().item())
            if n:
                self._hdr(k, S.OP_SPARSE_PUSH, sv["vid"], n, S.f2i(lr))
                self._send(k, local[m])
                self._send(k, vals[m].float().reshape(-1))

    # ------------------------------------------------------------------ state
    @property
    def global_step(self) -> int:
        self._hdr(0, S.OP_GET_STEP)
        return int(self._recv(0, 1, torch.int64).item())

    def state_dict(self) -> Dict[str, torch.Tensor]:
        self.pull()
        out = {s.name: self.var(s.name).clone() for s in self.variables()}
        for name, sv in self.sparse.items():
            out[name] = self.sparse_pull(name, torch.arange(sv["rows"])).view(sv["spec"].shape)
        out["global_step"] = torch.tensor(float(self.global_step))
        return out

    def load_state_dict(self, state) -> None:
        self.initialize(state)

    def done(self) -> None:
        """Post this worker's done-token to every ps (lr2.py:337-346's intent)."""
        self.server.post_done()
