"""ParameterStore (N3): the parameter server, sharded across the GPUs' HBM.

In the reference, variables created under ``replica_device_setter`` live in the ps process
and every ``sess.run`` pulls them and pushes gradients over gRPC (example.py:65-118,
lr2.py:359-396).  Here rank k is worker k AND ps shard k:

* every variable is a slice of a flat, symmetric (peer-mapped) buffer per dtype group;
  shard ownership is a contiguous range per rank (``placement.py``);
* group ``fp32``: the fp32 working copy IS the master (small models, biases);
  group ``bf16``: fp32 master on the owner + bf16 working copy on every rank (wide layers);
* ``push_sync``  = K1 reduce-scatter + mean + SGD + broadcast + token barrier,
  ``push_async`` = K3 peer ``red.add`` (Hogwild) followed by ``pull`` when the next step starts,
  ``pull``       = all-gather of owner slices;
* ``global_step`` is an int64 counter on the chief's shard (P9); it is saved as an fp32
  scalar for checkpoint parity (example.py:70-72 declares it f32).

The same object drives one real GPU per process (multi-process, CUDA IPC) or ``world``
virtual ranks on one device (tests, world=1).
"""
from __future__ import annotations

import zlib
from dataclasses import dataclass
from typing import Dict, List, Optional

import torch

from .. import ops
from ..parallel.symm import SymmBuffer, SymmPool, tensor_from_ptr
from .placement import Layout, VarSpec, plan_layout

_DT = {"fp32": torch.float32, "bf16": torch.bfloat16}


@dataclass
class _Group:
    name: str
    dtype: torch.dtype
    specs: List[VarSpec]
    layout: Layout
    work: SymmBuffer
    grads: SymmBuffer
    master: Optional[SymmBuffer]     # bf16 group only (fp32 master of the owned range)
    bounds: torch.Tensor
    t_masters: torch.Tensor
    t_outs: torch.Tensor
    t_grads: torch.Tensor
    t_ll_grads: Optional[torch.Tensor] = None     # low-latency packet landing buffers (small fp32 groups)
    t_ll_params: Optional[torch.Tensor] = None


@dataclass
class _SparseVar:
    """A row-sharded [rows, dim] fp32 variable with NO replicated working copy: workers touch
    only the rows a batch names, through peer pointers (lr2.py:384 keeps W on the ps; dim > 1 is
    the embedding table of the wide&deep sketch, lr2_debug.py:423-428)."""
    spec: VarSpec
    rows: int
    rows_per_shard: int
    shards: SymmBuffer
    table: torch.Tensor
    dim: int = 1


def _stream_id(name: str) -> int:
    return zlib.crc32(name.encode()) & 0x7FFFFFFF


class ParameterStore:
    LL_MAX_ELEMS = 1 << 19     # fp32 groups up to 2 MB use the low-latency packet push
    NVLS_MIN_BYTES = 1 << 20   # dense groups of >= 1 MB live in multicast-bound memory and push through the switch
    NVLS_MAX_BYTES = 1 << 29   # ... up to 512 MB: at 1 GB the P2P kernel is faster again (profiles/pushpull_sweep_n8.json)

    def __init__(self, pool: SymmPool, placement: str = "flat_shard", seed: int = 1,
                 num_shards: Optional[int] = None):
        self.pool = pool
        self.world = pool.world
        self.rank = pool.rank
        self.device = pool.device
        self.placement = placement
        self.seed = int(seed)
        self.num_shards = num_shards or pool.world
        if self.num_shards != pool.world:
            raise ValueError("GPU store: one parameter-server shard per rank")
        self._specs: List[VarSpec] = []
        self._order: Dict[str, int] = {}
        self.groups: Dict[str, _Group] = {}
        self.sparse: Dict[str, _SparseVar] = {}
        self._built = False
        import os
        self.use_ll = os.environ.get("B200PS_LL_PUSH", "1") != "0"
        self.use_nvls = os.environ.get("B200PS_NVLS", "1") != "0"

    # ------------------------------------------------------------------ registration
    def register(self, name: str, shape, *, group: str = "fp32", init: str = "zeros",
                 stddev: float = 1.0, mean: float = 0.0, value: float = 0.0,
                 trainable: bool = True) -> None:
        if self._built:
            raise RuntimeError("register() after build()")
        if name in self._order:
            raise ValueError(f"variable {name!r} registered twice")
        if group.split(":")[0] not in _DT and group != "sparse":
            raise ValueError(f"unknown group {group!r}")
        if group == "sparse" and not (len(shape) == 2 and int(shape[1]) >= 1):
            raise ValueError("sparse variables are [rows, dim] (lr2.py:384, lr2_debug.py:423)")
        self._order[name] = len(self._order)          # creation order drives round-robin placement
        self._specs.append(VarSpec(name, tuple(int(s) for s in shape), group, init, stddev, mean, value,
                                   trainable))

    def note_untracked_variable(self, name: str) -> None:
        """Account for a variable that takes a round-robin slot but is not stored here
        (``global_step`` is created first in both reference scripts)."""
        if name not in self._order:
            self._order[name] = len(self._order)

    def owner_of(self, name: str) -> int:
        for g in self.groups.values():
            if name in g.layout.owners:
                return g.layout.owners[name]
        raise KeyError(name)

    # ------------------------------------------------------------------ build
    def build(self) -> "ParameterStore":
        pool, dev = self.pool, self.device
        self.pads = pool.alloc("signal_pads", 4 * ops.signal_pad_words())
        self.gstep = pool.alloc("global_step", 8)
        self.t_pads = ops.device_ptr_table(self.pads.ptrs, dev)
        self.t_gsteps = ops.device_ptr_table(self.gstep.ptrs, dev)
        gnames = []
        for s in self._specs:          # "bf16:<tag>" = a bucket of its own (per-layer push / fused pull)
            if s.group != "sparse" and s.group not in gnames:
                gnames.append(s.group)
        for gname in gnames:
            dtype = _DT[gname.split(":")[0]]
            specs = [s for s in self._specs if s.group == gname and s.trainable]
            if not specs:
                continue
            layout = plan_layout(specs, self.world, self.placement, self._order)
            esz = 4 if dtype == torch.float32 else 2
            # large dense groups: memory bound to an NVSwitch multicast object (when the box has one), so the
            # synchronous push / collective pull run as multimem.ld_reduce / multimem.st (K1-NVLS)
            mc = (self.world > 1 and layout.total * esz >= self.NVLS_MIN_BYTES
                  and all(b % 8 == 0 for b in layout.bounds))
            work = pool.alloc(f"{gname}.work", layout.total * esz, multicast=mc)
            grads = pool.alloc(f"{gname}.grads", layout.total * esz, multicast=mc)
            if dtype == torch.bfloat16:
                master = pool.alloc(f"{gname}.master", max(layout.max_shard, 8) * 4)
                master_ptrs = [master.ptrs[r] - 4 * layout.bounds[r] for r in range(self.world)]
            else:
                master = None
                master_ptrs = list(work.ptrs)
            grp = _Group(
                gname, dtype, specs, layout, work, grads, master,
                torch.tensor(layout.bounds, dtype=torch.int64, device=dev),
                ops.device_ptr_table(master_ptrs, dev), ops.device_ptr_table(work.ptrs, dev),
                ops.device_ptr_table(grads.ptrs, dev))
            if dtype == torch.float32 and self.world > 1 and layout.total <= self.LL_MAX_ELEMS:
                llg = pool.alloc(f"{gname}.ll_grads", self.world * max(layout.max_shard, 8) * 8)
                llp = pool.alloc(f"{gname}.ll_params", layout.total * 8)
                grp.t_ll_grads = ops.device_ptr_table(llg.ptrs, dev)
                grp.t_ll_params = ops.device_ptr_table(llp.ptrs, dev)
                grp._ll_bufs = (llg, llp)
            self.groups[gname] = grp
        for s in self._specs:
            if s.group != "sparse":
                continue
            rows, dim = int(s.shape[0]), int(s.shape[1])
            if self.placement == "round_robin_variable":
                # whole variable on one ps task (the reference's only option)
                owner = self._order[s.name] % self.world
                rps = rows
                shards = pool.alloc(f"sparse.{s.name}", max(rows * dim, 8) * 4)
                ptrs = [shards.ptrs[owner]] * self.world    # every lookup resolves to the owner
                table = ops.device_ptr_table(ptrs, dev)
            else:
                rps = (rows + self.world - 1) // self.world
                p2 = 1
                while p2 < rps:
                    p2 <<= 1
                rps = p2 if p2 * self.world < rows * 2 and p2 * (self.world - 1) < rows else ((rps + 3) // 4 * 4)
                shards = pool.alloc(f"sparse.{s.name}", max(rps * dim, 8) * 4)
                table = ops.device_ptr_table(shards.ptrs, dev)
            self.sparse[s.name] = _SparseVar(s, rows, rps, shards, table, dim)
        self._built = True
        return self

    def sparse_var(self, name: str) -> _SparseVar:
        return self.sparse[name]

    def sparse_shard(self, name: str, rank: Optional[int] = None) -> torch.Tensor:
        sv = self.sparse[name]
        return sv.shards.view(self.pool.local_ranks[0] if rank is None else rank, torch.float32,
                              0, sv.rows_per_shard * sv.dim)

    def _sparse_owned(self, sv: _SparseVar, r: int):
        """Global row range stored in rank r's shard buffer."""
        if self.placement == "round_robin_variable":
            owner = self._order[sv.spec.name] % self.world
            return (0, sv.rows) if r == owner else (0, 0)
        lo = min(sv.rows, r * sv.rows_per_shard)
        return lo, min(sv.rows, lo + sv.rows_per_shard)

    def master_address(self, name: str) -> int:
        """Address (valid in this process) of a dense fp32 variable's master copy on its owner."""
        g = self._group_of(name)
        if g.master is not None:
            raise NotImplementedError
        owner = g.layout.owner_of(g.layout.offsets[name])
        return g.work.ptrs[owner] + 4 * g.layout.offsets[name]

    # ------------------------------------------------------------------ views
    def _group_of(self, name: str) -> _Group:
        for g in self.groups.values():
            if name in g.layout.offsets:
                return g
        raise KeyError(f"unknown variable {name!r}")

    def _spec(self, name: str) -> VarSpec:
        for s in self._specs:
            if s.name == name:
                return s
        raise KeyError(name)

    def _view(self, buf: SymmBuffer, g: _Group, name: str, rank: Optional[int]) -> torch.Tensor:
        rank = self.pool.local_ranks[0] if rank is None else rank
        s = self._spec(name)
        esz = 4 if g.dtype == torch.float32 else 2
        return buf.view(rank, g.dtype, g.layout.offsets[name] * esz, s.numel).view(s.shape)

    def var(self, name: str, rank: Optional[int] = None) -> torch.Tensor:
        """Working copy of a variable (what the model kernels read)."""
        g = self._group_of(name)
        return self._view(g.work, g, name, rank)

    def grad(self, name: str, rank: Optional[int] = None) -> torch.Tensor:
        """Gradient slot of a variable inside the symmetric push buffer."""
        g = self._group_of(name)
        return self._view(g.grads, g, name, rank)

    def flat_work(self, group: str, rank: Optional[int] = None) -> torch.Tensor:
        g = self.groups[group]
        return g.work.view(self.pool.local_ranks[0] if rank is None else rank, g.dtype)

    def flat_grads(self, group: str, rank: Optional[int] = None) -> torch.Tensor:
        g = self.groups[group]
        return g.grads.view(self.pool.local_ranks[0] if rank is None else rank, g.dtype)

    def variables(self) -> List[VarSpec]:
        return [s for s in self._specs if s.trainable and s.group != "sparse"]

    # ------------------------------------------------------------------ init / state
    def initialize(self) -> None:
        """Run the initialisers.  Values are a pure function of (seed, variable name, element),
        so every rank can fill its copy — and its owned master slice — without communication;
        the Supervisor still gates workers on the chief's "ready" (P10)."""
        for s in self.variables():
            full = self._initial_value(s)
            self._assign(s, full)
        for sv in self.sparse.values():
            for r in self.pool.local_ranks:
                lo, hi = self._sparse_owned(sv, r)
                if hi <= lo:
                    continue
                dst = sv.shards.view(r, torch.float32, 0, (hi - lo) * sv.dim)
                if sv.spec.init == "normal":
                    ops.init_normal(dst, mean=sv.spec.mean, stddev=sv.spec.stddev, seed=self.seed,
                                    stream_id=_stream_id(sv.spec.name), elem_offset=lo * sv.dim)
                else:
                    dst.fill_(sv.spec.value if sv.spec.init == "constant" else 0.0)
        for r in self.pool.local_ranks:
            self.gstep.view(r, torch.int64)[0] = 0
        torch.cuda.synchronize(self.device) if self.device.type == "cuda" else None

    def _initial_value(self, s: VarSpec) -> torch.Tensor:
        t = torch.empty(s.numel, dtype=torch.float32, device=self.device)
        if s.init == "normal":
            if self.device.type == "cuda":
                ops.init_normal(t, mean=s.mean, stddev=s.stddev, seed=self.seed, stream_id=_stream_id(s.name))
            else:
                gen = torch.Generator().manual_seed(self.seed * 1000003 + _stream_id(s.name))
                t = torch.randn(s.numel, generator=gen) * s.stddev + s.mean
        elif s.init == "constant":
            t.fill_(s.value)
        else:
            t.zero_()
        return t

    def _assign(self, s: VarSpec, full_fp32: torch.Tensor) -> None:
        g = self._group_of(s.name)
        off = g.layout.offsets[s.name]
        for r in self.pool.local_ranks:
            self._view(g.work, g, s.name, r).view(-1).copy_(full_fp32.to(g.dtype))
            if g.master is not None:
                lo, hi = g.layout.bounds[r], g.layout.bounds[r + 1]
                a, b = max(lo, off), min(hi, off + s.numel)
                if a < b:
                    g.master.view(r, torch.float32)[a - lo: b - lo].copy_(full_fp32[a - off: b - off])

    def state_dict(self) -> Dict[str, torch.Tensor]:
        """Full fp32 value of every variable, gathered from the shard owners (host tensors)."""
        out = {}
        for s in self.variables():
            g = self._group_of(s.name)
            off = g.layout.offsets[s.name]
            if g.master is None:
                # fp32 group: the owner's working copy is authoritative; read it through the mapping
                full = torch.empty(s.numel, dtype=torch.float32)
                for r in range(self.world):
                    lo, hi = g.layout.bounds[r], g.layout.bounds[r + 1]
                    a, b = max(lo, off), min(hi, off + s.numel)
                    if a < b:
                        src = tensor_from_ptr(g.work.ptrs[r] + 4 * a, 4 * (b - a), self.device).view(torch.float32)
                        full[a - off: b - off] = src.cpu()
            else:
                full = torch.empty(s.numel, dtype=torch.float32)
                for r in range(self.world):
                    lo, hi = g.layout.bounds[r], g.layout.bounds[r + 1]
                    a, b = max(lo, off), min(hi, off + s.numel)
                    if a < b:
                        src = tensor_from_ptr(g.master.ptrs[r] + 4 * (a - lo), 4 * (b - a), self.device).view(torch.float32)
                        full[a - off: b - off] = src.cpu()
            out[s.name] = full.view(s.shape)
        for name, sv in self.sparse.items():
            if name.startswith("_"):                  # scratch tables (e.g. pending coefficients): not model state
                continue
            full = torch.empty(sv.rows * sv.dim, dtype=torch.float32)
            for r in range(self.world):
                lo, hi = self._sparse_owned(sv, r)
                if hi > lo:
                    src = tensor_from_ptr(sv.shards.ptrs[r], 4 * (hi - lo) * sv.dim, self.device).view(torch.float32)
                    full[lo * sv.dim:hi * sv.dim] = src.cpu()
            out[name] = full.view(sv.spec.shape)
        out["global_step"] = torch.tensor(float(self.global_step), dtype=torch.float32)
        return out

    def load_state_dict(self, state: Dict[str, torch.Tensor]) -> None:
        for s in self.variables():
            if s.name not in state:
                raise KeyError(f"checkpoint lacks variable {s.name!r}")
            t = torch.as_tensor(state[s.name], dtype=torch.float32)
            if tuple(t.shape) != tuple(s.shape):
                raise ValueError(f"{s.name}: checkpoint shape {tuple(t.shape)} != {s.shape}")
            self._assign(s, t.reshape(-1).to(self.device))
        for name, sv in self.sparse.items():
            if name.startswith("_"):
                continue
            t = torch.as_tensor(state[name], dtype=torch.float32).reshape(-1)
            if t.numel() != sv.rows * sv.dim:
                raise ValueError(f"{name}: checkpoint has {t.numel()} elements, want {sv.rows * sv.dim}")
            for r in self.pool.local_ranks:
                lo, hi = self._sparse_owned(sv, r)
                if hi > lo:
                    sv.shards.view(r, torch.float32, 0, (hi - lo) * sv.dim).copy_(
                        t[lo * sv.dim:hi * sv.dim].to(self.device))
        if "global_step" in state:
            for r in self.pool.local_ranks:
                self.gstep.view(r, torch.int64)[0] = int(round(float(state["global_step"])))

    # ------------------------------------------------------------------ global_step
    @property
    def global_step(self) -> int:
        """The chief's counter (rank 0 shard)."""
        if 0 in self.pool.local_ranks:
            return int(self.gstep.view(0, torch.int64)[0].item())
        return int(tensor_from_ptr(self.gstep.ptrs[0], 8, self.device).view(torch.int64)[0].item())

    def global_step_tensor(self, rank: Optional[int] = None) -> torch.Tensor:
        return self.gstep.view(self.pool.local_ranks[0] if rank is None else rank, torch.int64)

    # ------------------------------------------------------------------ traffic
    def _ranks(self):
        return self.pool.local_ranks[0], len(self.pool.local_ranks)

    def zero_grads(self) -> None:
        for g in self.groups.values():
            for r in self.pool.local_ranks:
                g.grads.view(r, g.dtype).zero_()

    def _nvls(self, g: "_Group", prefer: Optional[bool] = None) -> bool:
        """Does this group's push / collective pull go through the switch?  Needs multicast-bound buffers; by default
        taken for 1 MB .. 512 MB groups (measured window in which it beats the P2P kernel on 8 GPUs); ``prefer``
        overrides the size rule either way."""
        if not (g.work.mc_ptr and g.grads.mc_ptr and len(self.pool.local_ranks) == 1 and self.use_nvls):
            return False
        if prefer is not None:
            return bool(prefer)
        return g.work.nbytes <= self.NVLS_MAX_BYTES

    def _store_nvls(self, g: "_Group", lr: float, reduce: bool, broadcast: bool, gs, cursor, cursor_step, cursor_wrap):
        r = self.pool.local_ranks[0]
        master = (g.work.ptrs[r] if g.master is None else g.master.ptrs[r] - 4 * g.layout.bounds[r])
        ops.store_nvls(self.t_pads, g.grads.mc_ptr if reduce else 0, g.work.mc_ptr, g.work.ptrs[r], master,
                       g.layout.bounds[r], g.layout.bounds[r + 1], dtype=g.dtype, world=self.world, rank=r, lr=lr,
                       reduce=reduce, broadcast=broadcast, global_step=None if gs is None else self.gstep.ptrs[r],
                       cursor=cursor, cursor_step=cursor_step, cursor_wrap=cursor_wrap)

    def pull(self, groups=None, include_self: bool = False, collective: bool = False) -> None:
        """Refresh the working copies from the owners' masters.  Default: every rank READS the owners' slices over
        NVLink whenever it likes (what an asynchronous worker does before a step).  ``collective=True`` (all ranks
        call it together): groups in multicast memory are broadcast by their OWNERS with ``multimem.st`` — every
        byte leaves its owner once and the switch fans it out."""
        base, n = self._ranks()
        for g in self._iter_groups(groups):
            if collective and self._nvls(g, True):
                self._store_nvls(g, 0.0, False, True, None, None, 0, 1)
                continue
            ops.pull(g.t_masters, g.t_outs, g.bounds, world=self.world, rank_base=base, local_ranks=n,
                     max_shard_elems=g.layout.max_shard, dtype=g.dtype,
                     include_self=include_self or g.master is not None)

    def push_sync(self, lr: float, replicas_to_aggregate: Optional[int] = None, groups=None,
                  count_step: bool = True, cursor=None, cursor_step: int = 0, cursor_wrap: int = 1,
                  broadcast: bool = True, nvls: Optional[bool] = None) -> None:
        base, n = self._ranks()
        R = replicas_to_aggregate or self.world
        gl = list(self._iter_groups(groups))
        for i, g in enumerate(gl):
            # one global_step per aggregated update: only the last group's launch counts it
            gs = self.t_gsteps if (count_step and i == len(gl) - 1) else None
            if R >= self.world and self._nvls(g, nvls):
                self._store_nvls(g, lr, True, broadcast, gs, cursor if i == len(gl) - 1 else None, cursor_step, cursor_wrap)
                continue
            if g.t_ll_grads is not None and R >= self.world and broadcast and self.use_ll:
                ops.push_sync_ll(self.t_pads, g.t_masters, g.t_grads, g.bounds, gs, g.t_ll_grads, g.t_ll_params,
                                 world=self.world, rank_base=base, local_ranks=n,
                                 max_shard_elems=max(g.layout.max_shard, 8), total_elems=g.layout.total, lr=lr,
                                 cursor=cursor if i == len(gl) - 1 else None, cursor_step=cursor_step,
                                 cursor_wrap=cursor_wrap)
                continue
            ops.push_sync_sgd(self.t_pads, g.t_masters, g.t_outs, g.t_grads, g.bounds, gs,
                              world=self.world, rank_base=base, local_ranks=n,
                              max_shard_elems=g.layout.max_shard, dtype=g.dtype,
                              replicas_to_aggregate=R, lr=lr,
                              cursor=cursor if i == len(gl) - 1 else None, cursor_step=cursor_step,
                              cursor_wrap=cursor_wrap, broadcast=broadcast)

    def push_async(self, lr: float, rank: Optional[int] = None, groups=None, count_step: bool = True,
                   cursor=None, cursor_step: int = 0, cursor_wrap: int = 1) -> None:
        """Hogwild push of ONE worker's gradients (fp32 groups only)."""
        gl = list(self._iter_groups(groups))
        r = self.pool.local_ranks[0] if rank is None else rank
        for i, g in enumerate(gl):
            if g.master is not None:
                raise NotImplementedError("async push is defined for fp32 groups")
            gs = self.t_gsteps if (count_step and i == len(gl) - 1) else None
            ops.push_async(g.t_masters, g.t_grads, g.bounds, gs, world=self.world, rank_base=r,
                           total_elems=g.layout.total, lr=lr,
                           cursor=cursor if i == len(gl) - 1 else None, cursor_step=cursor_step,
                           cursor_wrap=cursor_wrap)

    def gather_info(self, name: str, rank: Optional[int] = None, ctas: int = 0) -> dict:
        """Arguments of the fused pull (``ops.gemm(..., gather=...)``) for a [K,N] bf16 variable that
        is the only member of its group: every rank's working copy + the local panel flags."""
        g = self._group_of(name)
        s = self._spec(name)
        if g.dtype != torch.bfloat16 or len(g.specs) != 1 or len(s.shape) != 2:
            raise ValueError("fused pull needs a 2-D bf16 variable in a group of its own")
        rank = self.pool.local_ranks[0] if rank is None else rank
        key = (name, rank)
        if not hasattr(self, "_gather"):
            self._gather = {}
        if key not in self._gather:
            self._gather[key] = torch.zeros(s.shape[0] // 64 + 1, dtype=torch.int32, device=self.device)
        return dict(src_table=g.t_outs, flags=self._gather[key], world=self.world, rank=rank, ctas=ctas)

    def barrier(self) -> None:
        base, n = self._ranks()
        ops.barrier(self.t_pads, world=self.world, rank_base=base, local_ranks=n)

    def _iter_groups(self, groups):
        for name, g in self.groups.items():
            if groups is None or name in groups:
                yield g
