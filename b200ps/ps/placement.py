"""Variable placement over parameter-server shards (R3, P5, P6).

``round_robin_variable`` reproduces ``tf.train.replica_device_setter`` (example.py:65-67,
lr2.py:359-361): whole variables go to ps tasks round-robin in creation order, so with the
shipped creation order ``global_step, W1, W2, b1, b2`` and two ps tasks the owners are
ps0, ps1, ps0, ps1, ps0.  ``flat_shard`` is the B200-native layout: the flat
parameter space is cut into equal contiguous slices, one per GPU, regardless of variable
boundaries — this is what lets a single 4 GB ``[1e9,1]`` weight (run_lr2.sh:56) span all
GPUs, which whole-variable placement cannot do.

Both produce the same thing for the kernels: a flat element space in which each shard owns
one contiguous range ``[bounds[r], bounds[r+1])`` (multiples of ``ALIGN`` elements).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Sequence, Tuple

ALIGN = 8  # elements; keeps every variable 16-byte aligned in both fp32 and bf16


def _pad(n: int, a: int = ALIGN) -> int:
    return (n + a - 1) // a * a


@dataclass
class VarSpec:
    name: str
    shape: Tuple[int, ...]
    group: str = "fp32"            # "fp32" (master = working copy) | "bf16" (fp32 master + bf16 copy)
    init: str = "zeros"            # "zeros" | "normal" | "constant"
    stddev: float = 1.0
    mean: float = 0.0
    value: float = 0.0
    trainable: bool = True

    @property
    def numel(self) -> int:
        n = 1
        for s in self.shape:
            n *= int(s)
        return n


@dataclass
class Layout:
    placement: str
    num_shards: int
    total: int                                  # padded flat size (elements)
    bounds: List[int]                           # [num_shards + 1]
    offsets: Dict[str, int] = field(default_factory=dict)
    owners: Dict[str, int] = field(default_factory=dict)   # owner of a variable's FIRST element

    @property
    def max_shard(self) -> int:
        return max(self.bounds[r + 1] - self.bounds[r] for r in range(self.num_shards))

    def owner_of(self, elem: int) -> int:
        for r in range(self.num_shards):
            if self.bounds[r] <= elem < self.bounds[r + 1]:
                return r
        raise IndexError(elem)


def round_robin_owner(index_in_creation_order: int, num_ps: int) -> int:
    return index_in_creation_order % max(num_ps, 1)


def plan_layout(specs: Sequence[VarSpec], num_shards: int, placement: str,
                creation_index: Dict[str, int] = None) -> Layout:
    """Lay ``specs`` (one dtype group, creation order) out over ``num_shards`` owners.

    ``creation_index`` gives each variable's position among ALL variables of the model (the
    round-robin counter runs over every variable, including ones in other groups and
    non-trainable ones such as ``global_step``)."""
    num_shards = max(int(num_shards), 1)
    if placement == "flat_shard":
        offsets, off = {}, 0
        for s in specs:
            offsets[s.name] = off
            off += _pad(s.numel)
        shard = _pad((off + num_shards - 1) // num_shards) if off else ALIGN
        total = shard * num_shards
        bounds = [r * shard for r in range(num_shards + 1)]
        lay = Layout(placement, num_shards, total, bounds, offsets)
        for s in specs:
            lay.owners[s.name] = lay.owner_of(offsets[s.name])
        return lay
    if placement == "round_robin_variable":
        per_owner: List[List[VarSpec]] = [[] for _ in range(num_shards)]
        owners = {}
        for i, s in enumerate(specs):
            ci = creation_index[s.name] if creation_index else i
            o = round_robin_owner(ci, num_shards)
            owners[s.name] = o
            per_owner[o].append(s)
        offsets, bounds, off = {}, [0], 0
        for o in range(num_shards):
            for s in per_owner[o]:
                offsets[s.name] = off
                off += _pad(s.numel)
            bounds.append(off)
        total = max(off, ALIGN)
        if off == 0:
            bounds = [0] * num_shards + [0]
        return Layout(placement, num_shards, total, bounds, offsets, owners)
    raise ValueError(f"unknown placement {placement!r} (round_robin_variable | flat_shard)")
