"""Replica optimizers (N4) over a ParameterStore.

* :class:`AsyncSGD` — ``GradientDescentOptimizer.minimize`` under between-graph replication
  (example.py:108,118; lr2.py:395-396): every worker pushes ``-lr*g`` into the owning shard
  whenever it finishes a batch, unlocked (Hogwild); ``global_step`` counts ALL workers' updates.
* :class:`SyncReplicasSGD` — ``tf.train.SyncReplicasOptimizer`` (example.py:110-123,147-152):
  the mean of ``replicas_to_aggregate`` gradients is applied once per step, ``global_step``
  advances by one per aggregated update, and no replica starts step t+1 before the update of
  step t has landed (the token queue becomes the two flag exchanges inside the push kernel;
  the chief's queue-runner duties are the shard owners' reduce+apply).
"""
from __future__ import annotations

from typing import Optional


class AsyncSGD:
    def __init__(self, learning_rate: float):
        self.learning_rate = float(learning_rate)
        self.mode = "async"

    def prepare(self, store, rank: Optional[int] = None) -> None:
        """Refresh the worker's view of peer-owned slices (stale reads are fine)."""
        store.pull()

    def apply(self, store, rank: Optional[int] = None, **cursor_kw) -> None:
        store.push_async(self.learning_rate, rank=rank, **cursor_kw)


class SyncReplicasSGD:
    def __init__(self, learning_rate: float, replicas_to_aggregate: Optional[int] = None,
                 total_num_replicas: Optional[int] = None):
        self.learning_rate = float(learning_rate)
        self.replicas_to_aggregate = replicas_to_aggregate
        self.total_num_replicas = total_num_replicas
        self.mode = "sync"

    def prepare(self, store, rank: Optional[int] = None) -> None:
        pass   # push_sync broadcast the fresh values already: the pull is fused into the push

    def apply(self, store, rank: Optional[int] = None, **cursor_kw) -> None:
        total = self.total_num_replicas or store.world
        if total != store.world:
            raise ValueError(f"total_num_replicas={total} but {store.world} workers are running")
        store.push_sync(self.learning_rate, self.replicas_to_aggregate or total, **cursor_kw)
