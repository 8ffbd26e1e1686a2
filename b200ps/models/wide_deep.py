"""Sparse wide & deep logistic model (N5) — the extension sketched at lr2_debug.py:423-428.

The reference's scratch trainer adds, commented out, "two more layers" to the sparse LR::

    py_x  = embedding_lookup_sparse(W[F,1], ids, vals, 'sum')                      # wide (lr2.py:390)
    py_x += embedding_lookup_sparse(deep_w1[F,128], ids, vals, 'sum') @ deep_w2[128,1]

with ``init_weights(stddev=0.01)`` (lr2_debug.py:288-289) and the same sigmoid cross-entropy / SGD.
Here it is a working model: the three tables are row-sharded ``sparse`` variables (no replicated
copy anywhere) and one fused kernel (``b200ps_sparse_wd``) does gather, loss and all four Hogwild
updates per launch.  Variables, in creation order after ``global_step``: ``weights/Variable``
[F,1], ``deep/weights`` [F,D], ``deep/weights_1`` [D,1], ``bias/Variable`` [1].
"""
from __future__ import annotations

from typing import Dict, Optional

import torch

from .. import ops


def wide_deep_reference(W, E, w2, b, row_ptr, fid, fval, labels, train: bool = True):
    """float64 oracle.  Returns (loss, pred, grads) with dense grads dict {W, E, w2, b} (or None)."""
    B = labels.numel()
    n = int(row_ptr[-1])
    lens = (row_ptr[1:] - row_ptr[:-1]).long()
    row_of = torch.repeat_interleave(torch.arange(B), lens)
    fid = fid[:n].long()
    fval = fval[:n].double()
    E = E.double()
    w2 = w2.double().reshape(-1)
    h = torch.zeros(B, E.shape[1], dtype=torch.float64).index_add_(0, row_of, E[fid] * fval[:, None])
    z = h @ w2 + b.double().reshape(())
    if W is not None:
        z = z + torch.zeros(B, dtype=torch.float64).index_add_(0, row_of, W.reshape(-1).double()[fid] * fval)
    y = labels.double().reshape(-1)
    loss = (torch.clamp(z, min=0) - z * y + torch.log1p(torch.exp(-z.abs()))).mean()
    pred = torch.sigmoid(z)
    if not train:
        return loss, pred, None
    dz = (pred - y) / B
    grads = {"w2": h.T @ dz, "b": dz.sum(),
             "E": torch.zeros_like(E).index_add_(0, fid, (dz[row_of] * fval)[:, None] * w2[None, :])}
    if W is not None:
        grads["W"] = torch.zeros(W.shape[0], dtype=torch.float64).index_add_(0, fid, dz[row_of] * fval)
    return loss, pred, grads


class SparseWideDeep:
    W, E, W2, Bn = "weights/Variable", "deep/weights", "deep/weights_1", "bias/Variable"
    P, P2 = "_pending/coefficients", "_pending/deep_weights_1"      # scratch tables of the deferred update

    def __init__(self, features: int, dim: int = 128, wide: bool = True, stddev: float = 0.01,
                 deferred: Optional[bool] = None, owner: Optional[bool] = None):
        """``deferred`` (opt-in, also ``B200PS_WD_DEFERRED=1``): workers push one scalar coefficient
        per non-zero into a pending table and the owner of a shard folds them into W and E before its next
        step (the embedding update is rank one, dE[fid,:] = c[fid] * w2) — 4 instead of 516 bytes of remote
        atomics per non-zero."""
        import os
        if dim not in (128, 256, 512):
            raise ValueError("SparseWideDeep: dim must be 128, 256 or 512")
        self.features, self.dim, self.wide, self.stddev = int(features), int(dim), wide, stddev
        self.deferred = (os.environ.get("B200PS_WD_DEFERRED", "0") != "0") if deferred is None else bool(deferred)
        # owner-side step across GPUs (csrc/kernels/sparse_wd_owner.cu): rows never cross a link; the step is
        # COLLECTIVE (request / serve / loss rounds separated by cross-rank barriers).  Default for synchronous
        # replicas; for asynchronous workers it turns Hogwild into lockstep rounds of unaveraged updates, so there it
        # is opt-in (``owner=True`` / B200PS_WD_OWNER=1) — every rank must then take the same number of steps.
        env = os.environ.get("B200PS_WD_OWNER")
        self.owner = (None if env is None else env != "0") if owner is None else bool(owner)
        self._bufs: Dict[int, dict] = {}
        self._own: Dict[int, dict] = {}

    def register(self, store) -> None:
        if hasattr(store, "note_untracked_variable"):
            store.note_untracked_variable("global_step")
        if self.wide:
            store.register(self.W, (self.features, 1), group="sparse", init="normal", stddev=self.stddev)
        store.register(self.E, (self.features, self.dim), group="sparse", init="normal", stddev=self.stddev)
        store.register(self.W2, (self.dim, 1), group="sparse", init="normal", stddev=self.stddev)
        store.register(self.Bn, (1,), group="fp32", init="zeros")
        self._owner_capable = (hasattr(store, "sparse_shard") and store.placement == "flat_shard" and store.world > 1
                               and self.owner is not False)
        if (self.deferred or self._owner_capable) and hasattr(store, "sparse_shard") and store.placement == "flat_shard":
            # GPU store with row sharding only (the pending tables must be sharded exactly like W / E);
            # the gloo path and whole-variable placement apply their updates directly
            store.register(self.P, (self.features, 1), group="sparse", init="zeros", trainable=False)
            store.register(self.P2, (self.dim, 1), group="sparse", init="zeros", trainable=False)

    def names(self):
        return ((self.W,) if self.wide else ()) + (self.E, self.W2, self.Bn)

    def _buffers(self, B: int, device, key: int) -> dict:
        k = (B, key)
        if k not in self._bufs:
            self._bufs[k] = dict(pred=torch.empty(B, device=device), dz=torch.empty(B, device=device),
                                 loss=torch.zeros(1, device=device))
        return self._bufs[k]

    def _tables(self, store) -> dict:
        from ..parallel.symm import tensor_from_ptr
        if not hasattr(self, "_tab"):
            e, w2 = store.sparse_var(self.E), store.sparse_var(self.W2)
            w = store.sparse_var(self.W) if self.wide else None
            if w2.rows_per_shard % 4:
                raise ValueError("SparseWideDeep: w2 shard size must be a multiple of 4 rows")
            self._tab = dict(
                wide_table=None if w is None else w.table, wide_rows_per_shard=0 if w is None else w.rows_per_shard,
                emb_table=e.table, emb_rows_per_shard=e.rows_per_shard, dim=self.dim,
                w2_table=w2.table, w2_rows_per_shard=w2.rows_per_shard, features=self.features,
                bias=tensor_from_ptr(store.master_address(self.Bn), 4, store.device).view(torch.float32))
            if self.deferred and self.P in store.sparse:
                pe, p2 = store.sparse_var(self.P), store.sparse_var(self.P2)
                self._tab.update(pend_table=pe.table, pend_rows_per_shard=pe.rows_per_shard,
                                 pend2_table=p2.table, pend2_rows_per_shard=p2.rows_per_shard)
            self._has_pending = self.P in store.sparse
            self._gs0 = tensor_from_ptr(store.gstep.ptrs[0], 8, store.device).view(torch.int64)
        return self._tab

    def flush(self, store, ranks=None) -> None:
        """Deferred mode: fold the coefficients that have landed in the pending shards of ``ranks`` (default:
        every rank of this process) into their rows of W / E (phase 1, all ranks) and then their slices of w2
        (phase 2), so that every row update used the w2 of before.  A no-op in direct mode."""
        tabs = self._tables(store)
        if not getattr(self, "_has_pending", False) or not (self.deferred or getattr(self, "_owner_used", False)):
            return
        ranks = list(store.pool.local_ranks) if ranks is None else list(ranks)
        e, w2 = store.sparse_var(self.E), store.sparse_var(self.W2)
        for ph in (1, 2):
            for r in ranks:
                lo, hi = store._sparse_owned(e, r)
                lo2, hi2 = store._sparse_owned(w2, r)
                ops.sparse_wd_apply(store.sparse_shard(self.P, r), store.sparse_shard(self.W, r) if self.wide else None,
                                    store.sparse_shard(self.E, r), hi - lo, self.dim, w2.table, w2.rows_per_shard,
                                    store.sparse_shard(self.P2, r), store.sparse_shard(self.W2, r), hi2 - lo2, ph)

    # ---- owner-side step (collective): csrc/kernels/sparse_wd_owner.cu
    def _owner_buffers(self, store, cap: int) -> dict:
        if cap not in self._own:            # symmetric allocations: every rank gets here on its first (eager) step
            w, dev = store.world, store.device
            inbox = store.pool.alloc(f"wd.inbox.{cap}", w * cap * 8)
            cnt = store.pool.alloc(f"wd.inbox_cnt.{cap}", 4 * max(w, 64))
            resp = store.pool.alloc(f"wd.resp.{cap}", w * cap * 8)        # answers land in inbox order per owner
            pb = store.pool.alloc(f"wd.bias_pend.{cap}", 256)
            g = store._group_of(self.Bn)
            b_owner = g.layout.owner_of(g.layout.offsets[self.Bn])
            from ..parallel.symm import tensor_from_ptr
            self._own[cap] = dict(
                inbox=inbox, cnt=cnt, resp=resp, bias_pend_buf=pb, bias_owner=b_owner,
                bias_pend=tensor_from_ptr(pb.ptrs[b_owner], 4, dev).view(torch.float32),
                t_inbox=ops.device_ptr_table(inbox.ptrs, dev), t_cnt=ops.device_ptr_table(cnt.ptrs, dev),
                t_resp=ops.device_ptr_table(resp.ptrs, dev),
                count={r: torch.zeros(w + 1, dtype=torch.int32, device=dev) for r in store.pool.local_ranks},
                slot_tab={r: torch.zeros(cap, dtype=torch.int32, device=dev) for r in store.pool.local_ranks})
        return self._own[cap]

    def _step_owner(self, store, batches: dict, lr: float, sync: bool):
        """``batches``: {local rank: (row_ptr, fid, fval, labels)}.  Request / serve / loss rounds with the cross-rank
        barrier kernel in between, then the owners fold the pending coefficients (rows + dw2) and w2."""
        ranks = list(batches)
        self._owner_used = True
        e, w2 = store.sparse_var(self.E), store.sparse_var(self.W2)
        pe, p2 = store.sparse_var(self.P), store.sparse_var(self.P2)
        self._tables(store)
        N = store.world
        out = {}
        for r in ranks:
            row_ptr, fid, fval, labels = batches[r]
            ob = self._owner_buffers(store, fid.numel())
            ops.wd_owner_request(row_ptr, fid, world=N, rank=r, cap=fid.numel(), rows_per_shard=e.rows_per_shard,
                                 features=self.features, inbox_table=ob["t_inbox"], cnt_table=ob["t_cnt"],
                                 resp_table=ob["t_resp"], count=ob["count"][r], slot_tab=ob["slot_tab"][r])
        store.barrier()
        for r in ranks:
            ob = self._owner_buffers(store, batches[r][1].numel())
            ops.wd_owner_serve(world=N, rank=r, cap=batches[r][1].numel(), inbox_table=ob["t_inbox"], cnt_table=ob["t_cnt"],
                               resp_table=ob["t_resp"], emb_local=store.sparse_shard(self.E, r),
                               wide_local=store.sparse_shard(self.W, r) if self.wide else None, dim=self.dim,
                               w2_table=w2.table, w2_rows_per_shard=w2.rows_per_shard)
        store.barrier()
        for r in ranks:
            row_ptr, fid, fval, labels = batches[r]
            ob = self._owner_buffers(store, fid.numel())
            bufs = self._buffers(labels.numel(), labels.device, r)
            bufs["loss"].zero_()
            ops.wd_owner_loss(row_ptr, fid, fval, labels, world=N, rank=r, cap=fid.numel(),
                              rows_per_shard=e.rows_per_shard, features=self.features, resp_table=ob["t_resp"],
                              slot_tab=ob["slot_tab"][r], pend_table=pe.table, bias=self._tab["bias"], bias_pend=ob["bias_pend"],
                              pred=bufs["pred"], loss_sum=bufs["loss"],
                              global_step=store.global_step_tensor(r) if sync else self._gs0, update=True,
                              lr=lr / N if sync else lr)
            out[r] = bufs["loss"]
        store.barrier()                            # every coefficient has landed in its owner's pending table
        for r in ranks:                            # rows: driven by the inbox (the rows requested this round)
            ob = self._owner_buffers(store, batches[r][1].numel())
            ops.wd_owner_fold(world=N, rank=r, cap=batches[r][1].numel(), inbox_table=ob["t_inbox"], cnt_table=ob["t_cnt"],
                              pend_local=store.sparse_shard(self.P, r), emb_local=store.sparse_shard(self.E, r),
                              wide_local=store.sparse_shard(self.W, r) if self.wide else None, dim=self.dim,
                              w2_table=w2.table, pend2_table=p2.table, w2_rows_per_shard=w2.rows_per_shard)
        store.barrier()                            # every row used the old w2; every dw2 partial has landed
        self._fold_small(store, ranks)             # w2 and the bias; the next round's first barrier publishes them
        return out

    def _fold_small(self, store, ranks) -> None:
        """Owner-side step: fold the pending w2 vector (every rank its slice) and the pending bias (its owner)."""
        w2 = store.sparse_var(self.W2)
        for r in ranks:
            lo2, hi2 = store._sparse_owned(w2, r)
            ops.sparse_wd_apply(None, None, None, 0, self.dim, w2.table, w2.rows_per_shard,
                                store.sparse_shard(self.P2, r), store.sparse_shard(self.W2, r), hi2 - lo2, 2)
            for ob in self._own.values():
                if r == ob["bias_owner"]:
                    ops.sparse_wd_apply(None, None, None, 0, self.dim, w2.table, w2.rows_per_shard,
                                        ob["bias_pend_buf"].view(r, torch.float32), self._tab["bias"], 1, 2)

    def uses_owner_step(self, store, sync: bool) -> bool:
        if (not getattr(self, "_owner_capable", False) or store.device.type != "cuda"
                or len(store.pool.local_ranks) != 1):      # virtual ranks: call _step_owner with every rank's batch
            return False
        return True if (sync or self.owner) else False

    def step_async(self, store, batch, lr: float, rank: Optional[int] = None):
        """One fused Hogwild step — gather, loss, four scatter-updates, global_step — one launch."""
        if self.uses_owner_step(store, sync=False):
            r = store.pool.local_ranks[0] if rank is None else rank
            return self._step_owner(store, {r: batch}, lr, sync=False)[r]
        row_ptr, fid, fval, labels = batch
        bufs = self._buffers(labels.numel(), labels.device, 0 if rank is None else rank)
        bufs["loss"].zero_()
        self.flush(store, None if rank is None else [rank])   # owner side of the deferred update: what peers sent so far
        ops.sparse_wd(row_ptr, fid, fval, labels, **self._tables(store), pred=bufs["pred"],
                      loss_sum=bufs["loss"], global_step=self._gs0, forward=True, update=True, lr=lr)
        return bufs["loss"]

    def step_sync(self, store, batch, lr: float, rank: Optional[int] = None, replicas: Optional[int] = None):
        """Synchronous replicas: all read the same tables (barrier), then the mean gradient is applied."""
        if self.uses_owner_step(store, sync=True):
            r = store.pool.local_ranks[0] if rank is None else rank
            return self._step_owner(store, {r: batch}, lr, sync=True)[r]
        row_ptr, fid, fval, labels = batch
        bufs = self._buffers(labels.numel(), labels.device, 0 if rank is None else rank)
        bufs["loss"].zero_()
        N = replicas or store.world
        tabs = self._tables(store)
        ops.sparse_wd(row_ptr, fid, fval, labels, **tabs, pred=bufs["pred"], dz_out=bufs["dz"],
                      loss_sum=bufs["loss"], forward=True, update=False)
        store.barrier()
        ops.sparse_wd(row_ptr, fid, fval, labels, **tabs, dz_in=bufs["dz"],
                      global_step=store.global_step_tensor(rank), forward=False, update=True, lr=lr / N)
        store.barrier()
        if "pend_table" in tabs:                    # every replica's coefficients have landed: owners fold them in
            self.flush(store)
            store.barrier()
        return bufs["loss"]

    def predict(self, store, batch, rank: Optional[int] = None):
        row_ptr, fid, fval, labels = batch
        bufs = self._buffers(labels.numel(), labels.device, 0 if rank is None else rank)
        bufs["loss"].zero_()
        self.flush(store)
        ops.sparse_wd(row_ptr, fid, fval, labels, **self._tables(store), pred=bufs["pred"],
                      loss_sum=bufs["loss"], forward=True, update=False)
        return bufs["pred"], bufs["loss"]
