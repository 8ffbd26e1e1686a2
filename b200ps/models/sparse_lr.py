"""Sparse logistic regression (N5) — lr2.py:363-400.

``py_x = embedding_lookup_sparse(W, sp_fids, sp_fvals, combiner='sum') + b`` over a
``[features, 1]`` weight, mean sigmoid cross-entropy, plain SGD, streaming AUC for eval.
Variable names follow the checkpoint contract: ``weights/Variable``
``[features,1]`` ~ N(0,1) and ``bias/Variable`` ``[1]`` = 0, created after ``global_step``.

Batches are CSR (``row_ptr[B+1]``, ``fid[nnz]`` int64, ``fval[nnz]``) — the same information
as the reference's ``(sp_indices, fids, fvals)`` triple (lr2.py:69-84) without the redundant
row index per non-zero.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch

from .. import ops


def sparse_lr_reference(W: torch.Tensor, b: torch.Tensor, row_ptr, fid, fval, labels, lr: float = 0.0,
                        train: bool = True):
    """float64 restatement (CPU/gloo mode + oracle). Returns (loss, pred, dW as (ids, vals), db)."""
    B = labels.numel()
    n = int(row_ptr[-1])
    lens = (row_ptr[1:] - row_ptr[:-1]).long()
    row_of = torch.repeat_interleave(torch.arange(B), lens)
    fid = fid[:n].long()
    fval = fval[:n].double()
    z = torch.zeros(B, dtype=torch.float64).index_add_(0, row_of, W.reshape(-1).double()[fid] * fval)
    z = z + b.double().reshape(())
    y = labels.double().reshape(-1)
    loss = (torch.clamp(z, min=0) - z * y + torch.log1p(torch.exp(-z.abs()))).mean()
    pred = torch.sigmoid(z)
    if not train:
        return loss, pred, None, None
    dz = (pred - y) / B
    return loss, pred, (fid, dz[row_of] * fval), dz.sum()


class SparseLR:
    W, Bn = "weights/Variable", "bias/Variable"

    def __init__(self, features: int, stddev: float = 1.0, dedupe: Optional[bool] = None):
        import os
        self.features, self.stddev = int(features), float(stddev)
        # dedupe: make the batch's ids unique first (csrc/kernels/sparse_dedup.cu) — every distinct row is read once
        # and receives one update, as embedding_lookup_sparse + IndexedSlices do in the reference (lr2.py:390-396).
        # Pays when ids repeat inside a batch (click logs); with uniform ids the direct kernel is faster.
        self.dedupe = (os.environ.get("B200PS_SPARSE_DEDUPE", "0") != "0") if dedupe is None else bool(dedupe)
        self._bufs: Dict[int, dict] = {}
        self._dd: Dict[tuple, dict] = {}

    def register(self, store) -> None:
        if hasattr(store, "note_untracked_variable"):
            store.note_untracked_variable("global_step")
        store.register(self.W, (self.features, 1), group="sparse", init="normal", stddev=self.stddev)
        store.register(self.Bn, (1,), group="fp32", init="zeros")

    def _buffers(self, B: int, device, key: int) -> dict:
        k = (B, key)
        if k not in self._bufs:
            self._bufs[k] = dict(pred=torch.empty(B, device=device), dz=torch.empty(B, device=device),
                                 loss=torch.zeros(1, device=device))
        return self._bufs[k]

    def _dedup_scratch(self, cap: int, device, key: int) -> dict:
        """Per (rank, capacity): the id hash table (epoch-tagged keys, never cleared), the unique list and the
        gathered-weight / gradient buffers of the unique-id step."""
        k = (int(cap), key)
        if k not in self._dd:
            if torch.cuda.is_current_stream_capturing():
                # the epoch word must be set ONCE: a captured "state[1] = 1" would rewind it on every replay
                raise RuntimeError("SparseLR(dedupe=True): run one eager step (or predict) on this batch shape before "
                                   "capturing it into a CUDA graph")
            T = ops.sparse_dedup_table_size(cap)
            with torch.cuda.device(device):
                L = ops.sparse_dedup_list_size(cap)
                sms = torch.cuda.get_device_properties(device).multi_processor_count
            R = int(ops.native.load().b200ps_sparse_dedup_grad_copies())
            z = lambda n, dt: torch.zeros(n, dtype=dt, device=device)
            state = z(16, torch.int32)
            state[1] = 1                        # epoch 0 is what the zero-filled table holds
            self._dd[k] = dict(keys=z(T, torch.int64), slot_of=z(T, torch.int32), hpos=z(cap, torch.int32),
                               uniq=z(L, torch.int64), wbuf=z(L, torch.float32), gbuf=z(R * L, torch.float32),
                               seg_cnt=z(2 * sms, torch.int32), state=state, table_size=T, list_cap=L)
        return self._dd[k]

    def last_unique_counts(self, cap: int, key: int = 0):
        """(distinct ids, non-zeros) of the last finished unique-id step on that scratch (a device read)."""
        sc = self._dd[(int(cap), key)]
        return int(sc["seg_cnt"].sum().item()), int(sc["state"][3].item())

    def last_phase_times_ns(self, cap: int, key: int = 0):
        """CTA 0's globaltimer stamps of the last unique-id step as differences: (hash pass, sync, gather + convert,
        sync, examples, sync, scatter)."""
        st = [v & 0xffffffff for v in self._dd[(int(cap), key)]["state"].tolist()]
        t = st[4:12]
        return [(t[i + 1] - t[i]) & 0xffffffff for i in range(7)]

    def _run(self, store, batch, rank, **kw):
        """One launch of the step kernel: the direct one, or (dedupe) unique pass + cooperative step."""
        row_ptr, fid, fval, labels = batch
        sv = store.sparse_var(self.W)
        if not (self.dedupe and labels.device.type == "cuda"):
            kw.pop("unique", None), kw.pop("finish", None)
            return ops.sparse_lr(row_ptr, fid, fval, labels, sv.table, self._bias(store),
                                 rows_per_shard=sv.rows_per_shard, features=self.features, **kw)
        sc = self._dedup_scratch(fid.numel(), labels.device, 0 if rank is None else rank)
        return ops.sparse_lr_unique(row_ptr, fid, fval, labels, sv.table, self._bias(store), scratch=sc,
                                    rows_per_shard=sv.rows_per_shard, features=self.features, **kw)

    def _bias(self, store) -> torch.Tensor:
        from ..parallel.symm import tensor_from_ptr
        if not hasattr(self, "_bias_t"):
            self._bias_t = tensor_from_ptr(store.master_address(self.Bn), 4, store.device).view(torch.float32)
        return self._bias_t

    def step_async(self, store, batch, lr: float, rank: Optional[int] = None):
        """One fused Hogwild step: gather, loss, scatter-add, bias, global_step — one launch."""
        row_ptr, fid, fval, labels = batch
        sv = store.sparse_var(self.W)
        bufs = self._buffers(labels.numel(), labels.device, 0 if rank is None else rank)
        bufs["loss"].zero_()
        gs = store.gstep.ptrs[0]
        from ..parallel.symm import tensor_from_ptr
        if not hasattr(self, "_gs0"):
            self._gs0 = tensor_from_ptr(gs, 8, store.device).view(torch.int64)
        self._run(store, batch, rank, pred=bufs["pred"], loss_sum=bufs["loss"], global_step=self._gs0, forward=True,
                  update=True, lr=lr)
        return bufs["loss"]

    def step_sync(self, store, batch, lr: float, rank: Optional[int] = None, replicas: Optional[int] = None):
        """Synchronous replicas: every replica reads the same weights (barrier), then the mean of
        the replicas' gradients is applied (sum of -lr/N scatter-adds), then a second barrier."""
        row_ptr, fid, fval, labels = batch
        sv = store.sparse_var(self.W)
        bufs = self._buffers(labels.numel(), labels.device, 0 if rank is None else rank)
        bufs["loss"].zero_()
        N = replicas or store.world
        self._run(store, batch, rank, pred=bufs["pred"], dz_out=bufs["dz"], loss_sum=bufs["loss"], forward=True,
                  update=False, finish=False)   # (dedupe) the update launch below reuses this batch's unique list
        store.barrier()                       # nobody updates rows somebody is still reading
        gstep = store.global_step_tensor(rank)  # sync: every rank counts one step per aggregate
        self._run(store, batch, rank, dz_in=bufs["dz"], global_step=gstep, forward=False, update=True, lr=lr / N,
                  unique=False)
        store.barrier()                       # all updates landed before the next read
        return bufs["loss"]

    def predict(self, store, batch, rank: Optional[int] = None):
        """Forward only (the ``Test()`` helper and the AUC pass, lr2.py:307-315,455-468)."""
        row_ptr, fid, fval, labels = batch
        sv = store.sparse_var(self.W)
        bufs = self._buffers(labels.numel(), labels.device, 0 if rank is None else rank)
        bufs["loss"].zero_()
        self._run(store, batch, rank, pred=bufs["pred"], loss_sum=bufs["loss"], forward=True, update=False)
        return bufs["pred"], bufs["loss"]
