"""MLP models (N5): the MNIST sigmoid MLP of example.py and the wide bf16 stress model.

``SigmoidMLP`` is example.py:77-128: x[.,784] -> sigmoid(x W1 + b1) -> a2 W2 + b2 -> softmax,
mean cross-entropy, plain SGD; accuracy = mean(argmax == label).  Variable names, shapes and
creation order follow the checkpoint contract of SURVEY §2.7
(``weights/Variable`` [784,100], ``weights/Variable_1`` [100,10], ``biases/Variable`` [100],
``biases/Variable_1`` [10]); weights are stored [in,out] like the reference.

GPU step (per worker), all native kernels:
    z2   = x W1            tcgen05 tf32 GEMM, split-K with fp32 red.add          (K2/K8)
    tail = bias+sigmoid, head GEMM, softmax-CE, accuracy, dZ2, dW2, db2, db1     (K6)
    dW1  = x^T dZ2         tcgen05 tf32 GEMM straight into the push buffer       (K7)
    push = sync reduce-scatter+SGD+broadcast  |  async peer red.add              (K1/K3)
Deviation kept from SURVEY §7.1: the loss uses a stable log-softmax instead of the
reference's ``-sum(y * log(softmax))`` (example.py:98,103), equal wherever that is finite.

``WideMLP`` is BASELINE.json config 4 (8 x 8192 x 8192, bf16, sigmoid, softmax-CE head).
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch

from .. import ops


# --------------------------------------------------------------------------------------
# plain PyTorch restatement: CPU/gloo mode and the test oracle
# --------------------------------------------------------------------------------------
def mlp_reference_step(params: Dict[str, torch.Tensor], x: torch.Tensor, labels: torch.Tensor,
                       names=("weights/Variable", "weights/Variable_1", "biases/Variable",
                              "biases/Variable_1"), train: bool = True):
    """fp32 forward (+ manual backward) of the sigmoid MLP.  Returns (loss, grads, correct)."""
    w1, w2, b1, b2 = (params[n].float() for n in names)
    x = x.float()
    z2 = x @ w1 + b1
    a2 = torch.sigmoid(z2)
    z3 = a2 @ w2 + b2
    logp = torch.log_softmax(z3, dim=1)
    B = x.shape[0]
    idx = labels.long()
    loss = -logp[torch.arange(B), idx].mean()
    correct = int((z3.argmax(1) == idx).sum().item())
    if not train:
        return loss, None, correct
    dz3 = torch.softmax(z3, dim=1)
    dz3[torch.arange(B), idx] -= 1.0
    dz3 /= B
    grads = {names[1]: a2.t() @ dz3, names[3]: dz3.sum(0)}
    dz2 = (dz3 @ w2.t()) * a2 * (1 - a2)
    grads[names[0]] = x.t() @ dz2
    grads[names[2]] = dz2.sum(0)
    return loss, grads, correct


class SigmoidMLP:
    W1, W2, B1, B2 = "weights/Variable", "weights/Variable_1", "biases/Variable", "biases/Variable_1"

    def __init__(self, in_dim: int = 784, hidden: int = 100, classes: int = 10, split_k: int = 5):
        self.in_dim, self.hidden, self.classes = in_dim, hidden, classes
        self.split_k = split_k
        self._bufs: Dict[Tuple[int, int], dict] = {}

    # -- variables in the reference's creation order (example.py:70-90) --------------------
    def register(self, store) -> None:
        if hasattr(store, "note_untracked_variable"):
            store.note_untracked_variable("global_step")
        store.register(self.W1, (self.in_dim, self.hidden), init="normal", stddev=1.0)
        store.register(self.W2, (self.hidden, self.classes), init="normal", stddev=1.0)
        store.register(self.B1, (self.hidden,), init="zeros")
        store.register(self.B2, (self.classes,), init="zeros")

    def names(self):
        return (self.W1, self.W2, self.B1, self.B2)

    def _buffers(self, batch: int, device, key: int) -> dict:
        k = (batch, key)
        if k not in self._bufs:
            self._bufs[k] = dict(
                z2=torch.zeros(max(self.split_k, 1), batch, self.hidden, device=device),
                dz2=torch.empty(batch, self.hidden, device=device),
                loss=torch.zeros(1, device=device),
                correct=torch.zeros(1, device=device, dtype=torch.int32))
        return self._bufs[k]

    # -- GPU path -------------------------------------------------------------------------
    def forward_backward(self, store, x: torch.Tensor, labels: torch.Tensor, rank: Optional[int] = None,
                         global_batch: Optional[int] = None, row_offset: Optional[torch.Tensor] = None,
                         batch: Optional[int] = None, stats_out=None, fused: bool = False):
        """Forward + backward of one worker's batch; gradients land in ``store.grad(...)``.
        Returns the (device) loss tensor and the correct-count tensor (both overwritten per step).
        ``stats_out=(loss, correct)`` redirects the two words, e.g. into pinned mapped host memory.

        With ``row_offset`` (int32 device scalar) ``x``/``labels`` are a whole resident dataset and
        the batch is rows ``[*row_offset, *row_offset + batch)`` — read on the device, no copies."""
        B = x.shape[0] if row_offset is None else batch
        bufs = self._buffers(B, x.device, 0 if rank is None else rank)
        z2, dz2, loss, correct = bufs["z2"], bufs["dz2"], bufs["loss"], bufs["correct"]
        if stats_out is not None:
            loss, correct = stats_out
        w1 = store.var(self.W1, rank)
        if fused and self.supports_fused(B):
            # K13: the whole forward/backward in one launch (exact fp32), gradients for the push
            ops.mlp_fused_step(x, labels, w1, store.var(self.B1, rank), store.var(self.W2, rank),
                               store.var(self.B2, rank), loss=loss, correct=correct, batch=B, cursor=row_offset,
                               grads=(store.grad(self.W1, rank), store.grad(self.B1, rank),
                                      store.grad(self.W2, rank), store.grad(self.B2, rank)))
            return loss, correct
        off = dict(a_row_offset=row_offset, a_rows=B) if row_offset is not None else {}
        # split-K over CTAs, each writing its own slab of z2; the head kernel sums the slabs in a
        # fixed order (an atomic split-K would make the step irreproducible bit for bit)
        ksplits = min(max(self.split_k, 1), -(-self.in_dim // 32))
        ops.gemm(x, w1, z2[0], b_mn_major=True, split_k=ksplits, c_split_stride=z2.stride(0), **off)
        ops.mlp_tail(z2[0], store.var(self.B1, rank), store.var(self.W2, rank), store.var(self.B2, rank),
                     labels, loss_sum=loss, correct=correct, dz2=dz2, dw2=store.grad(self.W2, rank),
                     db2=store.grad(self.B2, rank), db1=store.grad(self.B1, rank), train=True,
                     grad_scale=1.0 / B, loss_scale=1.0 / B, overwrite_stats=True, label_offset=row_offset,
                     z2_splits=ksplits, z2_split_stride=z2.stride(0))
        ops.gemm(x, dz2, store.grad(self.W1, rank), a_mn_major=True, b_mn_major=True, **off)
        return loss, correct

    def supports_fused(self, batch: int) -> bool:
        """Shapes the one-launch step kernel (K13, csrc/kernels/mlp_fused.cu) covers."""
        return (batch <= 128 and self.hidden <= 128 and self.hidden % 4 == 0 and self.classes <= 16
                and self.in_dim % 16 == 0 and self.hidden * self.classes <= 2048)

    def fused_train_step(self, store, x: torch.Tensor, labels: torch.Tensor, lr: float, rank: Optional[int] = None,
                         row_offset: Optional[torch.Tensor] = None, batch: Optional[int] = None, stats_out=None,
                         cursor: Optional[torch.Tensor] = None, cursor_step: int = 0, cursor_wrap: int = 1):
        """Single-replica step entirely in K13: forward, backward, SGD in place, ``global_step += 1`` and the
        resident-dataset cursor advance — one launch (what example.py:168 asks of one ``sess.run``)."""
        B = x.shape[0] if row_offset is None else batch
        bufs = self._buffers(B, x.device, 0 if rank is None else rank)
        loss, correct = stats_out if stats_out is not None else (bufs["loss"], bufs["correct"])
        ops.mlp_fused_step(x, labels, store.var(self.W1, rank), store.var(self.B1, rank), store.var(self.W2, rank),
                           store.var(self.B2, rank), loss=loss, correct=correct, batch=B, cursor=row_offset, lr=lr,
                           global_step=store.global_step_tensor(rank), cursor_rw=cursor, cursor_step=cursor_step,
                           cursor_wrap=cursor_wrap)
        return loss, correct

    def supports_cluster(self, batch: int) -> bool:
        """Shapes (and devices) the resident-cluster kernel (K14, csrc/kernels/mlp_cluster.cu) covers."""
        return ops.mlp_cluster_supported(batch, self.in_dim, self.hidden, self.classes)

    def cluster_train_steps(self, store, x: torch.Tensor, labels: torch.Tensor, lr: float, steps: int = 1,
                            rank: Optional[int] = None, row_offset: Optional[torch.Tensor] = None,
                            batch: Optional[int] = None, stats_out=None, stat_stride: int = 0,
                            cursor: Optional[torch.Tensor] = None, cursor_step: Optional[int] = None,
                            cursor_wrap: Optional[int] = None):
        """``steps`` consecutive single-replica steps in ONE launch of K14 (one 16-CTA cluster, parameters resident
        in shared memory, tensor-core products, DSMEM exchanges): what ``steps`` ``sess.run(train_op)`` calls of
        example.py:168 do.  ``x``/``labels`` hold the batches back to back (or a whole dataset walked through
        ``row_offset``); step ``s`` writes ``stats_out[0][s*stat_stride]`` / ``stats_out[1][s*stat_stride]``."""
        B = x.shape[0] if (row_offset is None and steps == 1 and batch is None) else batch
        bufs = self._buffers(B, x.device, 0 if rank is None else rank)
        loss, correct = stats_out if stats_out is not None else (bufs["loss"], bufs["correct"])
        ops.mlp_cluster_steps(x, labels, store.var(self.W1, rank), store.var(self.B1, rank), store.var(self.W2, rank),
                              store.var(self.B2, rank), loss=loss, correct=correct, stat_stride=stat_stride,
                              steps=steps, lr=lr, batch=B, cursor=row_offset,
                              global_step=store.global_step_tensor(rank), cursor_rw=cursor,
                              cursor_step=cursor_step, cursor_wrap=cursor_wrap)
        return loss, correct

    def supports_fused_push(self, store) -> bool:
        """Can the synchronous push run inside K13?  One rank per process, all four variables alone in one
        fp32 group that has low-latency landing buffers."""
        try:
            g = store._group_of(self.W1)
        except Exception:
            return False
        names = {s.name for s in g.specs}
        return (store.world > 1 and len(store.pool.local_ranks) == 1 and g.t_ll_grads is not None and store.use_ll
                and names == set(self.names()) and g.layout.offsets[self.W1] % 4 == 0)

    def supports_fused_async(self, store) -> bool:
        """Can the asynchronous worker step run entirely inside K13 (parameters read from, and updates added
        to, the owners' masters)?  All four variables in one fp32 group with 4-aligned offsets."""
        try:
            g = store._group_of(self.W1)
        except Exception:
            return False
        off = g.layout.offsets
        return (g.master is None and {s.name for s in g.specs} == set(self.names())
                and all(off[n] % 4 == 0 for n in (self.W1, self.W2, self.B1)))

    def _push_tables(self, store, r: int, bufs: dict):
        from ..ops import native
        if "push" not in bufs:
            g = store._group_of(self.W1)
            off = g.layout.offsets
            bufs["push"] = native.FusedPush(
                pads=native.ptr(store.t_pads), masters=native.ptr(g.t_masters), bounds=native.ptr(g.bounds),
                ll_grads=native.ptr(g.t_ll_grads), ll_params=native.ptr(g.t_ll_params),
                global_steps=native.ptr(store.t_gsteps), max_shard_elems=max(g.layout.max_shard, 8),
                off_w1=off[self.W1], off_b1=off[self.B1], off_w2=off[self.W2], off_b2=off[self.B2],
                world=store.world, rank=r)
        return bufs["push"]

    def fused_async_step(self, store, x: torch.Tensor, labels: torch.Tensor, lr: float, rank: Optional[int] = None,
                         row_offset: Optional[torch.Tensor] = None, batch: Optional[int] = None, stats_out=None,
                         cursor: Optional[torch.Tensor] = None, cursor_step: int = 0, cursor_wrap: int = 1):
        """Asynchronous (Hogwild) worker step as ONE launch: parameters fetched from the owners' masters
        inside the kernel (the pull fused into the first GEMM), ``-lr * g`` added into them unlocked, the
        chief's ``global_step`` incremented — example.py:108,118 without a local parameter copy."""
        B = x.shape[0] if row_offset is None else batch
        r = store.pool.local_ranks[0] if rank is None else rank
        bufs = self._buffers(B, x.device, r)
        loss, correct = stats_out if stats_out is not None else (bufs["loss"], bufs["correct"])
        ops.mlp_fused_step(x, labels, store.var(self.W1, r), store.var(self.B1, r), store.var(self.W2, r),
                           store.var(self.B2, r), loss=loss, correct=correct, batch=B, cursor=row_offset, lr=lr,
                           cursor_rw=cursor, cursor_step=cursor_step, cursor_wrap=cursor_wrap,
                           push=self._push_tables(store, r, bufs), hogwild=True)
        return loss, correct

    def fused_push_step(self, store, x: torch.Tensor, labels: torch.Tensor, lr: float, rank: Optional[int] = None,
                        row_offset: Optional[torch.Tensor] = None, batch: Optional[int] = None, stats_out=None,
                        cursor: Optional[torch.Tensor] = None, cursor_step: int = 0, cursor_wrap: int = 1):
        """Synchronous-replica step as ONE launch per rank: forward, backward, and the low-latency push
        (gradient packets to the owners, reduce + SGD + publish, unpack) inside K13 — the fused
        compute + collective form of example.py:110-123."""
        B = x.shape[0] if row_offset is None else batch
        r = store.pool.local_ranks[0] if rank is None else rank
        bufs = self._buffers(B, x.device, r)
        loss, correct = stats_out if stats_out is not None else (bufs["loss"], bufs["correct"])
        ops.mlp_fused_step(x, labels, store.var(self.W1, r), store.var(self.B1, r), store.var(self.W2, r),
                           store.var(self.B2, r), loss=loss, correct=correct, batch=B, cursor=row_offset, lr=lr,
                           cursor_rw=cursor, cursor_step=cursor_step, cursor_wrap=cursor_wrap,
                           push=self._push_tables(store, r, bufs))
        return loss, correct

    def evaluate(self, store, x: torch.Tensor, labels: torch.Tensor, rank: Optional[int] = None):
        """Accuracy and mean loss over a (possibly large) batch, e.g. the 10 000-image test set
        in one call like example.py:187."""
        B = x.shape[0]
        z2 = torch.empty(B, self.hidden, device=x.device)
        loss = torch.zeros(1, device=x.device)
        correct = torch.zeros(1, device=x.device, dtype=torch.int32)
        ops.gemm(x, store.var(self.W1, rank), z2, b_mn_major=True)
        ops.mlp_tail(z2, store.var(self.B1, rank), store.var(self.W2, rank), store.var(self.B2, rank),
                     labels, loss_sum=loss, correct=correct, train=False, loss_scale=1.0 / B)
        return correct.item() / B, loss.item()


class WideMLP:
    """``layers`` dense layers of ``hidden x hidden`` (bf16 working copy, fp32 master), sigmoid
    between layers, softmax cross-entropy over the last layer's ``hidden`` outputs.

    Every weight matrix is a parameter-server bucket of its own (group ``bf16:<l>``), row-sharded
    over the ranks, which is what lets the multi-GPU step overlap its traffic with its math:

    * forward: layer l's GEMM pulls the other ranks' row shards of W_l *inside the launch* (K2,
      gather CTAs next to the GEMM CTAs) — no separate all-gather, no broadcast in the push;
    * backward: as soon as dW_l is written, its reduce-scatter + SGD (K1, ``broadcast=False``) is
      queued on a side stream and runs under the remaining backward GEMMs.
    """

    def __init__(self, layers: int = 8, hidden: int = 8192, fuse_sgd_in_wgrad: bool = True):
        self.layers, self.hidden = layers, hidden
        self.fuse_sgd_in_wgrad = fuse_sgd_in_wgrad
        self._acts = {}

    def wname(self, l: int) -> str:
        return "weights/Variable" + (f"_{l}" if l else "")

    def bname(self, l: int) -> str:
        return "biases/Variable" + (f"_{l}" if l else "")

    def wgroup(self, l: int) -> str:
        return f"bf16:{l}"

    def register(self, store) -> None:
        if hasattr(store, "note_untracked_variable"):
            store.note_untracked_variable("global_step")
        std = 4.0 / (self.hidden ** 0.5)   # keeps sigmoid layers away from saturation
        for l in range(self.layers):
            store.register(self.wname(l), (self.hidden, self.hidden), group=self.wgroup(l), init="normal", stddev=std)
        for l in range(self.layers):
            store.register(self.bname(l), (self.hidden,), group="fp32", init="zeros")

    def flops_per_step(self, batch: int) -> float:
        """Executed GEMM FLOPs: fwd + wgrad for every layer, dgrad for all but the first."""
        per = 2.0 * batch * self.hidden * self.hidden
        return per * (2 * self.layers + (self.layers - 1))

    def _buffers(self, batch: int, device, key: int):
        k = (batch, key)
        if k not in self._acts:
            D = self.hidden
            self._acts[k] = dict(
                acts=[torch.empty(batch, D, device=device, dtype=torch.bfloat16) for _ in range(self.layers + 1)],
                dz=[torch.empty(batch, D, device=device, dtype=torch.bfloat16) for _ in range(2)],
                loss=torch.zeros(1, device=device), correct=torch.zeros(1, device=device, dtype=torch.int32))
        return self._acts[k]

    def forward_backward(self, store, x: torch.Tensor, labels: torch.Tensor, rank: Optional[int] = None,
                         lr: float = 0.0, local_sgd: bool = False, fused_pull: bool = False,
                         on_wgrad_done=None):
        """One worker's forward/backward.

        ``local_sgd`` (world == 1): the weight update is fused into the wgrad GEMM's epilogue
        (master -= lr*dW, bf16 copy refreshed) so the 1 GB gradient never touches HBM.
        ``fused_pull`` (world > 1): forward GEMMs gather their weights from the shard owners in
        the same launch.  ``on_wgrad_done(l)`` is called right after layer l's gradients (weights
        and bias) have been queued, so the caller can start that layer's push."""
        B, L = x.shape[0], self.layers
        bufs = self._buffers(B, x.device, 0 if rank is None else rank)
        acts, dzs, loss, correct = bufs["acts"], bufs["dz"], bufs["loss"], bufs["correct"]
        loss.zero_(); correct.zero_()
        a = x
        for l in range(L):
            out = acts[l + 1]
            ops.gemm(a, store.var(self.wname(l), rank), out, b_mn_major=True,
                     bias=store.var(self.bname(l), rank),
                     act=ops.ACT_SIGMOID if l < L - 1 else ops.ACT_NONE,
                     gather=store.gather_info(self.wname(l), rank) if fused_pull else None)
            a = out
        dz = dzs[0]
        ops.softmax_ce(acts[L], labels, dlogits=dz, loss_sum=loss, correct=correct,
                       grad_scale=1.0 / B, loss_scale=1.0 / B)
        ops.colsum(dz, store.grad(self.bname(L - 1), rank))       # bias gradient of the head
        fuse_cs = False   # measured: the fused column sums add 28 us to the dgrad GEMM, the stand-alone pass costs 24 us
        for l in range(L - 1, -1, -1):
            a_in = x if l == 0 else acts[l]
            w = store.var(self.wname(l), rank)
            dz_prev = None
            if l > 0:   # dgrad first: it must see the pre-update weights.  Its epilogue also emits the
                        # column sums of dz_{l-1} = the bias gradient of layer l-1 (no extra pass over dz)
                dz_prev = dzs[1] if dz is dzs[0] else dzs[0]
                ops.gemm(dz, w, dz_prev, dsigmoid_of=acts[l],
                         colsum_out=store.grad(self.bname(l - 1), rank) if fuse_cs else None)
                if not fuse_cs:
                    ops.colsum(dz_prev, store.grad(self.bname(l - 1), rank))
            if local_sgd:
                g = store.groups[self.wgroup(l)]
                r0 = store.pool.local_ranks[0]
                lo = g.layout.offsets[self.wname(l)] - g.layout.bounds[r0]
                master = g.master.view(r0, torch.float32)[lo: lo + w.numel()].view_as(w)
                ops.gemm(a_in, dz, a_mn_major=True, b_mn_major=True, epilogue=ops.EPI_SGD,
                         sgd_master=master, sgd_lowp=w, sgd_lr=lr)
            else:
                ops.gemm(a_in, dz, store.grad(self.wname(l), rank), a_mn_major=True, b_mn_major=True)
            if on_wgrad_done is not None:
                on_wgrad_done(l)
            dz = dz_prev
        return loss, correct
