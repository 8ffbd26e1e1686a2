"""Training engines: the public step API of the GPU path.

A *trainer* owns a ParameterStore, a model and a replica optimizer, and exposes

    trainer.step(x, labels)            # inputs already on the device (or None = reuse)
    trainer.step_from_host(x, labels)  # pinned host inputs -> H2D -> step -> loss back on the host

The step body (pull/prepare, forward, backward, push/apply) is launch-bound for the small
models, so after a warm-up it is captured once into a CUDA graph and replayed; all kernels
in it are native (libb200ps.so).  Cross-rank synchronisation lives inside the push kernel,
so a replay needs no host-side collective.

Mirrors the worker loop of example.py:158-184 / lr2.py:434-451 minus the per-step host
fetches: the loss stays on the device unless the caller asks for it.
"""
from __future__ import annotations

from typing import Optional

import torch

from .models.mlp import SigmoidMLP, WideMLP
from .models.sparse_lr import SparseLR
from .optim.sgd import AsyncSGD, SyncReplicasSGD
from .ps.store import ParameterStore
from .utils.trace import nvtx_range


def _capture(body, device):
    """Capture ``body`` into a CUDA graph.  The cyclic garbage collector is parked for the duration:
    a collection that happened to run mid-capture could destroy unrelated CUDA objects (older
    graphs, their memory pools), and freeing device memory is illegal while a stream is capturing
    — it invalidates the capture."""
    import gc
    gc.collect()
    torch.cuda.synchronize(device)
    was_enabled = gc.isenabled()
    gc.disable()
    try:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            body()
    finally:
        if was_enabled:
            gc.enable()
    return g


class _GraphedStep:
    """Warm up eagerly, then capture ``body`` into a CUDA graph and replay it."""

    def __init__(self, body, device, use_graph: bool, warmup: int = 2):
        self.body, self.device = body, device
        self.use_graph = use_graph and device.type == "cuda"
        self.warmup, self.calls, self.graph = warmup, 0, None

    def __call__(self):
        if not self.use_graph:
            self.body()
            return
        if self.graph is None:
            if self.calls < self.warmup:
                self.calls += 1
                self.body()
                return
            self.graph = _capture(self.body, self.device)
        self.graph.replay()


class _TrainerBase:
    def __init__(self, ctx, mode: str, lr: float, replicas_to_aggregate: Optional[int], placement: str,
                 seed: int):
        self.ctx = ctx
        self.device = ctx.device
        self.mode = mode
        self.lr = float(lr)
        self.store = ParameterStore(ctx.pool, placement=placement, seed=seed)
        if mode == "sync":
            self.opt = SyncReplicasSGD(lr, replicas_to_aggregate, ctx.world)
        elif mode == "async":
            self.opt = AsyncSGD(lr)
        else:
            raise ValueError("mode must be 'sync' or 'async'")
        self.local_ranks = list(ctx.pool.local_ranks)
        self.local_step = 0
        self._loss_host = torch.zeros(1).pin_memory() if self.device.type == "cuda" else torch.zeros(1)

    @property
    def global_step(self) -> int:
        return self.store.global_step

    def read_loss(self) -> float:
        """Device -> pinned host copy of the last step's loss (synchronises the stream)."""
        self._loss_host.copy_(self.last_loss, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        return float(self._loss_host.item())

    # -- pipelined end-to-end stepping: the host queues step t+1 (H2D copy, graph launch, D2H loss copy)
    #    before it blocks on the loss of step t, so PCIe, the GPU and the host loop overlap.  Every
    #    step's inputs still come from pinned host memory and every step's loss still reaches the host.
    def step_from_host_async(self, *host_inputs):
        if not hasattr(self, "_ring"):
            self._ring = [(torch.zeros(1).pin_memory(), torch.cuda.Event()) for _ in range(4)]
            self._ring_i = 0
        self.step(*host_inputs)
        slot, ev = self._ring[self._ring_i % len(self._ring)]
        self._ring_i += 1
        slot.copy_(self.last_loss, non_blocking=True)
        ev.record(torch.cuda.current_stream(self.device))
        return _LossHandle(slot, ev)

    def close(self):
        self.ctx.close()


class _LossHandle:
    """Result of ``step_from_host_async``: ``result()`` blocks until that step's loss is on the host."""

    def __init__(self, slot, event):
        self._slot, self._event = slot, event

    def result(self) -> float:
        self._event.synchronize()
        return float(self._slot.item())


class MLPTrainer(_TrainerBase):
    """MNIST sigmoid MLP (example.py) under sync or async PS updates."""

    def __init__(self, ctx, batch_size: int = 100, lr: float = 0.0005, mode: str = "sync",
                 replicas_to_aggregate: Optional[int] = None, placement: str = "flat_shard",
                 seed: int = 1, use_graph: bool = True, model: Optional[SigmoidMLP] = None,
                 fused_step: Optional[bool] = None):
        super().__init__(ctx, mode, lr, replicas_to_aggregate, placement, seed)
        self.model = model or SigmoidMLP()
        self.batch_size = batch_size
        # K13: forward+backward (and, for a single replica, the update) as one launch.  Default on
        # where the kernel applies; B200PS_FUSED_STEP=0 or fused_step=False selects the 4-kernel
        # tcgen05 chain (the path for larger batches / widths).
        import os
        if fused_step is None:
            fused_step = os.environ.get("B200PS_FUSED_STEP", "1") != "0"
        self.fused = bool(fused_step) and self.device.type == "cuda" and self.model.supports_fused(batch_size)
        self.fused_apply = self.fused and mode == "sync" and ctx.world == 1 and (replicas_to_aggregate in (None, 1))
        # K14: the single-replica step on one resident thread-block cluster (tensor cores, DSMEM bulk copies); many
        # steps per launch over a resident dataset or a host chunk.  Opt-in (B200PS_CLUSTER_STEP=1): measured at
        # 21 us/step against 13.4 us for K13 — 16 SMs x 8 warps cannot hide the latency that 98 CTAs hide
        # (profiles/EXPERIMENTS.md, profiles/ncu_cluster.txt).
        self.cluster = (self.fused_apply and os.environ.get("B200PS_CLUSTER_STEP", "0") == "1"
                        and self.model.supports_cluster(batch_size))
        self._want_fused_push = (self.fused and mode == "sync" and ctx.world > 1 and
                                 replicas_to_aggregate in (None, ctx.world) and
                                 os.environ.get("B200PS_FUSED_PUSH", "1") != "0")
        self.fused_push = False
        self.model.register(self.store)
        self.store.build()
        self.store.initialize()
        # synchronous replicas, one rank per GPU: the push runs inside the step kernel (one launch per step)
        self.fused_push = self._want_fused_push and self.model.supports_fused_push(self.store)
        # asynchronous workers: the whole Hogwild step (pull from the owners, compute, unlocked add) in one launch
        self.fused_async = (self.fused and mode == "async" and os.environ.get("B200PS_FUSED_ASYNC", "1") != "0"
                            and self.model.supports_fused_async(self.store))
        dev = self.device
        self.x = {r: torch.zeros(batch_size, self.model.in_dim, device=dev) for r in self.local_ranks}
        self.labels = {r: torch.zeros(batch_size, device=dev, dtype=torch.int32) for r in self.local_ranks}
        self._losses = {}
        self._corrects = {}
        self._step = _GraphedStep(self._body, dev, use_graph)
        if ctx.group is not None:
            ctx.control_barrier()

    def _one_step(self, r, x, labels, row_offset=None, stats_out=None, **cursor_kw):
        """One replica step on rank ``r``'s batch: returns (loss, correct) device/host words."""
        st = self.store
        if self.cluster:
            with nvtx_range("cluster_step"):
                return self.model.cluster_train_steps(st, x, labels, self.lr, steps=1, rank=r, row_offset=row_offset,
                                                      batch=self.batch_size, stats_out=stats_out,
                                                      cursor=cursor_kw.get("cursor"),
                                                      cursor_step=cursor_kw.get("cursor_step") or self.batch_size,
                                                      cursor_wrap=cursor_kw.get("cursor_wrap") or max(x.shape[0], 1))
        if self.fused_apply:
            ck = dict(cursor=cursor_kw.get("cursor"), cursor_step=cursor_kw.get("cursor_step", 0),
                      cursor_wrap=cursor_kw.get("cursor_wrap", 1))
            with nvtx_range("fused_step"):
                return self.model.fused_train_step(st, x, labels, self.lr, rank=r, row_offset=row_offset,
                                                   batch=self.batch_size, stats_out=stats_out, **ck)
        if self.fused_async:
            ck = dict(cursor=cursor_kw.get("cursor"), cursor_step=cursor_kw.get("cursor_step", 0),
                      cursor_wrap=cursor_kw.get("cursor_wrap", 1))
            with nvtx_range("fused_step+pull+push_async"):
                return self.model.fused_async_step(st, x, labels, self.lr, rank=r, row_offset=row_offset,
                                                   batch=self.batch_size, stats_out=stats_out, **ck)
        if self.fused_push:
            ck = dict(cursor=cursor_kw.get("cursor"), cursor_step=cursor_kw.get("cursor_step", 0),
                      cursor_wrap=cursor_kw.get("cursor_wrap", 1))
            with nvtx_range("fused_step+push_sync"):
                return self.model.fused_push_step(st, x, labels, self.lr, rank=r, row_offset=row_offset,
                                                  batch=self.batch_size, stats_out=stats_out, **ck)
        with nvtx_range("forward_backward"):
            out = self.model.forward_backward(st, x, labels, rank=r, row_offset=row_offset, batch=self.batch_size,
                                              stats_out=stats_out, fused=self.fused)
        with nvtx_range("push"):
            if self.mode == "async":
                self.opt.apply(st, rank=r, **cursor_kw)
            elif len(self.local_ranks) == 1:
                self.opt.apply(st, **cursor_kw)
        return out

    def _prepare(self):
        """Refresh the worker's parameter view before a step (async pull; nothing to do when the step kernel
        reads the owners' masters itself, or in sync mode where the push already broadcast)."""
        if not self.fused_async:
            with nvtx_range("pull"):
                self.opt.prepare(self.store)

    def _body(self):
        st = self.store
        self._prepare()
        for r in self.local_ranks:
            self._losses[r], self._corrects[r] = self._one_step(r, self.x[r], self.labels[r])
        if self.mode == "sync" and len(self.local_ranks) > 1:
            self.opt.apply(st)                     # virtual ranks: one push launch for all of them

    # -- resident dataset: the whole (per-worker) epoch lives on the device and the captured step
    #    walks it through a device-side cursor — no per-step input copies, no host work but the
    #    graph launch (the on-device replacement of mnist.train.next_batch, example.py:165)
    def attach_dataset(self, data: torch.Tensor, labels: torch.Tensor, steps_per_graph: int = 8) -> None:
        if len(self.local_ranks) != 1:
            raise NotImplementedError("resident datasets need one rank per process")
        n = (data.shape[0] // self.batch_size) * self.batch_size
        if n == 0 or labels.dtype != torch.int32 or labels.shape[0] < n:
            raise ValueError("attach_dataset: need >= batch_size rows and int32 labels")
        self._ds = (data, labels, n)
        self._cursor = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._step_res = _GraphedStep(self._body_resident, self.device, self._step.use_graph)
        # the cursor lives on the device, so consecutive steps chain inside ONE graph: a replay of
        # `steps_per_graph` steps costs one host launch
        self._spg = max(1, steps_per_graph)
        self._step_res_multi = _GraphedStep(self._body_resident_multi, self.device, self._step.use_graph)

    def _body_resident(self):
        data, labels, n = self._ds
        r = self.local_ranks[0]
        self._prepare()                          # async: refresh the peer-owned slices (no-op for sync)
        self._losses[r], self._corrects[r] = self._one_step(r, data, labels, row_offset=self._cursor,
                                                            cursor=self._cursor, cursor_step=self.batch_size,
                                                            cursor_wrap=n)

    def _body_resident_multi(self):
        if self.cluster:
            self._resident_launch(self._spg)
            return
        for _ in range(self._spg):
            self._body_resident()

    def _resident_launch(self, steps: int) -> None:
        """K14: ``steps`` steps over the attached dataset in one launch (the cluster stays resident)."""
        data, labels, n = self._ds
        r = self.local_ranks[0]
        self._losses[r], self._corrects[r] = self.model.cluster_train_steps(
            self.store, data, labels, self.lr, steps=steps, rank=r, row_offset=self._cursor, batch=self.batch_size,
            cursor=self._cursor, cursor_step=self.batch_size, cursor_wrap=n)

    def step_resident(self) -> None:
        """One step on the next batch of the attached dataset."""
        self._step_res()
        self.local_step += 1

    def run_resident(self, steps: int) -> None:
        """``steps`` consecutive steps over the attached dataset (multi-step graph replays + remainder)."""
        n = steps
        if self.cluster and steps > 0:
            self._resident_launch(steps)         # one launch, however many steps
            self.local_step += steps
            return
        while n >= self._spg and self._spg > 1:
            self._step_res_multi()
            n -= self._spg
        for _ in range(n):
            self._step_res()
        self.local_step += steps

    @property
    def last_loss(self) -> torch.Tensor:
        return self._losses[self.local_ranks[0]]

    @property
    def last_correct(self) -> torch.Tensor:
        return self._corrects[self.local_ranks[0]]

    def step(self, x: Optional[torch.Tensor] = None, labels: Optional[torch.Tensor] = None, rank=None):
        r = self.local_ranks[0] if rank is None else rank
        if x is not None:
            self.x[r].copy_(x, non_blocking=True)
        if labels is not None:
            self.labels[r].copy_(labels, non_blocking=True)
        self._step()
        self.local_step += 1

    def step_from_host(self, x_pinned: torch.Tensor, labels_pinned: torch.Tensor) -> float:
        """The end-to-end call: pinned host batch in, loss (python float) out."""
        return self.step_from_host_async(x_pinned, labels_pinned).result()

    def step_from_host_async(self, x_pinned: torch.Tensor, labels_pinned: torch.Tensor):
        """Queue one end-to-end step and return a handle whose ``result()`` is that step's loss.
        Once the trainer is warm the whole host side is ONE native call into the ``HostFeeder``
        (copy-stream H2D into a staging slot, event hand-off, graph launch; the loss is written by
        the step kernel into pinned host memory)."""
        f = self.__dict__.get("_feeder1")
        if f is None:
            if not self._step.use_graph or self._step.calls < self._step.warmup or len(self.local_ranks) != 1:
                return super().step_from_host_async(x_pinned, labels_pinned)
            f = self._feeder1 = self.host_feeder(steps_per_call=1, slots=8)
        self.local_step += 1
        return f.submit(x_pinned, labels_pinned)

    def host_feeder(self, steps_per_call: int = 1, slots: int = 4):
        """A ``HostFeeder`` whose unit is ``steps_per_call`` consecutive training steps on a pinned host
        chunk ``x[steps_per_call*B, in_dim]``, ``labels[steps_per_call*B]``.  ``submit(x, labels).result()``
        is the loss of the step (a float) or the ``steps_per_call`` losses (numpy array).  The trainer
        must have been warmed up (two eager ``step`` calls) so that nothing allocates during capture."""
        from .runtime import HostFeeder
        if len(self.local_ranks) != 1:
            raise NotImplementedError("host_feeder needs one rank per process")
        if self._step.use_graph and self._step.calls < self._step.warmup:
            raise RuntimeError("host_feeder: run %d eager step() calls first (warm-up)" % self._step.warmup)
        K, B, r, dev = max(1, steps_per_call), self.batch_size, self.local_ranks[0], self.device
        trainer = self

        def make_slot(slot, result):
            x = torch.zeros(K * B, self.model.in_dim, device=dev)
            y = torch.zeros(K * B, device=dev, dtype=torch.int32)
            words = result.view(K, 2)
            correct_words = words.view(torch.int32)

            def body():
                if trainer.cluster:              # K steps of the chunk in ONE launch, one (loss, correct) pair per step
                    trainer.model.cluster_train_steps(trainer.store, x, y, trainer.lr, steps=K, rank=r, batch=B,
                                                      stats_out=(words[0, 0:1], correct_words[0, 1:2]),
                                                      stat_stride=2, cursor_step=B, cursor_wrap=K * B)
                    return
                for i in range(K):
                    trainer._prepare()
                    trainer._one_step(r, x[i * B:(i + 1) * B], y[i * B:(i + 1) * B],
                                      stats_out=(words[i, 0:1], correct_words[i, 1:2]))
            return _capture(body, dev), (x, y)

        if K == 1:
            def decode(out):
                return float(out[0])
            decode.first_word = True          # lets the feeder return the loss straight from C
        else:
            decode = lambda out: out.reshape(K, 2)[:, 0].copy()                 # noqa: E731
        f = HostFeeder(make_slot, (K * B * self.model.in_dim * 4, K * B * 4), 2 * K, slots=slots, decode=decode)
        f.steps_per_call = K
        return f

    def fit_from_host(self, x_pinned: torch.Tensor, labels_pinned: torch.Tensor, steps_per_call: int = 8,
                      on_loss=None) -> int:
        """Train once over a pinned host dataset ``x[N, in_dim]``, ``labels[N]`` (the ``for i in
        range(batch_count): next_batch(); sess.run(...)`` epoch of example.py:163-170) through a
        chunked feeder: ``steps_per_call`` batches per H2D copy / graph launch, every step's loss
        delivered to ``on_loss(step_index, loss)``.  Returns the number of steps taken."""
        K, B = max(1, steps_per_call), self.batch_size
        key = "_feederK%d" % K
        if getattr(self, key, None) is None:
            setattr(self, key, self.host_feeder(steps_per_call=K))
        f = getattr(self, key)
        n_chunks = x_pinned.shape[0] // (K * B)
        pending, done = [], 0
        for c in range(n_chunks):
            pending.append(f.submit(x_pinned[c * K * B:(c + 1) * K * B], labels_pinned[c * K * B:(c + 1) * K * B]))
            if len(pending) > f.slots - 2:
                losses = pending.pop(0).result()
                if on_loss is not None:
                    for v in (losses if K > 1 else [losses]):
                        on_loss(done, float(v)); done += 1
                else:
                    done += K
        for t in pending:
            losses = t.result()
            if on_loss is not None:
                for v in (losses if K > 1 else [losses]):
                    on_loss(done, float(v)); done += 1
            else:
                done += K
        self.local_step += n_chunks * K
        # the remainder (fewer than K batches) goes through the per-step feeder
        rest = x_pinned.shape[0] // B - n_chunks * K
        for i in range(rest):
            lo = (n_chunks * K + i) * B
            v = self.step_from_host(x_pinned[lo:lo + B], labels_pinned[lo:lo + B])
            if on_loss is not None:
                on_loss(done, v)
            done += 1
        return done

    def evaluate(self, x: torch.Tensor, labels: torch.Tensor):
        if self.fused_async:
            self.store.pull()                    # the in-kernel Hogwild step keeps no local copy up to date
        return self.model.evaluate(self.store, x, labels, rank=self.local_ranks[0])


class WideMLPTrainer(_TrainerBase):
    """BASELINE.json config 4: 8 x 8192 x 8192 bf16 sigmoid MLP, synchronous PS."""

    def __init__(self, ctx, batch_size: int = 4096, layers: int = 8, hidden: int = 8192, lr: float = 0.01,
                 replicas_to_aggregate: Optional[int] = None, seed: int = 1, use_graph: bool = True,
                 fused_pull: bool = False, overlap: bool = True):
        super().__init__(ctx, "sync", lr, replicas_to_aggregate, "flat_shard", seed)
        self.model = WideMLP(layers, hidden)
        self.batch_size = batch_size
        self.model.register(self.store)
        self.store.build()
        self.store.initialize()
        dev = self.device
        self.x = {r: torch.zeros(batch_size, hidden, device=dev, dtype=torch.bfloat16) for r in self.local_ranks}
        self.labels = {r: torch.zeros(batch_size, device=dev, dtype=torch.int32) for r in self.local_ranks}
        self._losses, self._corrects = {}, {}
        self.local_sgd = ctx.world == 1
        # K2 (pull fused into the forward GEMMs) needs the row shards to be whole 64-row panels.  It is
        # off by default: on this compute-bound model the broadcast folded into the overlapped push is
        # cheaper (measured 12.1 vs 13.3 ms/step on 8 GPUs)
        self.fused_pull = fused_pull and ctx.world > 1 and hidden % (64 * ctx.world) == 0
        self.overlap = overlap
        # per-layer pushes that run on the side stream UNDER the backward GEMMs stay on the P2P kernel: measured on
        # 8 GPUs, 11.6 ms/step against 12.0-13.3 ms through the switch (tools/wide_nvls_sweep.py); B200PS_WIDE_NVLS=1
        # opts in.  Pushes that are not overlapped take the store's default (NVLS for 1 MB .. 512 MB groups).
        import os
        self._nvls_overlapped = True if os.environ.get("B200PS_WIDE_NVLS", "0") == "1" else False
        self._side = torch.cuda.Stream(device=dev) if dev.type == "cuda" else None
        self._step = _GraphedStep(self._body, dev, use_graph)
        if ctx.group is not None:
            ctx.control_barrier()

    def _body(self):
        st, m = self.store, self.model
        if self.local_sgd:
            for r in self.local_ranks:
                loss, correct = m.forward_backward(st, self.x[r], self.labels[r], rank=r, lr=self.lr, local_sgd=True)
                self._losses[r], self._corrects[r] = loss, correct
            st.push_sync(self.lr, groups=("fp32",))            # biases; weights were fused into wgrad
            return
        R = self.opt.replicas_to_aggregate
        if len(self.local_ranks) > 1 or not self.overlap:
            # virtual ranks (tests) / overlap off: same kernels, pushes after the whole backward
            for r in self.local_ranks:
                loss, correct = m.forward_backward(st, self.x[r], self.labels[r], rank=r, fused_pull=self.fused_pull)
                self._losses[r], self._corrects[r] = loss, correct
            for l in range(m.layers - 1, -1, -1):
                st.push_sync(self.lr, R, groups=(m.wgroup(l),), count_step=False, broadcast=not self.fused_pull)
            st.push_sync(self.lr, R, groups=("fp32",))
            return
        # one rank per process: layer l's push runs on the side stream under the rest of the backward
        main = torch.cuda.current_stream(self.device)
        side = self._side
        r = self.local_ranks[0]

        def on_wgrad_done(l):
            side.wait_stream(main)
            with torch.cuda.stream(side):
                st.push_sync(self.lr, R, groups=(m.wgroup(l),), count_step=False, broadcast=not self.fused_pull,
                             nvls=self._nvls_overlapped)

        loss, correct = m.forward_backward(st, self.x[r], self.labels[r], rank=r, fused_pull=self.fused_pull,
                                           on_wgrad_done=on_wgrad_done)
        self._losses[r], self._corrects[r] = loss, correct
        side.wait_stream(main)
        with torch.cuda.stream(side):                           # all pushes share the signal pads:
            st.push_sync(self.lr, R, groups=("fp32",))          # keep them on ONE stream, in order
        main.wait_stream(side)

    @property
    def last_loss(self) -> torch.Tensor:
        return self._losses[self.local_ranks[0]]

    def step(self, x=None, labels=None, rank=None):
        r = self.local_ranks[0] if rank is None else rank
        if x is not None:
            self.x[r].copy_(x, non_blocking=True)
        if labels is not None:
            self.labels[r].copy_(labels, non_blocking=True)
        self._step()
        self.local_step += 1

    def step_from_host(self, x_pinned, labels_pinned) -> float:
        self.step(x_pinned, labels_pinned)
        return self.read_loss()


class SparseLRTrainer(_TrainerBase):
    """lr2.py's trainer over a row-sharded weight: async (default, like the reference) or sync."""

    def __init__(self, ctx, features: int, lr: float = 0.001, mode: str = "async",
                 placement: str = "flat_shard", seed: int = 1, use_graph: bool = True, model=None,
                 dedupe: Optional[bool] = None):
        super().__init__(ctx, mode, lr, None, placement, seed)
        self.model = model or SparseLR(features, dedupe=dedupe)    # dedupe: unique-id step (sparse_dedup.cu)
        self.model.register(self.store)
        self.store.build()
        self.store.initialize()
        self._last = None
        self.use_graph = use_graph and self.device.type == "cuda"
        self._graphs = {}
        if ctx.group is not None:
            ctx.control_barrier()

    @property
    def last_loss(self) -> torch.Tensor:
        return self._last

    def _eager(self, batch, rank=None):
        if self.mode == "async":
            self._last = self.model.step_async(self.store, batch, self.lr, rank=rank)
        else:
            self._last = self.model.step_sync(self.store, batch, self.lr, rank=rank)

    def step(self, batch, rank=None):
        """``batch`` = (row_ptr int32[B+1], fid int64[nnz], fval f32[nnz], labels f32[B]) on the device.

        The step is one or two kernels, i.e. host-launch-bound; a batch whose buffers are reused
        (a prefetch ring) gets its own CUDA graph on its second visit."""
        key = tuple(t.data_ptr() for t in batch) + (batch[3].numel(),)
        ent = self._graphs.get(key) if self.use_graph else None
        if ent is None:
            self._eager(batch, rank)
            if self.use_graph:
                self._graphs[key] = 1 if key not in self._graphs else self._graphs[key]
        elif ent == 1:
            g = _capture(lambda: self._eager(batch, rank), self.device)
            self._graphs[key] = (g, batch)          # keep the batch tensors alive with the graph
            g.replay()
        else:
            ent[0].replay()
        self.local_step += 1

    def step_from_host(self, batch_pinned, dev_bufs) -> float:
        for src, dst in zip(batch_pinned, dev_bufs):
            dst[: src.numel()].copy_(src, non_blocking=True)
        n = batch_pinned[3].numel()
        self.step((dev_bufs[0][: n + 1], dev_bufs[1], dev_bufs[2], dev_bufs[3][:n]))
        return self.read_loss()

    def predict(self, batch):
        return self.model.predict(self.store, batch)

    def read_loss(self) -> float:
        """Loss of the last step; raises ``ValueError`` if any step since the last call saw a feature id
        outside ``[0, features)`` (the kernels skip and count them)."""
        v = super().read_loss()
        from . import ops
        ops.check_sparse_ids()
        return v

    # -- packed batches (b200ps.data.native_loader): one pinned buffer per batch, one H2D copy, one graph
    def host_feeder(self, layout, slots: int = 4):
        """A ``HostFeeder`` whose unit is one training step on a *packed batch* (``PackedLayout``): the
        buffer is copied to a staging slot on the copy stream, the slot's captured step graph reads
        ``row_ptr | fid | fval | labels`` at the layout's fixed offsets, and the loss lands in pinned host
        memory.  ``submit(packed_uint8_tensor).result()`` is the step's loss."""
        from . import ops
        from .runtime import HostFeeder
        dev = self.device
        trainer = self

        def make_slot(slot, result):
            staging = torch.zeros(layout.bytes, dtype=torch.uint8, device=dev)
            batch = layout.views(staging)
            if slot == 0:
                trainer.model.predict(trainer.store, batch)      # loads the kernel before any capture (no update)
                torch.cuda.synchronize(dev)

            def body():
                trainer._eager(batch)
                ops.copy_words(result, trainer._last, 1)
            return _capture(body, dev), (staging,)

        def decode(out):
            return float(out[0])
        decode.first_word = True
        f = HostFeeder(make_slot, (layout.bytes,), 1, slots=slots, decode=decode)
        f.layout = layout
        return f

    def fit_from_loader(self, loader, max_steps: Optional[int] = None, on_loss=None, feeder=None) -> int:
        """Drain a ``NativeLibsvmLoader`` through a feeder: the loader's reader threads parse and batch, each
        packed batch costs one H2D copy and one graph launch, losses come back through ``on_loss(step, loss)``.
        Returns the number of steps.  (The training loop of lr2.py:434-451 without a Python batch formatter.)"""
        f = feeder or getattr(self, "_loader_feeder", None)
        if f is None or f.layout != loader.layout:
            f = self._loader_feeder = self.host_feeder(loader.layout, slots=4)
        pending, steps = [], 0
        # batches in flight are ring buffers the reader threads cannot reuse: leave one per reader plus one
        depth = max(1, min(f.slots - 1, loader.ring - loader.threads - 1))

        def drain(n_keep):
            nonlocal steps
            while len(pending) > n_keep:
                ticket, batch = pending.pop(0)
                loss = ticket.result()
                batch.release()                      # the H2D copy of this buffer has completed
                if on_loss is not None:
                    on_loss(steps, loss)
                steps += 1

        submitted = 0
        while max_steps is None or submitted < max_steps:
            batch = loader.next()
            if batch is None:
                break
            pending.append((f.submit(batch.tensor), batch))
            submitted += 1
            drain(depth)
        drain(0)
        self.local_step += steps
        return steps


class SparseWideDeepTrainer(SparseLRTrainer):
    """The wide & deep extension of the LR trainer (lr2_debug.py:423-428): same step API and modes,
    model = wide [F,1] + embedding [F,dim] x [dim,1]."""

    def __init__(self, ctx, features: int, dim: int = 128, wide: bool = True, lr: float = 0.001,
                 mode: str = "async", placement: str = "flat_shard", seed: int = 1, use_graph: bool = True,
                 deferred: Optional[bool] = None, owner: Optional[bool] = None):
        from .models.wide_deep import SparseWideDeep
        super().__init__(ctx, features, lr=lr, mode=mode, placement=placement, seed=seed, use_graph=use_graph,
                         model=SparseWideDeep(features, dim=dim, wide=wide, deferred=deferred, owner=owner))
        self.owner_step = self.model.uses_owner_step(self.store, sync=(mode == "sync"))

    def flush(self) -> None:
        """Fold pending (deferred) updates of this process's shards into the tables — before checkpoints and
        evaluations that must see every update sent so far."""
        self.model.flush(self.store)
