"""The measured baseline: a PyTorch + NCCL (+ cuBLAS) implementation of the same
parameter-server semantics, using only stock library calls — the reference itself (TensorFlow
0.8-0.12, Python 2) cannot be installed here.

    sharded PS            : each rank owns 1/N of the flat parameter vector (fp32 master)
    pull                  : dist.all_gather_into_tensor of the owner slices
    sync push + apply     : dist.reduce_scatter_tensor(mean) -> slice -= lr * g  (then the pull)
    async push (emulated) : per-worker dist.reduce_scatter_tensor(sum) applied with the full lr
                            (every worker's update lands; NCCL has no barrier-free peer atomics)
    MLP math              : torch.addmm / sigmoid / cross_entropy + autograd (cuBLAS GEMMs)
    sparse LR             : row-sharded W, dist.all_to_all_single of ids and rows, index_add_

It is driven by ``bench.py --impl baseline`` with exactly the protocol of the b200ps arm
(device-resident synthetic epoch streamed per step, CUDA-event timing, max over ranks; e2e
from pinned host memory with the loss read back each step).  Nothing from b200ps's kernels
or engine is used here except the synthetic data generators that produce the input tensors.
"""
from __future__ import annotations

import datetime
import json
import os
import sys
import time

import torch
import torch.distributed as dist
import torch.nn.functional as Fnn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MNIST_TRAIN = 55000


class ShardedPS:
    """Flat fp32 parameters, 1/N master slice per rank, NCCL collectives for pull/push."""

    def __init__(self, shapes, rank, world, device, seed=1, lowp=None, group=None):
        self.rank, self.world, self.device = rank, world, device
        self.group = group
        self.shapes = shapes
        sizes = [int(torch.Size(s).numel()) for s in shapes.values()]
        total = sum(sizes)
        self.shard = (total + world - 1) // world
        self.total = self.shard * world
        g = torch.Generator(device="cpu").manual_seed(seed)
        full = torch.zeros(self.total)
        off = 0
        self.views = {}
        for (name, shp), n in zip(shapes.items(), sizes):
            if len(shp) == 2:
                full[off:off + n] = torch.randn(n, generator=g) * (1.0 if lowp is None else 4.0 / shp[0] ** 0.5)
            self.views[name] = (off, n, shp)
            off += n
        self.master = full[rank * self.shard:(rank + 1) * self.shard].to(device)       # owner slice (fp32)
        self.lowp = lowp
        self.work = torch.zeros(self.total, device=device, dtype=lowp or torch.float32)  # pulled replica
        self.grads = torch.zeros(self.total, device=device, dtype=lowp or torch.float32)
        self._gslice = torch.zeros(self.shard, device=device, dtype=lowp or torch.float32)
        self.global_step = torch.zeros(1, device=device, dtype=torch.int64)
        self.pull()

    def var(self, name):
        off, n, shp = self.views[name]
        return self.work[off:off + n].view(shp)

    def grad(self, name):
        off, n, shp = self.views[name]
        return self.grads[off:off + n].view(shp)

    def pull(self):
        src = self.master if self.lowp is None else self.master.to(self.lowp)
        if self.world == 1:
            self.work[:self.shard].copy_(src)
        else:
            dist.all_gather_into_tensor(self.work, src, group=self.group)

    def push_sync(self, lr):
        if self.world == 1:
            self.master.add_(self.grads[:self.shard].float(), alpha=-lr)
        else:
            dist.reduce_scatter_tensor(self._gslice, self.grads, op=dist.ReduceOp.AVG, group=self.group)
            self.master.add_(self._gslice.float(), alpha=-lr)
        self.global_step += 1
        self.pull()

    def push_async(self, lr):
        if self.world == 1:
            self.master.add_(self.grads[:self.shard].float(), alpha=-lr)
        else:
            dist.reduce_scatter_tensor(self._gslice, self.grads, op=dist.ReduceOp.SUM, group=self.group)
            self.master.add_(self._gslice.float(), alpha=-lr)
        self.global_step += self.world
        self.pull()


def mlp_step(ps, x, y, lr, mode):
    w1, w2, b1, b2 = (ps.var(n).detach().requires_grad_(True) for n in ("W1", "W2", "b1", "b2"))
    a2 = torch.sigmoid(torch.addmm(b1, x, w1))
    loss = Fnn.cross_entropy(torch.addmm(b2, a2, w2), y)
    g = torch.autograd.grad(loss, (w1, w2, b1, b2))
    for n, gi in zip(("W1", "W2", "b1", "b2"), g):
        ps.grad(n).copy_(gi)
    (ps.push_sync if mode == "sync" else ps.push_async)(lr)
    return loss.detach()


def wide_step(ps, x, y, lr, layers):
    ws = [ps.var(f"W{l}").detach().requires_grad_(True) for l in range(layers)]
    bs = [ps.var(f"b{l}").detach().float().requires_grad_(True) for l in range(layers)]
    a = x
    for l in range(layers):
        z = torch.addmm(bs[l].to(a.dtype), a, ws[l])
        a = torch.sigmoid(z) if l < layers - 1 else z
    loss = Fnn.cross_entropy(a.float(), y)
    g = torch.autograd.grad(loss, ws + bs)
    for l in range(layers):
        ps.grad(f"W{l}").copy_(g[l])
        ps.grad(f"b{l}").copy_(g[layers + l])
    ps.push_sync(lr)
    return loss.detach()


class SparsePS:
    def __init__(self, F, rank, world, device):
        self.rows = (F + world - 1) // world
        self.rank, self.world, self.device = rank, world, device
        self.W = torch.randn(self.rows, device=device)
        self.b = torch.zeros(1, device=device)
        self.global_step = 0

    def step(self, batch, lr):
        rp, fid, fv, y = batch
        B = y.numel()
        if self.world == 1:
            return self._step_local(rp, fid, fv, y, lr)
        n = int(rp[-1])          # the variable-split all_to_all API needs the counts on the host
        fid, fv = fid[:n], fv[:n]
        row_of = torch.repeat_interleave(torch.arange(B, device=self.device), (rp[1:] - rp[:-1]).long())
        if self.world == 1:
            w = self.W[fid]
        else:
            owner = fid // self.rows
            order = torch.argsort(owner)
            counts = torch.bincount(owner, minlength=self.world)
            recv_counts = torch.empty_like(counts)
            dist.all_to_all_single(recv_counts, counts)
            sc, rc = counts.tolist(), recv_counts.tolist()
            ids_out = (fid - owner * self.rows)[order]
            ids_in = torch.empty(sum(rc), dtype=torch.int64, device=self.device)
            dist.all_to_all_single(ids_in, ids_out, rc, sc)
            rows_out = self.W[ids_in]
            rows_in = torch.empty(n, device=self.device)
            dist.all_to_all_single(rows_in, rows_out, sc, rc)
            w = torch.empty(n, device=self.device)
            w[order] = rows_in
        z = torch.zeros(B, device=self.device).index_add_(0, row_of, w * fv) + self.b
        loss = Fnn.binary_cross_entropy_with_logits(z, y)
        dz = (torch.sigmoid(z) - y) / B
        gv = dz[row_of] * fv
        if self.world == 1:
            self.W.index_add_(0, fid, gv, alpha=-lr)
            self.b -= lr * dz.sum()
        else:
            g_in = torch.empty(sum(rc), device=self.device)
            dist.all_to_all_single(g_in, gv[order], rc, sc)
            self.W.index_add_(0, ids_in, g_in, alpha=-lr)
            db = dz.sum()
            dist.all_reduce(db)
            self.b -= lr * db
        self.global_step += self.world
        return loss.detach()


    def _step_local(self, rp, fid, fv, y, lr):
        """One GPU: no host read anywhere (capturable).  The CSR buffers are padded to a fixed capacity;
        slots past rp[-1] are masked on the device."""
        B, cap = y.numel(), fid.numel()
        pos = torch.arange(cap, device=self.device, dtype=torch.int32)
        live = pos < rp[-1]
        row_of = torch.searchsorted(rp[1:].contiguous(), pos, right=True).clamp_(max=B - 1).long()
        fidc = torch.where(live, fid, torch.zeros_like(fid))
        fvm = torch.where(live, fv, torch.zeros_like(fv))
        z = torch.zeros(B, device=self.device).index_add_(0, row_of, self.W[fidc] * fvm) + self.b
        loss = Fnn.binary_cross_entropy_with_logits(z, y)
        dz = (torch.sigmoid(z) - y) / B
        self.W.index_add_(0, fidc, dz[row_of] * fvm, alpha=-lr)
        self.b -= lr * dz.sum()
        self.global_step += 1
        return loss.detach()


# ---------------------------------------------------------------------------------------------
class WideDeepPS:
    """The wide & deep extension with stock ops: row-sharded W [F,1] and E [F,D], replicated w2 [D,1];
    ids / rows / gradient rows move with all_to_all_single, w2 and b gradients with all_reduce."""

    def __init__(self, F, D, rank, world, device):
        self.rows = (F + world - 1) // world
        self.rank, self.world, self.device, self.D = rank, world, device, D
        g = torch.Generator(device=device).manual_seed(1)
        self.W = torch.randn(self.rows, device=device, generator=g) * 0.01
        self.E = torch.randn(self.rows, D, device=device, generator=g) * 0.01
        self.w2 = torch.randn(D, device=device, generator=g) * 0.01
        self.b = torch.zeros(1, device=device)
        self.global_step = 0

    def step(self, batch, lr):
        rp, fid, fv, y = batch
        n = int(rp[-1])
        fid, fv = fid[:n], fv[:n]
        B, D = y.numel(), self.D
        row_of = torch.repeat_interleave(torch.arange(B, device=self.device), (rp[1:] - rp[:-1]).long())
        if self.world == 1:
            w, e = self.W[fid], self.E[fid]
        else:
            owner = fid // self.rows
            order = torch.argsort(owner)
            counts = torch.bincount(owner, minlength=self.world)
            recv_counts = torch.empty_like(counts)
            dist.all_to_all_single(recv_counts, counts)
            sc, rc = counts.tolist(), recv_counts.tolist()
            ids_in = torch.empty(sum(rc), dtype=torch.int64, device=self.device)
            dist.all_to_all_single(ids_in, (fid - owner * self.rows)[order], rc, sc)
            rows_in = torch.empty(n, D + 1, device=self.device)
            dist.all_to_all_single(rows_in, torch.cat([self.E[ids_in], self.W[ids_in, None]], 1), sc, rc)
            both = torch.empty(n, D + 1, device=self.device)
            both[order] = rows_in
            e, w = both[:, :D], both[:, D]
        h = torch.zeros(B, D, device=self.device).index_add_(0, row_of, e * fv[:, None])
        z = h @ self.w2 + torch.zeros(B, device=self.device).index_add_(0, row_of, w * fv) + self.b
        loss = Fnn.binary_cross_entropy_with_logits(z, y)
        dz = (torch.sigmoid(z) - y) / B
        gv = dz[row_of] * fv
        ge = gv[:, None] * self.w2[None, :]
        gw2, db = h.t() @ dz, dz.sum()
        if self.world == 1:
            self.W.index_add_(0, fid, gv, alpha=-lr)
            self.E.index_add_(0, fid, ge, alpha=-lr)
        else:
            g_in = torch.empty(sum(rc), D + 1, device=self.device)
            dist.all_to_all_single(g_in, torch.cat([ge, gv[:, None]], 1)[order], rc, sc)
            self.E.index_add_(0, ids_in, g_in[:, :D], alpha=-lr)
            self.W.index_add_(0, ids_in, g_in[:, D], alpha=-lr)
            small = torch.cat([gw2, db[None]])
            dist.all_reduce(small)
            gw2, db = small[:D], small[D]
        self.w2 -= lr * gw2
        self.b -= lr * db
        self.global_step += self.world
        return loss.detach()


class GraphedMLPStep:
    """The same `mlp_step` (autograd + NCCL collectives) captured once into a CUDA graph on static input
    buffers and replayed: the launch-overhead-free form of the baseline."""

    def __init__(self, ps, B, lr, mode, device):
        self.ps, self.lr, self.mode = ps, lr, mode
        self.x = torch.zeros(B, 784, device=device)
        self.y = torch.zeros(B, device=device, dtype=torch.int64)
        self.loss = torch.zeros((), device=device)
        side = torch.cuda.Stream(device=device)
        side.wait_stream(torch.cuda.current_stream(device))
        with torch.cuda.stream(side):
            for _ in range(3):
                self.loss.copy_(mlp_step(ps, self.x, self.y, lr, mode))
        torch.cuda.current_stream(device).wait_stream(side)
        torch.cuda.synchronize(device)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.loss.copy_(mlp_step(ps, self.x, self.y, lr, mode))

    def __call__(self, x, y):
        self.x.copy_(x, non_blocking=True)
        self.y.copy_(y, non_blocking=True)
        self.graph.replay()
        return self.loss


def run_mnist(args, rank, world, dev, mode, group=None, graphed=True):
    """MNIST 784-100-10, batch 100 per worker, by bench.py's protocol.  Returns eager numbers and, under
    "graphed", the CUDA-graph replay of the same step."""
    sys.path.insert(0, ROOT)
    from b200ps import ops                      # input tensors only
    W, B = max(args.warmup, 3), 100
    ps = ShardedPS({"W1": (784, 100), "W2": (100, 10), "b1": (100,), "b2": (10,)}, rank, world, dev, group=group)
    data = torch.empty(MNIST_TRAIN, 784, device=dev); lab = torch.empty(MNIST_TRAIN, device=dev, dtype=torch.int32)
    ops.synth_mnist(data, lab, seed=1, offset=rank * MNIST_TRAIN)
    xs, ys = data.view(-1, B, 784), lab.long().view(-1, B)
    nb = xs.shape[0]
    hx, hy = xs[:64].cpu().pin_memory(), ys[:64].cpu().pin_memory()
    sx, sy = torch.empty(B, 784, device=dev), torch.empty(B, device=dev, dtype=torch.int64)

    def measure(step_dev, step_host):
        for i in range(W):
            step_dev(i)
        ms = _timed(args.steps, step_dev, world, group)
        for i in range(W):
            step_host(i)
        ms2 = _wall(args.steps, step_host, world, group)
        return {"value": B * world * args.steps / (ms / 1e3), "ms_per_step": ms / args.steps,
                "e2e": {"value": B * world * args.steps / (ms2 / 1e3), "unit": "examples/s",
                        "h2d_bytes_per_step": B * 784 * 4 + B * 8, "d2h_bytes_per_step": 4}}

    def e2e_eager(i):
        sx.copy_(hx[i % 64], non_blocking=True); sy.copy_(hy[i % 64], non_blocking=True)
        return float(mlp_step(ps, sx, sy, 0.0005, mode).item())

    out = measure(lambda i: mlp_step(ps, xs[i % nb], ys[i % nb], 0.0005, mode), e2e_eager)
    if graphed:
        try:
            g = GraphedMLPStep(ps, B, 0.0005, mode, dev)
            out["graphed"] = measure(lambda i: g(xs[i % nb], ys[i % nb]),
                                     lambda i: float(g(hx[i % 64], hy[i % 64]).item()))
        except Exception as e:                 # capture of the collectives can fail on some NCCL builds
            out["graphed"] = {"error": f"{type(e).__name__}: {e}"}
    return out


def _timed(steps, fn, world, group=None):
    if world > 1:
        dist.barrier(group=group)
    torch.cuda.synchronize()
    st, en = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    st.record()
    for i in range(steps):
        fn(i)
    en.record()
    torch.cuda.synchronize()
    ms = torch.tensor([st.elapsed_time(en)], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX, group=group)
    return float(ms.item())


def _wall(steps, fn, world, group=None):
    if world > 1:
        dist.barrier(group=group)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(steps):
        fn(i)
    torch.cuda.synchronize()
    ms = torch.tensor([(time.perf_counter() - t0) * 1e3], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX, group=group)
    return float(ms.item())


def main_from_bench(args) -> int:
    sys.path.insert(0, ROOT)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", timeout=datetime.timedelta(seconds=300), device_id=dev)
    from b200ps import ops                      # input tensors only
    out = {}
    W = max(args.warmup, 3)
    cfgs = (["mnist_mlp_sync", "mnist_mlp_async", "sparse_lr", "sparse_wide_deep", "wide_mlp"] if args.config == "all"
            else [args.config])
    for cfg in cfgs:
        if cfg.startswith("mnist"):
            out[cfg] = run_mnist(args, rank, world, dev, "sync" if cfg.endswith("_sync") else "async")
        elif cfg == "wide_mlp":
            B, L, D = args.wide_batch, args.wide_layers, args.wide_hidden
            shapes = {**{f"W{l}": (D, D) for l in range(L)}, **{f"b{l}": (D,) for l in range(L)}}
            ps = ShardedPS(shapes, rank, world, dev, lowp=torch.bfloat16)
            xs = torch.empty(4, B, D, device=dev, dtype=torch.bfloat16); ys = torch.empty(4, B, device=dev, dtype=torch.int32)
            for i in range(4):
                ops.synth_dense_bf16(xs[i], ys[i], classes=D, seed=3, offset=(rank * 4 + i) * B)
            yl = ys.long()
            steps = max(1, min(args.steps, args.wide_max_steps))
            step = lambda i: wide_step(ps, xs[i % 4], yl[i % 4], 0.01, L)
            for i in range(3):
                step(i)
            ms = _timed(steps, step, world)
            out[cfg] = {"value": B * world * steps / (ms / 1e3), "ms_per_step": ms / steps, "steps": steps}
        elif cfg in ("sparse_lr", "sparse_wide_deep"):
            deep = cfg == "sparse_wide_deep"
            F, B, nnz, spread = (1 << 22 if deep else 1 << 24), args.sparse_batch, 100, 20
            sp = WideDeepPS(F, 128, rank, world, dev) if deep else SparsePS(F, rank, world, dev)
            batches = []
            for i in range(64):
                rp = torch.empty(B + 1, device=dev, dtype=torch.int32); fid = torch.empty(B * (nnz + spread), device=dev, dtype=torch.int64)
                fv = torch.empty(B * (nnz + spread), device=dev); y = torch.empty(B, device=dev)
                ops.synth_sparse(rp, fid, fv, y, nnz_mean=nnz, spread=spread, features=F, seed=5, offset=(rank * 64 + i) * B)
                batches.append((rp, fid, fv, y))
            step = lambda i: sp.step(batches[i % 64], 0.05)
            for i in range(W):
                step(i)
            ms = _timed(args.steps, step, world)
            out[cfg] = {"value": B * world * args.steps / (ms / 1e3), "ms_per_step": ms / args.steps}
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
    if rank == 0:
        primary = "mnist_mlp_sync" if args.config == "all" else args.config
        line = {"impl": "baseline (PyTorch+NCCL+cuBLAS, same PS semantics)", "metric": "examples/sec (device-timed, max over ranks)",
                "value": out[primary]["value"], "unit": "examples/s", "n_gpus": world, "steps": args.steps, "warmup": W,
                "ms_per_step": out[primary]["ms_per_step"], "higher_is_better": True, "scaling": "weak",
                "data": "synthetic", "config": {"model": primary}, "e2e": out[primary].get("e2e"),
                "graphed": out[primary].get("graphed"),
                "extra": {k: v for k, v in out.items() if k != primary}}
        print(json.dumps(line))
        sys.stdout.flush()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0
