"""Tiny eager driver for ncu: a few un-graphed steps of one config so every kernel is visible."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ps import ops  # noqa: E402
from b200ps.parallel.context import init_from_env  # noqa: E402
from b200ps.train import MLPTrainer, SparseLRTrainer, SparseWideDeepTrainer, WideMLPTrainer  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--config", default="mnist")
ap.add_argument("--steps", type=int, default=4)
a = ap.parse_args()
ctx = init_from_env()
dev = ctx.device
if a.config == "mnist":
    tr = MLPTrainer(ctx, batch_size=100, lr=0.0005, mode="sync", use_graph=False)
    x = torch.empty(100, 784, device=dev); y = torch.empty(100, device=dev, dtype=torch.int32)
    ops.synth_mnist(x, y, seed=1)
    for _ in range(a.steps):
        tr.step(x, y)
elif a.config == "wide":
    tr = WideMLPTrainer(ctx, batch_size=4096, layers=8, hidden=8192, lr=0.01, use_graph=False)
    x = torch.empty(4096, 8192, device=dev, dtype=torch.bfloat16); y = torch.empty(4096, device=dev, dtype=torch.int32)
    ops.synth_dense_bf16(x, y, classes=8192, seed=3)
    for _ in range(a.steps):
        tr.step(x, y)
elif a.config == "sparse":
    F, B, nnz, spread = 1 << 24, 500, 100, 20
    tr = SparseLRTrainer(ctx, features=F, lr=0.05, mode="async")
    rp = torch.empty(B + 1, device=dev, dtype=torch.int32); fid = torch.empty(B * (nnz + spread), device=dev, dtype=torch.int64)
    fv = torch.empty(B * (nnz + spread), device=dev); yy = torch.empty(B, device=dev)
    ops.synth_sparse(rp, fid, fv, yy, nnz_mean=nnz, spread=spread, features=F, seed=5)
    for _ in range(a.steps):
        tr.step((rp, fid, fv, yy))
elif a.config == "mnist_chain":           # the four-kernel tcgen05 chain (fused step off)
    tr = MLPTrainer(ctx, batch_size=100, lr=0.0005, mode="sync", use_graph=False, fused_step=False)
    x = torch.empty(100, 784, device=dev); y = torch.empty(100, device=dev, dtype=torch.int32)
    ops.synth_mnist(x, y, seed=1)
    for _ in range(a.steps):
        tr.step(x, y)
elif a.config == "wd":
    F, B, nnz, spread = 1 << 22, 500, 100, 20
    tr = SparseWideDeepTrainer(ctx, features=F, dim=128, lr=0.05, mode="async", use_graph=False)
    rp = torch.empty(B + 1, device=dev, dtype=torch.int32); fid = torch.empty(B * (nnz + spread), device=dev, dtype=torch.int64)
    fv = torch.empty(B * (nnz + spread), device=dev); yy = torch.empty(B, device=dev)
    ops.synth_sparse(rp, fid, fv, yy, nnz_mean=nnz, spread=spread, features=F, seed=5)
    for _ in range(a.steps):
        tr.step((rp, fid, fv, yy))
torch.cuda.synchronize()
print("ok", tr.read_loss())
