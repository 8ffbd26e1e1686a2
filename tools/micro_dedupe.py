"""Sparse LR step time, direct kernel vs the unique-id step (B200PS_SPARSE_DEDUPE=1), on uniform or skewed ids.

    python tools/micro_dedupe.py [--skew P]      (torchrun for several GPUs)

ids = F * u^P: P = 1 is the bench's uniform draw, P = 6 puts 10 % of a batch's non-zeros on one id.
Device-timed (CUDA events around 200 graph replays over a ring of 16 batches, max over ranks)."""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from b200ps import ops
from b200ps.parallel.context import init_from_env
from b200ps.train import SparseLRTrainer

ap = argparse.ArgumentParser()
ap.add_argument("--skew", type=float, default=1.0)
ap.add_argument("--features", type=int, default=1 << 24)
ap.add_argument("--mode", default="async")
args = ap.parse_args()
ctx = init_from_env()
dev = ctx.device
F, B, nnz, spread = args.features, 500, 100, 20
tr = SparseLRTrainer(ctx, features=F, lr=0.05, mode=args.mode)
cap = B * (nnz + spread)
ring = []
for i in range(16):
    rp = torch.empty(B + 1, device=dev, dtype=torch.int32); fid = torch.empty(cap, device=dev, dtype=torch.int64)
    fv = torch.empty(cap, device=dev); y = torch.empty(B, device=dev)
    ops.synth_sparse(rp, fid, fv, y, nnz_mean=nnz, spread=spread, features=F, seed=5, offset=(ctx.rank * 16 + i) * B)
    if args.skew != 1.0:
        g = torch.Generator(device=dev).manual_seed(100 + ctx.rank * 16 + i)
        u = torch.rand(cap, device=dev, generator=g, dtype=torch.float64)
        fid.copy_((u.pow(args.skew) * F).long().clamp_(0, F - 1))
    ring.append((rp, fid, fv, y))
for _ in range(3):
    for b in ring:
        tr.step(b)
torch.cuda.synchronize(); ctx.control_barrier()
n = 200
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
best = 1e9
for trial in range(5):
    ctx.control_barrier(); torch.cuda.synchronize()
    e0.record()
    for i in range(n):
        tr.step(ring[i % 16])
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) * 1e3 / n])
    if ctx.world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=ctx.group)
    best = min(best, t.item())
loss = tr.read_loss()
if ctx.rank == 0:
    m = tr.model
    extra = ""
    if m.dedupe:
        d, nz = m.last_unique_counts(cap, 0)
        extra = f" distinct/nnz {d}/{nz} = {d / max(nz, 1):.3f} phases(ns: hash,sync,gather,sync,examples,sync,scatter) {m.last_phase_times_ns(cap, 0)}"
    print(f"sparse LR {args.mode} world={ctx.world} skew={args.skew} dedupe={int(m.dedupe)}: {best:.2f} us/step "
          f"({B * ctx.world / best:.1f} M examples/s) loss {loss:.4f}{extra}")
ctx.shutdown()
