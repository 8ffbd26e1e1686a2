"""Owner-side wide & deep step on 2 VIRTUAL ranks of one GPU at the benchmark's sizes — the command profiled under ncu."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200ps import ops
from b200ps.models.wide_deep import SparseWideDeep
from b200ps.parallel.context import init_from_env
from b200ps.ps.store import ParameterStore

world = int(os.environ.get("WD_WORLD", "2"))
ctx = init_from_env(virtual_world=world)
F, B, nnz, spread = 1 << 22, 500, 100, 20
store = ParameterStore(ctx.pool)
model = SparseWideDeep(F, dim=128, wide=True, owner=True)
model.register(store); store.build(); store.initialize()
batches = {}
for r in range(world):
    cap = B * (nnz + spread)
    rp = torch.empty(B + 1, device="cuda", dtype=torch.int32); fid = torch.empty(cap, device="cuda", dtype=torch.int64)
    fv = torch.empty(cap, device="cuda"); y = torch.empty(B, device="cuda")
    ops.synth_sparse(rp, fid, fv, y, nnz_mean=nnz, spread=spread, features=F, seed=5, offset=r * B)
    batches[r] = (rp, fid, fv, y)
for _ in range(4):
    out = model._step_owner(store, batches, 0.05, sync=False)
torch.cuda.synchronize()
print("loss", out[0].item())
