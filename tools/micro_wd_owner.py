"""Per-phase device time of the owner-side wide & deep step (torchrun, one rank per GPU)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200ps import ops
from b200ps.parallel.context import init_from_env
from b200ps.train import SparseWideDeepTrainer

ctx = init_from_env()
dev = ctx.device
F, B, nnz, spread = 1 << 22, 500, 100, 20
tr = SparseWideDeepTrainer(ctx, features=F, dim=128, lr=0.05, mode="async", owner=True, use_graph=False)
m, st = tr.model, tr.store
cap = B * (nnz + spread)
rp = torch.empty(B + 1, device=dev, dtype=torch.int32); fid = torch.empty(cap, device=dev, dtype=torch.int64)
fv = torch.empty(cap, device=dev); y = torch.empty(B, device=dev)
ops.synth_sparse(rp, fid, fv, y, nnz_mean=nnz, spread=spread, features=F, seed=5, offset=ctx.rank * B)
for _ in range(3):
    tr.step((rp, fid, fv, y))
torch.cuda.synchronize()
r = ctx.rank
e, w2 = st.sparse_var(m.E), st.sparse_var(m.W2); pe, p2 = st.sparse_var(m.P), st.sparse_var(m.P2)
ob = m._owner_buffers(st, cap)
bufs = m._buffers(B, dev, r)
lo, hi = st._sparse_owned(e, r); lo2, hi2 = st._sparse_owned(w2, r)
E_l, W_l, P_l, P2_l, W2_l = (st.sparse_shard(n, r) for n in (m.E, m.W, m.P, m.P2, m.W2))
gs = m._gs0
kw = dict(world=ctx.world, rank=r, cap=cap)
own = dict(inbox_table=ob["t_inbox"], cnt_table=ob["t_cnt"])
scratch = torch.zeros(1, device=dev)
# every argument is evaluated out here: the timed interval holds one wrapper call and its kernel
phases = {
    "null": lambda: scratch.add_(1.0),
    "request": lambda: ops.wd_owner_request(rp, fid, rows_per_shard=e.rows_per_shard, features=F, resp_table=ob["t_resp"],
                                            count=ob["count"][r], slot_tab=ob["slot_tab"][r], **kw, **own),
    "barrier": lambda: st.barrier(),
    "serve": lambda: ops.wd_owner_serve(resp_table=ob["t_resp"], emb_local=E_l, wide_local=W_l, dim=128, w2_table=w2.table,
                                        w2_rows_per_shard=w2.rows_per_shard, **kw, **own),
    "loss": lambda: ops.wd_owner_loss(rp, fid, fv, y, rows_per_shard=e.rows_per_shard, features=F, resp_table=ob["t_resp"],
                                      slot_tab=ob["slot_tab"][r], pend_table=pe.table, bias=m._tab["bias"],
                                      bias_pend=ob["bias_pend"], pred=bufs["pred"], loss_sum=bufs["loss"], global_step=gs,
                                      update=True, lr=0.05, **kw),
    "fold1": lambda: ops.wd_owner_fold(pend_local=P_l, emb_local=E_l, wide_local=W_l, dim=128, w2_table=w2.table,
                                       pend2_table=p2.table, w2_rows_per_shard=w2.rows_per_shard, **kw, **own),
    "fold2": lambda: ops.sparse_wd_apply(None, None, None, 0, 128, w2.table, w2.rows_per_shard, P2_l, W2_l, hi2 - lo2, 2),
}
order = ["null", "request", "barrier", "serve", "barrier", "loss", "barrier", "fold1", "barrier", "fold2", "barrier"]
acc = {k: 0.0 for k in phases}
n = 30
for it in range(n):
    for name in order:
        ctx.control_barrier() if name == "barrier" else None
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); phases[name](); e1.record(); torch.cuda.synchronize()
        acc[name] += e0.elapsed_time(e1) * 1e3
if ctx.rank == 0:
    print({k: round(v / n / (order.count(k)), 1) for k, v in acc.items()}, "us per launch (eager, one wrapper call + kernel between two events; null = a 1-element torch add)")
ctx.shutdown()
