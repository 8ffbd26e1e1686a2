"""Effective NVLink bandwidth of the K2 gather role: a fused pull+GEMM with a tiny M, so the launch time
is the time to stream the remote row panels into local HBM.  Run under torchrun."""
import os, sys, json
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ps import ops
from b200ps.parallel.context import init_from_env

ctx = init_from_env()
dev, N, r = ctx.device, ctx.world, ctx.rank
K = Nn = 8192
buf = ctx.pool.alloc("w", K * Nn * 2)
w = buf.view(r, torch.bfloat16).view(K, Nn)
w.normal_()
table = ops.device_ptr_table(buf.ptrs, dev)
flags = torch.zeros(K // 64 + 1, device=dev, dtype=torch.int32)
x = torch.randn(128, K, device=dev).bfloat16()
out = torch.empty(128, Nn, device=dev, dtype=torch.bfloat16)
ctx.control_barrier()
remote = K * Nn * 2 * (N - 1) / N
for G in (4, 8, 12, 16, 24, 32, 48):
    fn = lambda: ops.gemm(x, w, out, b_mn_major=True, gather=dict(src_table=table, flags=flags, world=N, rank=r, ctas=G))
    for _ in range(3):
        fn()
    ctx.control_barrier(); torch.cuda.synchronize()
    st, en = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    st.record()
    for _ in range(10):
        fn()
    en.record(); torch.cuda.synchronize()
    us = ctx.all_reduce_max(st.elapsed_time(en) / 10 * 1e3)
    if r == 0:
        print(json.dumps({"gather_ctas": G, "us": round(us, 1), "remote_MB": round(remote / 1e6, 1), "GBs": round(remote / us / 1e3, 1)}), flush=True)
    ctx.control_barrier()
ctx.shutdown()
