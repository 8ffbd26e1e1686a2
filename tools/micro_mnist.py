"""Where do the microseconds of the MNIST step go?  Times each piece inside CUDA graphs."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ps import ops  # noqa: E402
from b200ps.parallel.context import init_from_env  # noqa: E402
from b200ps.train import MLPTrainer  # noqa: E402


def graph_time(fn, reps=50, iters=20):
    fn(); fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    g.replay(); torch.cuda.synchronize()
    st, en = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    st.record()
    for _ in range(iters):
        g.replay()
    en.record(); torch.cuda.synchronize()
    return st.elapsed_time(en) / iters / reps * 1e3


ctx = init_from_env()
dev = ctx.device
tr = MLPTrainer(ctx, batch_size=100, lr=0.0005, mode="sync", use_graph=False)
st = tr.store
m = tr.model
x = torch.empty(100, 784, device=dev); y = torch.empty(100, device=dev, dtype=torch.int32)
ops.synth_mnist(x, y, seed=1)
tr.step(x, y)
b = m._buffers(100, dev, 0)
z2s, dz2, loss, correct = b["z2"], b["dz2"], b["loss"], b["correct"]
z2 = z2s[0]
w1 = st.var(m.W1)
res = {}
res["fill z2"] = graph_time(lambda: z2.zero_())
res["gemm1 splitk5 atomic"] = graph_time(lambda: ops.gemm(x, w1, z2, b_mn_major=True, epilogue=ops.EPI_ATOMIC_ADD, split_k=5))
res["gemm1 splitk1 store"] = graph_time(lambda: ops.gemm(x, w1, z2, b_mn_major=True))
res["gemm1 splitk5 slabs"] = graph_time(lambda: ops.gemm(x, w1, z2, b_mn_major=True, split_k=5, c_split_stride=z2s.stride(0)))
res["gemm1 splitk13 atomic"] = graph_time(lambda: ops.gemm(x, w1, z2, b_mn_major=True, epilogue=ops.EPI_ATOMIC_ADD, split_k=13))
res["tail"] = graph_time(lambda: ops.mlp_tail(z2, st.var(m.B1), st.var(m.W2), st.var(m.B2), y, loss_sum=loss, correct=correct,
                                              dz2=dz2, dw2=st.grad(m.W2), db2=st.grad(m.B2), db1=st.grad(m.B1), overwrite_stats=True,
                                              z2_splits=5, z2_split_stride=z2s.stride(0)))
res["gemm dW1"] = graph_time(lambda: ops.gemm(x, dz2, st.grad(m.W1), a_mn_major=True, b_mn_major=True))
res["push_sync"] = graph_time(lambda: st.push_sync(0.0005))
res["sgd_apply (plain)"] = graph_time(lambda: ops.sgd_apply(st.flat_work("fp32"), st.flat_grads("fp32"), 0.0005))
res["whole body"] = graph_time(lambda: tr._body(), reps=10)
res["copy x D2D"] = graph_time(lambda: tr.x[0].copy_(x))
for k, v in res.items():
    print(f"{k:28s} {v:8.2f} us")

