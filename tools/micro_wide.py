"""Per-kernel timing of the wide-MLP step pieces at the real shapes (CUDA events, back-to-back)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ps import ops

def t(fn, it=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(it): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / it

B, D = 4096, 8192
dev = "cuda"
x = torch.randn(B, D, device=dev).bfloat16(); w = (torch.randn(D, D, device=dev) * 0.02).bfloat16()
out = torch.empty(B, D, device=dev, dtype=torch.bfloat16); dz = (torch.randn(B, D, device=dev) * 0.01).bfloat16()
act = torch.rand(B, D, device=dev).bfloat16(); bias = torch.zeros(D, device=dev); cs = torch.zeros(D, device=dev)
master = torch.randn(D, D, device=dev) * 0.02; lowp = master.bfloat16(); grad = torch.empty(D, D, device=dev, dtype=torch.bfloat16)
labels = torch.randint(0, D, (B,), device=dev, dtype=torch.int32); loss = torch.zeros(1, device=dev); corr = torch.zeros(1, device=dev, dtype=torch.int32)
fl = 2.0 * B * D * D
res = {
 "fwd plain": t(lambda: ops.gemm(x, w, out, b_mn_major=True)),
 "fwd bias+sigmoid": t(lambda: ops.gemm(x, w, out, b_mn_major=True, bias=bias, act=ops.ACT_SIGMOID)),
 "dgrad plain": t(lambda: ops.gemm(dz, w, out)),
 "dgrad dsigmoid": t(lambda: ops.gemm(dz, w, out, dsigmoid_of=act)),
 "dgrad dsigmoid+colsum": t(lambda: ops.gemm(dz, w, out, dsigmoid_of=act, colsum_out=cs)),
 "wgrad store": t(lambda: ops.gemm(x, dz, grad, a_mn_major=True, b_mn_major=True)),
 "wgrad fused SGD": t(lambda: ops.gemm(x, dz, a_mn_major=True, b_mn_major=True, epilogue=ops.EPI_SGD, sgd_master=master, sgd_lowp=lowp, sgd_lr=0.01)),
 "cublas fwd": t(lambda: torch.matmul(x, w, out=out)),
 "softmax_ce": t(lambda: ops.softmax_ce(out, labels, dlogits=dz, loss_sum=loss, correct=corr)),
 "colsum": t(lambda: ops.colsum(dz, cs)),
 "sgd_apply 67M": t(lambda: ops.sgd_apply(master.view(-1), grad.view(-1), 0.01, lowp=lowp.view(-1))),
}
for k, v in res.items():
    print(f"{k:26s} {v*1e3:9.1f} us  {fl/v/1e9 if 'grad' in k or 'fwd' in k else 0:8.1f} TFLOP/s")
