"""Phase timing of the fused MNIST step kernel (K13): CTA-0 SM-clock stamps + CUDA-event time per launch."""
import torch
from b200ps import ops
from b200ps.ops import native

dev = "cuda"
B, D, H, C = 100, 784, 100, 10
rows = 55000
x = torch.rand(rows, D, device=dev); y = torch.randint(0, C, (rows,), device=dev, dtype=torch.int32)
w1 = torch.randn(D, H, device=dev); b1 = torch.zeros(H, device=dev); w2 = torch.randn(H, C, device=dev); b2 = torch.zeros(C, device=dev)
loss = torch.zeros(1, device=dev); corr = torch.zeros(1, device=dev, dtype=torch.int32)
cur = torch.zeros(1, device=dev, dtype=torch.int32); gs = torch.zeros(1, device=dev, dtype=torch.int64)
stamps = torch.zeros(24, device=dev, dtype=torch.int64)
lib = native.load()


def step():
    ops.mlp_fused_step(x, y, w1, b1, w2, b2, loss=loss, correct=corr, batch=B, cursor=cur, lr=1e-4, global_step=gs,
                       cursor_rw=cur, cursor_step=B, cursor_wrap=rows - rows % B)


for _ in range(20):
    step()
torch.cuda.synchronize()
lib.b200ps_mlp_fused_step_set_timing(stamps.data_ptr())
acc = torch.zeros(8, dtype=torch.float64); accb = torch.zeros(7, dtype=torch.float64)
n = 50
for _ in range(n):
    step(); torch.cuda.synchronize()
    s = stamps.cpu().double()
    acc += s[1:9] - s[:8]
    accb += torch.stack([s[9] - s[4], s[10] - s[9], s[11] - s[10], s[12] - s[11], s[13] - s[12], s[14] - s[13], s[5] - s[14]])
lib.b200ps_mlp_fused_step_set_timing(None)
names = ["pdl_wait", "fetch A", "phase A", "barrier 1", "phase B", "barrier 2", "phase C", "merge"]
clk = 1.9
print("CTA0 phases (us at %.1f GHz):" % clk, {k: round(v / n / clk / 1e3, 2) for k, v in zip(names, acc.tolist())},
      "total", round(acc.sum().item() / n / clk / 1e3, 2))
print("phase C: dz2 fetch %.2f, product+apply %.2f" % ((s[15] - s[6]) / clk / 1e3, (s[7] - s[15]) / clk / 1e3))
f = lambda a, b: round(float(s[a] - s[b]) / clk / 1e3, 2)
print("fetch A: issue all groups", f(19, 1), "wait g0", f(2, 19), "| phase C dz2: issue", f(17, 6), "wait", f(18, 17), "sync", f(15, 18))
print("phase B detail:", dict(zip(["wait+sync", "slab loads", "sigmoid+head+halving", "softmax", "dz2", "syncthreads", "partials"],
                                  [round(v / n / clk / 1e3, 2) for v in accb.tolist()])))
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    for _ in range(8):
        step()
for _ in range(5):
    g.replay()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); e0.record()
for _ in range(100):
    g.replay()
e1.record(); torch.cuda.synchronize()
print("graph of 8 steps: %.2f us/step" % (e0.elapsed_time(e1) * 1e3 / 800))
import os
os.environ["B200PS_PDL"] = "0"
