"""A few launches of the resident-cluster MNIST step (K14), 50 steps each — the command profiled under ncu."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200ps import ops

dev = "cuda"
B, D, H, C = 100, 784, 100, 10
rows = 55000
x = torch.rand(rows, D, device=dev); y = torch.randint(0, C, (rows,), device=dev, dtype=torch.int32)
w1 = torch.randn(D, H, device=dev) * 0.1; b1 = torch.zeros(H, device=dev); w2 = torch.randn(H, C, device=dev) * 0.1; b2 = torch.zeros(C, device=dev)
loss = torch.zeros(1, device=dev); corr = torch.zeros(1, device=dev, dtype=torch.int32)
cur = torch.zeros(1, device=dev, dtype=torch.int32); gs = torch.zeros(1, device=dev, dtype=torch.int64)
for _ in range(4):
    ops.mlp_cluster_steps(x, y, w1, b1, w2, b2, loss=loss, correct=corr, steps=50, lr=1e-3, batch=B, cursor=cur,
                          global_step=gs, cursor_rw=cur, cursor_step=B, cursor_wrap=rows - rows % B)
torch.cuda.synchronize()
print("loss", loss.item(), "steps", gs.item())
