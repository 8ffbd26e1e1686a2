import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ps import ops
from b200ps.ops import native as _n
dev = "cuda"
B, H, C = 100, 100, 10
z2 = torch.randn(B, H, device=dev); b1 = torch.zeros(H, device=dev); w2 = torch.randn(H, C, device=dev); b2 = torch.zeros(C, device=dev)
y = torch.randint(0, C, (B,), device=dev, dtype=torch.int32)
loss = torch.zeros(1, device=dev); correct = torch.zeros(1, device=dev, dtype=torch.int32)
dz2 = torch.empty(B, H, device=dev); dw2 = torch.empty(H, C, device=dev); db2 = torch.empty(C, device=dev); db1 = torch.empty(H, device=dev)
dbg = torch.zeros(16 * 160, dtype=torch.int64, device=dev)
_n.load().b200ps_mlp_tail_set_debug(dbg.data_ptr())
for _ in range(3):
    dbg.zero_()
    ops.mlp_tail(z2, b1, w2, b2, y, loss_sum=loss, correct=correct, dz2=dz2, dw2=dw2, db2=db2, db1=db1)
torch.cuda.synchronize()
t = dbg.cpu().view(-1, 16)
t0 = min(int(r[0]) for r in t if r[0])
for b in range(14):
    r = [int(v) for v in t[b] if v]
    if r: print(b, "start+%d" % (r[0] - t0), [r[i + 1] - r[i] for i in range(len(r) - 1)], "total", r[-1] - r[0])
