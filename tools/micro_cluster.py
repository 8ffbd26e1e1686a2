"""Phase timing of the resident-cluster MNIST step (K14): CTA-0 SM-clock stamps of the last step of a launch +
CUDA-event time per step for launches of 1 / 8 / 200 steps, 3xTF32 and plain tf32."""
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


def run(steps, plain=False):
    ops.mlp_cluster_steps(x, y, w1, b1, w2, b2, loss=loss, correct=corr, steps=steps, lr=1e-4, batch=B, cursor=cur,
                          global_step=gs, cursor_rw=cur, cursor_step=B, cursor_wrap=rows - rows % B, plain_tf32=plain)


names = ["labels+X wait", "forward+send", "sync1", "head a+send", "sync2", "head b+send", "sync3", "backward", "apply+replicas"]
clk = 1.965
for plain in (False, True):
    for _ in range(3):
        run(20, plain)
    torch.cuda.synchronize()
    lib.b200ps_mlp_cluster_set_timing(stamps.data_ptr())
    acc = torch.zeros(9, dtype=torch.float64)
    n = 20
    for _ in range(n):
        run(20, plain); torch.cuda.synchronize()
        s = stamps.cpu().double()
        acc += s[1:10] - s[:9]
    lib.b200ps_mlp_cluster_set_timing(None)
    print("plain tf32" if plain else "3xTF32", "CTA0 phases of the last step (us at %.3f GHz):" % clk,
          {k: round(v / n / clk / 1e3, 2) for k, v in zip(names, acc.tolist())}, "total", round(acc.sum().item() / n / clk / 1e3, 2))
    for steps in (1, 8, 200):
        for _ in range(3):
            run(steps, plain)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        reps = max(1, 400 // steps)
        for _ in range(reps):
            run(steps, plain)
        e1.record(); torch.cuda.synchronize()
        print("  launches of %d steps: %.2f us/step" % (steps, e0.elapsed_time(e1) * 1e3 / (reps * steps)))
