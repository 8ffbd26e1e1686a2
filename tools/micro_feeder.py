"""Per-step host feeding rate of MLPTrainer.step_from_host_async (HostFeeder: graph + events) by pipeline depth."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200ps import ops
from b200ps.parallel.context import init_from_env
from b200ps.train import MLPTrainer

mode = "graph"
tr = MLPTrainer(init_from_env(virtual_world=1), batch_size=100, lr=0.0005, mode="sync", use_graph=True)
x = torch.empty(64, 100, 784, device="cuda"); y = torch.empty(64, 100, device="cuda", dtype=torch.int32)
ops.synth_mnist(x.view(-1, 784), y.view(-1), seed=1)
hx, hy = x.cpu().pin_memory(), y.cpu().pin_memory()
for i in range(3):
    tr.step(x[i], y[i])
batches = [(hx[i], hy[i]) for i in range(64)]
for depth in (1, 2, 4, 6):
    pending = []
    for rep in range(2):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        n = 2000
        for i in range(n):
            pending.append(tr.step_from_host_async(*batches[i & 63]))
            if len(pending) >= depth:
                pending.pop(0).result()
        while pending:
            pending.pop(0).result()
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print(mode, "depth", depth, "us/step %.2f" % (dt / n * 1e6), tr._feeder1.stats())
