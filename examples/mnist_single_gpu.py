"""MNIST-shaped sigmoid MLP (example.py's model) on one GPU through the three step APIs.

    python examples/mnist_single_gpu.py [--mode sync|async]

1. device batches            tr.step(x, y)
2. device-resident epoch     tr.attach_dataset(X, Y); tr.run_resident(n)      (no host work per step)
3. pinned host batches       tr.step_from_host_async(x_pinned, y_pinned)      (HostFeeder pipeline)
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ps import ops  # noqa: E402
from b200ps.parallel.context import init_from_env  # noqa: E402
from b200ps.train import MLPTrainer  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--mode", default="sync", choices=["sync", "async"])
args = ap.parse_args()

ctx = init_from_env()                      # single process; under torchrun: one rank per GPU
dev = ctx.device
B, n_batches = 100, 550                    # example.py:43-45: batch 100, 55 000 training images
X = torch.empty(n_batches * B, 784, device=dev)
Y = torch.empty(n_batches * B, device=dev, dtype=torch.int32)
ops.synth_mnist(X, Y, seed=1)              # on-device stand-in for the MNIST download (no network)

tr = MLPTrainer(ctx, batch_size=B, lr=0.01, mode=args.mode)
for i in range(3):                         # 1. device batches (also the eager warm-up before graphs are captured)
    tr.step(X[i * B:(i + 1) * B], Y[i * B:(i + 1) * B])
print("loss after 3 steps: %.4f  (one-launch step: %s)" % (tr.read_loss(), tr.fused))

tr.attach_dataset(X, Y)                    # 2. resident epoch: a device cursor walks it, 8 steps per graph launch
tr.run_resident(2 * n_batches)
print("loss after 2 resident epochs: %.4f  global_step %d" % (tr.read_loss(), tr.global_step))

hx = X[:64 * B].view(64, B, 784).cpu().pin_memory()      # 3. host batches: H2D on a copy stream under the kernels
hy = Y[:64 * B].view(64, B).cpu().pin_memory()
pending, last = [], None
for i in range(256):
    pending.append(tr.step_from_host_async(hx[i % 64], hy[i % 64]))
    if len(pending) > 4:
        last = pending.pop(0).result()
last = [t.result() for t in pending][-1]
acc, test_loss = tr.evaluate(X[:10000], Y[:10000])
print("loss %.4f  accuracy on 10 000 images %.3f  global_step %d" % (last, acc, tr.global_step))
ctx.close()
