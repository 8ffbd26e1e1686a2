"""Sparse logistic regression (lr2.py's model) from libsvm files through the native loader.

    python examples/sparse_lr_from_files.py [--deep_dim 128]

Reader threads parse and batch the files into packed pinned buffers; every batch costs one H2D copy and one
graph launch (SparseLRTrainer.fit_from_loader); a streaming-AUC pass over held-out batches closes the run.
"""
import argparse
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ps.data.native_loader import NativeLibsvmLoader  # noqa: E402
from b200ps.data.synthetic import write_libsvm_files  # noqa: E402
from b200ps.metrics.auc import StreamingAUC  # noqa: E402
from b200ps.parallel.context import init_from_env  # noqa: E402
from b200ps.train import SparseLRTrainer, SparseWideDeepTrainer  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--deep_dim", type=int, default=0, help="128/256/512 adds the embedding + dense layer (wide & deep)")
ap.add_argument("--features", type=int, default=1 << 16)
args = ap.parse_args()

d = tempfile.mkdtemp(prefix="b200ps_example_")
files = write_libsvm_files(os.path.join(d, "data"), 5, 5000, nnz_mean=30, spread=10, features=args.features, seed=5)
train, test = files[:4], files[4:]                    # same generator (same hidden teacher), held-out file

ctx = init_from_env()
if args.deep_dim:
    tr = SparseWideDeepTrainer(ctx, args.features, dim=args.deep_dim, lr=1.0, mode="async")
else:
    tr = SparseLRTrainer(ctx, args.features, lr=1.0, mode="async")
tr.store.sparse_shard(tr.model.W).mul_(0.01)          # start near zero so that a short run shows learning

for epoch in range(8):
    loader = NativeLibsvmLoader(train, batch_size=500, threads=4, max_nnz_per_row=64, epochs=1, seed=epoch)
    losses = []
    steps = tr.fit_from_loader(loader, on_loss=lambda i, v: losses.append(v))
    print("epoch %d: %d steps, loss %.4f -> %.4f, loader %s" % (epoch, steps, losses[0], losses[-1], loader.stats()))
    loader.close()

auc = StreamingAUC(200, device=ctx.device)
loader = NativeLibsvmLoader(test, batch_size=500, threads=1, max_nnz_per_row=64, epochs=1)
for batch in loader:
    dev_batch = tuple(t.to(ctx.device) for t in batch.arrays())
    n = batch.nnz
    pred, _ = tr.predict((dev_batch[0], dev_batch[1][:n], dev_batch[2][:n], dev_batch[3]))
    auc.update(pred, dev_batch[3])
loader.close()
print("test AUC %.4f  global_step %d" % (auc.result(), tr.global_step))
ctx.close()
