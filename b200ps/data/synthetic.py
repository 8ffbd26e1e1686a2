"""Synthetic data (P23): stands in for the MNIST download and the HDFS libsvm files.

Device generators are the K9 kernels (``ops.synth_mnist`` / ``ops.synth_sparse``); the CPU
versions below serve the gloo plumbing config and the tests, and ``write_libsvm_files``
materialises text shards so the loaders/parsers can be exercised end to end offline.
"""
from __future__ import annotations

import os
from typing import List

import numpy as np
import torch


class SyntheticMNIST:
    """``mnist.train.next_batch(n)`` look-alike over a learnable synthetic 28x28 problem:
    each class lights a fixed pseudo-random third of the pixels (same rule as the device kernel)."""

    def __init__(self, num_examples: int = 55000, dim: int = 784, classes: int = 10, seed: int = 1,
                 offset: int = 0, device="cpu", one_hot: bool = False):
        self.num_examples, self.dim, self.classes, self.seed, self.offset = num_examples, dim, classes, seed, offset
        self.device = torch.device(device)
        self.one_hot = one_hot
        self._epoch_pos = 0
        self.epochs_completed = 0
        self._images = None
        self._labels = None

    def _materialize(self):
        if self._images is not None:
            return
        if self.device.type == "cuda":
            from .. import ops
            x = torch.empty(self.num_examples, self.dim, device=self.device)
            y = torch.empty(self.num_examples, device=self.device, dtype=torch.int32)
            ops.synth_mnist(x, y, classes=self.classes, seed=self.seed, offset=self.offset)
        else:
            rng = np.random.RandomState((self.seed * 7919 + self.offset) % (2 ** 31))
            yl = rng.randint(0, self.classes, size=self.num_examples).astype(np.int32)
            pat = np.random.RandomState(self.seed).randint(0, 3, size=(self.classes, self.dim)) == 0
            noise = rng.rand(self.num_examples, self.dim).astype(np.float32)
            xs = np.where(pat[yl], 0.55 + 0.45 * noise, 0.25 * noise * noise).astype(np.float32)
            x, y = torch.from_numpy(xs), torch.from_numpy(yl)
        self._images, self._labels = x, y
        self._perm = torch.arange(self.num_examples, device=self.device)

    @property
    def images(self):
        self._materialize()
        return self._images

    @property
    def labels(self):
        self._materialize()
        if self.one_hot:
            return torch.nn.functional.one_hot(self._labels.long(), self.classes).float()
        return self._labels

    def next_batch(self, batch_size: int):
        """Sequential batches with a reshuffle at each epoch boundary (MNIST tutorial loader semantics)."""
        self._materialize()
        if self._epoch_pos + batch_size > self.num_examples:
            self.epochs_completed += 1
            self._perm = torch.randperm(self.num_examples, device=self.device)
            self._epoch_pos = 0
        idx = self._perm[self._epoch_pos:self._epoch_pos + batch_size]
        self._epoch_pos += batch_size
        x, y = self._images[idx], self._labels[idx]
        if self.one_hot:
            y = torch.nn.functional.one_hot(y.long(), self.classes).float()
        return x, y


class Datasets:
    def __init__(self, train, test):
        self.train, self.test = train, test


def read_data_sets(data_dir: str = "", one_hot: bool = False, device="cpu", seed: int = 1) -> Datasets:
    """``input_data.read_data_sets`` (example.py:60-62): real idx files when ``data_dir`` has them,
    otherwise the synthetic 55 000 / 10 000 split (there is no network to download MNIST)."""
    from . import mnist_idx
    real = mnist_idx.try_load(data_dir, one_hot=one_hot, device=device)
    if real is not None:
        return real
    return Datasets(SyntheticMNIST(55000, seed=seed, offset=0, device=device, one_hot=one_hot),
                    SyntheticMNIST(10000, seed=seed, offset=10_000_000, device=device, one_hot=one_hot))


def synthetic_sparse_batch(B: int, nnz_mean: int, spread: int, features: int, seed: int = 5, offset: int = 0):
    """CPU CSR batch with the same label model as the device generator (hash-defined teacher)."""
    rng = np.random.RandomState((seed * 104729 + offset) % (2 ** 31))
    lens = rng.randint(nnz_mean - spread, nnz_mean + spread + 1, size=B)
    row_ptr = np.zeros(B + 1, dtype=np.int32)
    np.cumsum(lens, out=row_ptr[1:])
    n = int(row_ptr[-1])
    fid = rng.randint(0, features, size=n).astype(np.int64)
    fval = (0.5 + rng.rand(n)).astype(np.float32)
    teacher = np.random.RandomState(seed).uniform(-1, 1, size=min(features, 1 << 22)).astype(np.float32)
    z = np.add.reduceat(teacher[fid % teacher.size] * fval, row_ptr[:-1]) if n else np.zeros(B)
    logit = 3.0 * z / np.sqrt(np.maximum(lens, 1))
    labels = (rng.rand(B) < 1 / (1 + np.exp(-logit))).astype(np.float32)
    return row_ptr, fid, fval, labels


def write_libsvm_files(directory: str, num_files: int, rows_per_file: int, nnz_mean: int = 20, spread: int = 5,
                       features: int = 1000, seed: int = 5, prefix: str = "part-") -> List[str]:
    """Write ``part-00000`` ... libsvm shards; returns their paths."""
    os.makedirs(directory, exist_ok=True)
    paths = []
    for k in range(num_files):
        rp, fid, fv, y = synthetic_sparse_batch(rows_per_file, nnz_mean, spread, features, seed, offset=k * rows_per_file)
        path = os.path.join(directory, f"{prefix}{k:05d}")
        with open(path, "w") as f:
            for i in range(rows_per_file):
                a, b = rp[i], rp[i + 1]
                sep = " " if i % 2 == 0 else "\t"          # the reference splits on either (lr2.py:58)
                f.write(str(int(y[i])) + sep + sep.join(f"{int(fid[j])}:{fv[j]:.6g}" for j in range(a, b)) + "\n")
        paths.append(path)
    return paths
