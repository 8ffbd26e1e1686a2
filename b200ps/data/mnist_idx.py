"""Local MNIST idx reader (R17): used when ``MNIST_data/{job}_{task}`` already holds the four
idx(.gz) files; nothing is ever downloaded (example.py:62 downloads)."""
from __future__ import annotations

import gzip
import os
import struct

import numpy as np
import torch

_FILES = {"train_images": "train-images-idx3-ubyte", "train_labels": "train-labels-idx1-ubyte",
          "test_images": "t10k-images-idx3-ubyte", "test_labels": "t10k-labels-idx1-ubyte"}


def _open(path):
    if os.path.exists(path):
        return open(path, "rb")
    if os.path.exists(path + ".gz"):
        return gzip.open(path + ".gz", "rb")
    return None


def read_idx(f) -> np.ndarray:
    zero, dtype, ndim = struct.unpack(">HBB", f.read(4))
    if zero != 0 or dtype != 0x08:
        raise ValueError("not an unsigned-byte idx file")
    shape = struct.unpack(">" + "I" * ndim, f.read(4 * ndim))
    return np.frombuffer(f.read(), dtype=np.uint8).reshape(shape)


class _ArrayDataSet:
    def __init__(self, images, labels, one_hot, device):
        self.images = torch.from_numpy(images.reshape(len(images), -1).astype(np.float32) / 255.0).to(device)
        self._labels = torch.from_numpy(labels.astype(np.int32)).to(device)
        self.one_hot, self.num_examples = one_hot, len(images)
        self._perm = torch.randperm(self.num_examples, device=device)
        self._pos, self.epochs_completed = 0, 0

    @property
    def labels(self):
        return torch.nn.functional.one_hot(self._labels.long(), 10).float() if self.one_hot else self._labels

    def next_batch(self, n):
        if self._pos + n > self.num_examples:
            self.epochs_completed += 1
            self._perm = torch.randperm(self.num_examples, device=self.images.device)
            self._pos = 0
        idx = self._perm[self._pos:self._pos + n]
        self._pos += n
        y = self._labels[idx]
        return self.images[idx], (torch.nn.functional.one_hot(y.long(), 10).float() if self.one_hot else y)


def try_load(data_dir: str, one_hot: bool = False, device="cpu", validation_size: int = 5000):
    if not data_dir or not os.path.isdir(data_dir):
        return None
    arrs = {}
    for key, name in _FILES.items():
        f = _open(os.path.join(data_dir, name))
        if f is None:
            return None
        with f:
            arrs[key] = read_idx(f)
    from .synthetic import Datasets
    tr_x, tr_y = arrs["train_images"][validation_size:], arrs["train_labels"][validation_size:]
    return Datasets(_ArrayDataSet(tr_x, tr_y, one_hot, device),
                    _ArrayDataSet(arrs["test_images"], arrs["test_labels"], one_hot, device))
