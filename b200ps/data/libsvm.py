"""libsvm text input (P18, P24): ``Sample`` and the batch formatter of lr2.py:51-84, a native
whole-buffer CSR parser, and the local-file stand-in for ``tf.gfile.GFile`` (lr2.py:104,122).

HDFS itself is out of scope (no network in this environment): ``hdfs://host:port/path`` is
mapped onto ``$B200PS_HDFS_ROOT/path`` when that variable is set, and rejected otherwise.
"""
from __future__ import annotations

import ctypes as C
import os
import re
from typing import Iterator, List, Sequence

import numpy as np

_SPLIT = re.compile("[ \t]")


class Sample:
    """One libsvm line: ``label idx:val idx:val ...`` (lr2.py:51-67)."""
    __slots__ = ("label", "indices", "values")

    def __init__(self):
        self.label = -1
        self.indices: List[int] = []
        self.values: List[float] = []

    def parse_line_libsvm(self, line: str) -> "Sample":
        toks = _SPLIT.split(line.strip())
        self.label = int(toks[0])
        self.indices, self.values = [], []
        for item in toks[1:]:
            if not item:
                continue
            index, value = item.split(":")
            self.indices.append(int(index))
            self.values.append(float(value))
        return self

    @staticmethod
    def format_samples_sparse(samples: Sequence["Sample"]):
        """(labels[B,1], fids[nnz], fvals[nnz], sp_indices[nnz,2]=[row,fid], B) — lr2.py:69-84."""
        labels, fids, fvals, sp_indices = [], [], [], []
        for i, s in enumerate(samples):
            labels.append(s.label)
            fids += s.indices
            fvals += s.values
            sp_indices += [[i, idx] for idx in s.indices]
        size = len(samples)
        return np.reshape(np.asarray(labels), (size, 1)), fids, fvals, sp_indices, size

    @staticmethod
    def format_samples_csr(samples: Sequence["Sample"]):
        """The same batch as CSR arrays (what the GPU kernels consume)."""
        lens = np.fromiter((len(s.indices) for s in samples), dtype=np.int64, count=len(samples))
        row_ptr = np.zeros(len(samples) + 1, dtype=np.int32)
        np.cumsum(lens, out=row_ptr[1:])
        fid = np.fromiter((i for s in samples for i in s.indices), dtype=np.int64, count=int(row_ptr[-1]))
        fval = np.fromiter((v for s in samples for v in s.values), dtype=np.float32, count=int(row_ptr[-1]))
        labels = np.fromiter((s.label for s in samples), dtype=np.float32, count=len(samples))
        return row_ptr, fid, fval, labels


def csr_from_sparse_triple(sp_indices, fids, fvals, labels):
    """Convert the reference's feed tuple into CSR; rows are the first column of sp_indices."""
    rows = np.asarray([r for r, _ in sp_indices], dtype=np.int64)
    B = len(labels)
    row_ptr = np.zeros(B + 1, dtype=np.int32)
    np.add.at(row_ptr[1:], rows, 1)
    np.cumsum(row_ptr, out=row_ptr)
    return (row_ptr, np.asarray(fids, dtype=np.int64), np.asarray(fvals, dtype=np.float32),
            np.asarray(labels, dtype=np.float32).reshape(-1))


def parse_buffer(buf: bytes):
    """Native one-pass parse of a libsvm text buffer -> (row_ptr, fid, fval, labels)."""
    from ..ops import native
    lib = native.load()
    rows, nnz = C.c_uint64(), C.c_uint64()
    lib.b200ps_libsvm_count(buf, len(buf), C.byref(rows), C.byref(nnz))
    labels = np.empty(rows.value, dtype=np.float32)
    row_ptr = np.empty(rows.value + 1, dtype=np.int32)
    fid = np.empty(nnz.value, dtype=np.int64)
    fval = np.empty(nnz.value, dtype=np.float32)
    rc = lib.b200ps_libsvm_parse(buf, len(buf), labels.ctypes.data, row_ptr.ctypes.data, fid.ctypes.data,
                                 fval.ctypes.data, rows.value, nnz.value)
    if rc != 0:
        raise ValueError(f"malformed libsvm input at line {-rc}")
    n = int(row_ptr[-1]) if rows.value else 0
    return row_ptr, fid[:n], fval[:n], labels


def resolve_path(path: str) -> str:
    """Local path for a reference-style URI."""
    if path.startswith("hdfs://"):
        root = os.environ.get("B200PS_HDFS_ROOT")
        if not root:
            raise IOError(f"{path}: HDFS is not reachable offline; set B200PS_HDFS_ROOT to a local mirror "
                          f"or pass local paths")
        rest = path[len("hdfs://"):]
        rest = rest[rest.index("/"):] if "/" in rest else "/"
        return os.path.join(root, rest.lstrip("/"))
    if path.startswith("file://"):
        return path[len("file://"):]
    return path


class GFile:
    """Line-iterable file, the subset of ``tf.gfile.GFile`` the reference uses."""

    def __init__(self, name: str, mode: str = "r"):
        self.name = resolve_path(name)
        self._f = open(self.name, mode)

    def __iter__(self) -> Iterator[str]:
        return iter(self._f)

    def read(self, n: int = -1):
        return self._f.read(n)

    def close(self):
        self._f.close()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def list_files(pattern_or_dir: str, min_bytes: int = 0) -> List[str]:
    """Files under a directory (or matching a glob), sorted; ``min_bytes`` mirrors the
    ``$5>1000`` size filter of run_lr2.sh:42-43."""
    import glob
    p = resolve_path(pattern_or_dir)
    if os.path.isdir(p):
        cands = [os.path.join(dp, f) for dp, _, fs in os.walk(p) for f in fs]
    else:
        cands = glob.glob(p)
    return sorted(f for f in cands if os.path.isfile(f) and os.path.getsize(f) > min_bytes)


def count_lines(path: str) -> int:
    """What hdfs_test.py:11-16 does per thread."""
    n = 0
    with GFile(path) as f:
        for _ in f:
            n += 1
    return n
