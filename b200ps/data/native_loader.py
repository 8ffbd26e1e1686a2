"""Native libsvm batch loader (csrc/data/loader.cpp) and the *packed batch* it produces.

The reference reads, samples and batches its input in Python threads and re-formats every batch in
CPython before ``sess.run`` (lr2.py:69-169).  ``NativeLibsvmLoader`` does all of that in C++ reader
threads and hands out packed batches — one pinned buffer per batch with ``row_ptr | fid | fval |
labels`` at fixed 16-byte-aligned offsets — that a single H2D copy (``HostFeeder``) turns into the
operands of the sparse kernels.

    loader = NativeLibsvmLoader(files, batch_size=500, threads=8, epochs=1, sampling_rate=0.5)
    for batch in loader:                       # PackedBatch; returned to the ring when the loop moves on
        row_ptr, fid, fval, labels = batch.arrays()
    loader.stats()                             # lines, kept, sampled_out, malformed, ...

File sharding follows the reference: worker ``k`` of ``n`` takes ``files[k::n]`` (lr2.py:302-304, done by
the caller / ``shard_files``) and reader thread ``t`` takes ``files[t::threads]`` (lr2.py:232-233, done natively).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Iterator, List, Optional, Sequence

import numpy as np
import torch

from ..ops import native

MAGIC = 0xB200C5A0


@dataclass(frozen=True)
class PackedLayout:
    """Byte layout of a packed batch (identical for every batch of a loader / feeder)."""
    bytes: int
    off_rowptr: int
    off_fid: int
    off_fval: int
    off_labels: int
    rows: int
    cap_nnz: int

    @staticmethod
    def make(rows: int, cap_nnz: int) -> "PackedLayout":
        out = (C.c_int64 * 7)()
        native.check(native.load().b200ps_packed_layout(rows, cap_nnz, out), "packed_layout")
        return PackedLayout(*[int(v) for v in out])

    def views(self, buf: torch.Tensor):
        """(row_ptr int32[rows+1], fid int64[cap_nnz], fval f32[cap_nnz], labels f32[rows]) as views of a
        uint8 tensor (host or device) holding one packed batch."""
        r, z = self.rows, self.cap_nnz
        return (buf[self.off_rowptr:self.off_rowptr + 4 * (r + 1)].view(torch.int32),
                buf[self.off_fid:self.off_fid + 8 * z].view(torch.int64),
                buf[self.off_fval:self.off_fval + 4 * z].view(torch.float32),
                buf[self.off_labels:self.off_labels + 4 * r].view(torch.float32))

    def pack(self, row_ptr, fid, fval, labels, out: Optional[torch.Tensor] = None, pin: bool = True) -> torch.Tensor:
        """Pack CSR arrays (tensors or ndarrays, any device) into a host buffer of this layout."""
        if out is None:
            out = torch.zeros(self.bytes, dtype=torch.uint8)
            if pin and torch.cuda.is_available():
                out = out.pin_memory()
        rp, f, v, y = (torch.as_tensor(a).cpu() for a in (row_ptr, fid, fval, labels))
        n = int(rp[-1])
        if rp.numel() != self.rows + 1 or n > self.cap_nnz:
            raise ValueError("pack: batch does not fit the layout (rows %d, nnz %d)" % (rp.numel() - 1, n))
        d = self.views(out)
        d[0].copy_(rp.to(torch.int32)); d[1][:n].copy_(f[:n].to(torch.int64)); d[2][:n].copy_(v[:n].to(torch.float32))
        d[3].copy_(y.to(torch.float32).reshape(-1))
        hdr = out[:64].view(torch.int32)
        hdr[0] = MAGIC - (1 << 32)
        hdr[1], hdr[2], hdr[3], hdr[4] = self.rows, n, self.rows, self.cap_nnz
        hdr[5], hdr[6], hdr[7], hdr[8], hdr[9] = self.off_rowptr, self.off_fid, self.off_fval, self.off_labels, self.bytes
        return out


class PackedBatch:
    """One batch handed out by the loader: a host uint8 view of the ring buffer plus its header fields."""

    def __init__(self, loader: "NativeLibsvmLoader", ptr: int):
        self._loader, self._ptr = loader, ptr
        lay = loader.layout
        self._keep = (C.c_uint8 * lay.bytes).from_address(ptr)
        self.tensor = torch.frombuffer(self._keep, dtype=torch.uint8)
        hdr = np.frombuffer(self._keep, dtype=np.int32, count=16)
        if (int(hdr[0]) & 0xFFFFFFFF) != MAGIC:
            raise native.NativeError("packed batch without magic number")
        self.rows, self.nnz, self.epoch, self.thread = int(hdr[1]), int(hdr[2]), int(hdr[10]), int(hdr[11])
        self.seq = int(np.frombuffer(self._keep, dtype=np.int64, count=7)[6])

    def arrays(self):
        rp, fid, fval, y = self._loader.layout.views(self.tensor)
        return rp, fid[:self.nnz], fval[:self.nnz], y

    def release(self) -> None:
        if self._ptr:
            self._loader._release(self._ptr)
            self._ptr = 0


def shard_files(files: Sequence[str], num_workers: int, task_index: int) -> List[str]:
    """``files[task_index::num_workers]`` — the worker split of lr2.py:302-304."""
    return list(files[task_index::num_workers])


class NativeLibsvmLoader:
    def __init__(self, files: Sequence[str], batch_size: int, threads: int = 2, cap_nnz: Optional[int] = None,
                 max_nnz_per_row: int = 256, ring: int = 8, sampling_rate: float = 1.0, epochs: int = 1,
                 seed: int = 0, pinned: bool = True):
        from .libsvm import resolve_path
        self._lib = native.load()
        paths = [resolve_path(f).encode() for f in files]
        arr = (C.c_char_p * max(1, len(paths)))(*paths)
        cap = int(cap_nnz or batch_size * max_nnz_per_row)
        self._h = C.c_void_p()
        native.check(self._lib.b200ps_loader_create(C.byref(self._h), arr, len(paths), max(1, threads), batch_size, cap,
                                                    max(2, ring), float(sampling_rate), seed, epochs, int(pinned)),
                     "loader_create")
        out, pin = (C.c_int64 * 7)(), C.c_int32()
        native.check(self._lib.b200ps_loader_layout(self._h, out, C.byref(pin)), "loader_layout")
        self.layout = PackedLayout(*[int(v) for v in out])
        self.pinned = bool(pin.value)
        self.ring, self.threads = max(2, ring), max(1, threads)
        self._current: Optional[PackedBatch] = None

    def next(self, timeout_ms: int = -1) -> Optional[PackedBatch]:
        """The next packed batch, ``None`` at the end of the data; raises ``TimeoutError`` on timeout.  The
        caller releases it (``batch.release()``); iterating releases the previous batch automatically."""
        p = C.c_void_p()
        rc = self._lib.b200ps_loader_next(self._h, C.byref(p), timeout_ms)
        if rc == 1:
            return None
        if rc == 2:
            raise TimeoutError("native loader: no batch within %d ms" % timeout_ms)
        native.check(rc, "loader_next")
        return PackedBatch(self, p.value)

    def __iter__(self) -> Iterator[PackedBatch]:
        while True:
            if self._current is not None:
                self._current.release()
            self._current = self.next()
            if self._current is None:
                return
            yield self._current

    def _release(self, ptr: int) -> None:
        if self._h:
            native.check(self._lib.b200ps_loader_release(self._h, ptr), "loader_release")

    def stats(self) -> dict:
        out = (C.c_int64 * 7)()
        native.check(self._lib.b200ps_loader_stats(self._h, out), "loader_stats")
        names = ("lines", "kept", "sampled_out", "malformed", "dropped_nnz_cap", "batches", "dropped_incomplete_batch_rows")
        return dict(zip(names, (int(v) for v in out)))

    def close(self) -> None:
        if self._h:
            self._lib.b200ps_loader_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
