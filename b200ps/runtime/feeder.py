"""HostFeeder — pipelined host->device feeding of captured step graphs (csrc/runtime/feeder.cu).

The reference's worker loop hands every batch to the runtime through a feed_dict and blocks until
the fetches come back (example.py:165-170).  A ``HostFeeder`` keeps ``slots`` units in flight: the
H2D copy of unit t+1 runs on a copy stream under the kernels of unit t, graphs replay in order on
the compute stream, and the loss words are written by the step kernels straight into pinned host
memory — the host thread only blocks when it asks for a result (``Ticket.result()``).

A *unit* is whatever the bound graphs do with one filled staging slot: one training step, or a chunk
of K steps (``MLPTrainer.host_feeder(steps_per_call=K)``).
"""
from __future__ import annotations

import ctypes as C
from typing import Callable, Sequence

import numpy as np
import torch

from ..ops import native


def _load_fastfeed():
    """The optional CPython extension with the per-step fast paths (csrc/python/fastfeed.c)."""
    import importlib.util
    import os
    path = os.path.join(os.path.dirname(native.lib_path()), "_fastfeed.so")
    if not os.path.exists(path) or os.environ.get("B200PS_FASTFEED", "1") == "0":
        return None
    try:
        native.load()
        spec = importlib.util.spec_from_file_location("b200ps._fastfeed", path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        return mod
    except Exception:
        return None


_FF = None
_FF_TRIED = False


class Ticket:
    """Handle of one submitted unit; ``result()`` blocks until its result block is on the host."""
    __slots__ = ("_feeder", "_id")

    def __init__(self, feeder: "HostFeeder", ident: int):
        self._feeder, self._id = feeder, ident

    def result(self):
        f = self._feeder
        if f._fast_wait is not None:
            try:
                return f._fast_wait(f._hv, self._id)
            except RuntimeError as e:
                raise native.NativeError(str(e)) from None
        return f.wait(self._id)


class HostFeeder:
    """Ring of ``slots`` (staging buffers, graph, events) fed from pinned host memory.

    ``make_slot(slot, result)`` is called once per slot with a pinned+mapped ``result`` tensor
    (float32 view of the slot's result block); it must return ``(graph, staging)`` where ``graph``
    is a ``torch.cuda.CUDAGraph`` that consumes the ``staging`` device tensors (one or two) and
    writes its results into ``result``.
    """

    def __init__(self, make_slot: Callable, nbytes: Sequence[int], result_words: int, slots: int = 4,
                 decode: Callable = None):
        self._lib = native.load()
        self._h = C.c_void_p()
        nb = list(nbytes) + [0] * (2 - len(nbytes))
        native.check(self._lib.b200ps_feeder_create(C.byref(self._h), slots, nb[0], nb[1], 4 * result_words),
                     "feeder_create")
        self.slots, self.result_words = slots, result_words
        self._decode = decode
        self._keep = []
        self._out = np.zeros(result_words, dtype=np.float32)
        self._out_ptr = self._out.ctypes.data
        for s in range(slots):
            p = C.c_void_p()
            native.check(self._lib.b200ps_feeder_result_ptr(self._h, s, C.byref(p)), "feeder_result_ptr")
            buf = (C.c_float * result_words).from_address(p.value)
            result = torch.frombuffer(buf, dtype=torch.float32)        # host view; data_ptr is device-usable (UVA)
            graph, staging = make_slot(s, result)
            st = list(staging) + [None] * (2 - len(staging))
            native.check(self._lib.b200ps_feeder_bind(self._h, s, graph.raw_cuda_graph_exec(),
                                                      native.ptr(st[0]), native.ptr(st[1])), "feeder_bind")
            self._keep.append((graph, staging, result, buf))
        self._ticket = C.c_int64()
        self._submit = self._lib.b200ps_feeder_submit
        self._tref = C.byref(self._ticket)
        self._stream = native.stream_ptr()           # the compute stream the graphs were captured for
        global _FF, _FF_TRIED
        if not _FF_TRIED:
            _FF, _FF_TRIED = _load_fastfeed(), True
        self._hv = self._h.value
        self._fast_submit = _FF.submit if _FF is not None else None
        # a float straight from C when the unit's result is one loss word (first word of a block <= 256 bytes)
        self._fast_wait = _FF.wait_first if (_FF is not None and result_words <= 64 and getattr(decode, "first_word", False)) else None

    def submit(self, src0: torch.Tensor, src1: torch.Tensor = None, stream=None) -> Ticket:
        """Queue one unit from pinned host tensors; returns immediately."""
        if self._fast_submit is not None and stream is None:
            try:
                return Ticket(self, self._fast_submit(self._hv, src0.data_ptr(), 0 if src1 is None else src1.data_ptr(),
                                                      self._stream))
            except RuntimeError as e:
                raise native.NativeError(str(e)) from None
        rc = self._submit(self._h, src0.data_ptr(), 0 if src1 is None else src1.data_ptr(),
                          self._stream if stream is None else native.stream_ptr(stream), self._tref)
        if rc:
            native.check(rc, "feeder_submit")
        return Ticket(self, self._ticket.value)

    def stats(self) -> dict:
        """Host microseconds per submitted unit spent in each phase of ``submit`` (driver calls)."""
        ns, n = (C.c_double * 5)(), C.c_int64()
        native.check(self._lib.b200ps_feeder_stats(self._h, ns, C.byref(n)), "feeder_stats")
        k = max(1, n.value)
        names = ("wait_slot", "h2d_copies", "handoff", "graph_launch", "done_record")
        return {"units": n.value, **{name: ns[i] / k / 1e3 for i, name in enumerate(names)}}

    def wait(self, ticket: int):
        native.check(self._lib.b200ps_feeder_wait(self._h, ticket, self._out_ptr), "feeder_wait")
        return self._decode(self._out) if self._decode else self._out.copy()

    def close(self) -> None:
        if self._h:
            self._lib.b200ps_feeder_destroy(self._h)
            self._h = C.c_void_p()
            self._hv, self._fast_submit, self._fast_wait = 0, None, None
            self._keep.clear()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
