"""TensorBoard event files (R14, P14) without TensorFlow.

example.py writes scalar summaries ``cost`` and ``accuracy`` every step into
``./logs/mnist/{job}_{task}`` (example.py:48-51,131-135,154,171).  This writer emits the same
on-disk format — TFRecord framing (length, masked CRC-32C of the length, payload, masked
CRC-32C of the payload) around ``Event`` protos, first record ``file_version: "brain.Event:2"``
— from a background thread, so the training loop never blocks on file I/O (the reference
does the write synchronously inside its hot loop).
"""
from __future__ import annotations

import os
import queue
import socket
import struct
import threading
import time
from typing import Dict, Iterator, List, Optional, Tuple

from . import proto as pb
from .crc32c import masked_crc32c


def _native_lib():
    """libb200ps.so when built (native writer, csrc/io/events.cpp); B200PS_NATIVE_EVENTS=0 forces Python."""
    if os.environ.get("B200PS_NATIVE_EVENTS", "1") == "0":
        return None
    try:
        from ..ops import native
        return native if native.available() else None
    except Exception:
        return None


def _record(payload: bytes) -> bytes:
    hdr = struct.pack("<Q", len(payload))
    return hdr + struct.pack("<I", masked_crc32c(hdr)) + payload + struct.pack("<I", masked_crc32c(payload))


def _event(wall_time: float, step: int = 0, file_version: Optional[str] = None,
           summary: Optional[bytes] = None) -> bytes:
    out = pb.f_double(1, wall_time, keep_zero=True) + pb.f_varint(2, step)
    if file_version is not None:
        out += pb.f_bytes(3, file_version)
    if summary is not None:
        out += pb.f_msg(5, summary)
    return out


def _scalar_summary(values: Dict[str, float]) -> bytes:
    out = b""
    for tag, v in values.items():
        out += pb.f_msg(1, pb.f_bytes(1, tag) + pb.f_float(2, float(v), keep_zero=True))
    return out


class SummaryWriter:
    """``tf.train.SummaryWriter(logdir)``: ``add_summary({'cost': c, 'accuracy': a}, step)``."""

    def __init__(self, logdir: str, max_queue: int = 1000, flush_secs: float = 2.0, filename_suffix: str = ""):
        os.makedirs(logdir, exist_ok=True)
        self.logdir = logdir
        name = "events.out.tfevents.%010d.%s%s" % (int(time.time()), socket.gethostname(), filename_suffix)
        self.path = os.path.join(logdir, name)
        self._closed = False
        self._native = None
        nat = _native_lib()
        if nat is not None:                  # csrc/io/events.cpp: encoding, framing and the writer thread in C++
            import ctypes as C
            self._C, self._lib, self._check = C, nat.load(), nat.check
            h = C.c_void_p()
            nat.check(self._lib.b200ps_events_open(C.byref(h), self.path.encode(), float(flush_secs)), "events_open")
            self._native = h
            return
        self._f = open(self.path, "wb")
        self._q: "queue.Queue" = queue.Queue(max_queue)
        self._flush_secs = flush_secs
        self._closed = False
        self._f.write(_record(_event(time.time(), 0, file_version="brain.Event:2")))
        self._f.flush()
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()

    def _run(self):
        last = time.time()
        while True:
            try:
                item = self._q.get(timeout=self._flush_secs)
            except queue.Empty:
                item = b""
            if item is None:
                break
            if item:
                self._f.write(item)
            if time.time() - last >= self._flush_secs or self._q.empty():
                self._f.flush()
                last = time.time()
        self._f.flush()

    def add_summary(self, values: Dict[str, float], global_step: int, wall_time: Optional[float] = None) -> None:
        if self._closed:
            raise RuntimeError("SummaryWriter is closed")
        if self._native is not None:
            C, n = self._C, len(values)
            tags = (C.c_char_p * max(1, n))(*[str(t).encode("utf-8") for t in values])
            vals = (C.c_float * max(1, n))(*[float(v) for v in values.values()])
            self._check(self._lib.b200ps_events_add_scalars(self._native, -1.0 if wall_time is None else float(wall_time),
                                                            int(global_step), tags, vals, n), "events_add_scalars")
            return
        ev = _event(time.time() if wall_time is None else wall_time, int(global_step),
                    summary=_scalar_summary(values))
        self._q.put(_record(ev))

    add_scalars = add_summary

    def flush(self) -> None:
        if self._native is not None:
            self._check(self._lib.b200ps_events_flush(self._native), "events_flush")
            return
        while not self._q.empty():
            time.sleep(0.005)
        self._f.flush()

    def close(self) -> None:
        if not self._closed and self._native is not None:
            self._closed = True
            self._lib.b200ps_events_close(self._native)
            self._native = None
            return
        if not self._closed:
            self._closed = True
            self._q.put(None)
            self._t.join(timeout=10)
            self._f.close()


def read_events(path: str) -> Iterator[Tuple[float, int, Dict[str, float], Optional[str]]]:
    """Yield (wall_time, step, {tag: value}, file_version) for every record; verifies both CRCs."""
    with open(path, "rb") as f:
        data = f.read()
    pos = 0
    while pos < len(data):
        hdr = data[pos:pos + 8]
        (n,) = struct.unpack("<Q", hdr)
        (c1,) = struct.unpack("<I", data[pos + 8:pos + 12])
        if c1 != masked_crc32c(hdr):
            raise ValueError("event file: bad length checksum")
        payload = data[pos + 12:pos + 12 + n]
        (c2,) = struct.unpack("<I", data[pos + 12 + n:pos + 16 + n])
        if c2 != masked_crc32c(payload):
            raise ValueError("event file: bad payload checksum")
        pos += 16 + n
        m = pb.decode(payload)
        wall = pb.as_double(m[1][0]) if 1 in m else 0.0
        step = m.get(2, [0])[0]
        fv = m[3][0].decode() if 3 in m else None
        vals = {}
        for sm in m.get(5, []):
            for v in pb.decode(sm).get(1, []):
                vm = pb.decode(v)
                vals[vm[1][0].decode()] = pb.as_float(vm.get(2, [0])[0])
        yield wall, step, vals, fv


def find_event_files(logdir: str) -> List[str]:
    return sorted(os.path.join(logdir, f) for f in os.listdir(logdir) if f.startswith("events.out.tfevents."))
