"""Minimal protobuf wire-format encoder/decoder (no protoc, no TensorFlow).

Enough of proto3 to hand-build the handful of TensorFlow messages the reference's outputs
are made of: ``BundleHeaderProto`` / ``BundleEntryProto`` (Saver V2 checkpoints,
model_export.py:53), ``Event`` / ``Summary`` (example.py:131-135,171) and a skeletal
``MetaGraphDef`` with serving signatures (model_export.py:54-66).
"""
from __future__ import annotations

import struct
from typing import Dict, List, Tuple, Union

WT_VARINT, WT_64, WT_LEN, WT_32 = 0, 1, 2, 5


def varint(n: int) -> bytes:
    if n < 0:
        n += 1 << 64
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def tag(field: int, wt: int) -> bytes:
    return varint((field << 3) | wt)


def f_varint(field: int, v: int, keep_zero: bool = False) -> bytes:
    if v == 0 and not keep_zero:
        return b""
    return tag(field, WT_VARINT) + varint(int(v))


def f_bytes(field: int, b: Union[bytes, str], keep_empty: bool = False) -> bytes:
    if isinstance(b, str):
        b = b.encode("utf-8")
    if not b and not keep_empty:
        return b""
    return tag(field, WT_LEN) + varint(len(b)) + b


def f_msg(field: int, payload: bytes) -> bytes:
    """Sub-message with explicit presence (written even when empty)."""
    return tag(field, WT_LEN) + varint(len(payload)) + payload


def f_fixed32(field: int, v: int) -> bytes:
    return tag(field, WT_32) + struct.pack("<I", v & 0xFFFFFFFF)


def f_float(field: int, v: float, keep_zero: bool = False) -> bytes:
    if v == 0.0 and not keep_zero:
        return b""
    return tag(field, WT_32) + struct.pack("<f", v)


def f_double(field: int, v: float, keep_zero: bool = False) -> bytes:
    if v == 0.0 and not keep_zero:
        return b""
    return tag(field, WT_64) + struct.pack("<d", v)


def read_varint(buf: bytes, pos: int) -> Tuple[int, int]:
    shift, out = 0, 0
    while True:
        if pos >= len(buf):
            raise ValueError("truncated varint")
        b = buf[pos]
        pos += 1
        out |= (b & 0x7F) << shift
        if not b & 0x80:
            return out, pos
        shift += 7
        if shift > 70:
            raise ValueError("varint too long")


def decode(buf: bytes) -> Dict[int, List[Union[int, bytes]]]:
    """field number -> list of raw values (ints for varint/fixed, bytes for length-delimited)."""
    out: Dict[int, List[Union[int, bytes]]] = {}
    pos = 0
    while pos < len(buf):
        key, pos = read_varint(buf, pos)
        field, wt = key >> 3, key & 7
        if wt == WT_VARINT:
            v, pos = read_varint(buf, pos)
        elif wt == WT_64:
            v = struct.unpack_from("<Q", buf, pos)[0]
            pos += 8
        elif wt == WT_LEN:
            n, pos = read_varint(buf, pos)
            v = bytes(buf[pos:pos + n])
            if len(v) != n:
                raise ValueError("truncated field")
            pos += n
        elif wt == WT_32:
            v = struct.unpack_from("<I", buf, pos)[0]
            pos += 4
        else:
            raise ValueError(f"unsupported wire type {wt}")
        out.setdefault(field, []).append(v)
    return out


def as_float(u32: int) -> float:
    return struct.unpack("<f", struct.pack("<I", u32))[0]


def as_double(u64: int) -> float:
    return struct.unpack("<d", struct.pack("<Q", u64))[0]
