"""CRC-32C + TF's masking, native (libb200ps.so, SSE4.2/slice-by-8) with a pure-Python fallback."""
from __future__ import annotations

_MASK_DELTA = 0xA282EAD8
_table = None
_native = None


def _py_table():
    global _table
    if _table is None:
        t = []
        for i in range(256):
            c = i
            for _ in range(8):
                c = (c >> 1) ^ 0x82F63B78 if c & 1 else c >> 1
            t.append(c)
        _table = t
    return _table


def _crc_py(data: bytes, crc: int = 0) -> int:
    t = _py_table()
    c = crc ^ 0xFFFFFFFF
    for b in data:
        c = (c >> 8) ^ t[(c ^ b) & 0xFF]
    return c ^ 0xFFFFFFFF


def _load_native():
    global _native
    if _native is None:
        try:
            from ..ops import native
            lib = native.load()
            _native = lib.b200ps_crc32c_extend
        except Exception:
            _native = False
    return _native


def crc32c(data, crc: int = 0) -> int:
    data = bytes(data) if not isinstance(data, (bytes, bytearray)) else data
    fn = _load_native()
    if fn:
        import ctypes
        buf = (ctypes.c_char * len(data)).from_buffer_copy(data) if len(data) else None
        return int(fn(crc, buf, len(data)))
    return _crc_py(data, crc)


def masked_crc32c(data) -> int:
    c = crc32c(data)
    return (((c >> 15) | (c << 17)) + _MASK_DELTA) & 0xFFFFFFFF


def unmask(masked: int) -> int:
    rot = (masked - _MASK_DELTA) & 0xFFFFFFFF
    return ((rot >> 17) | (rot << 15)) & 0xFFFFFFFF
