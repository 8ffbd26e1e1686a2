"""TensorFlow tensor-bundle (Saver V2) writer/reader without TensorFlow (R12, P25).

``tf.train.Saver().save`` (model_export.py:53; implicit in Supervisor) writes

    <prefix>.index                  leveldb-style sorted table:
                                      ""      -> BundleHeaderProto{num_shards, endianness, version}
                                      <name>  -> BundleEntryProto{dtype, shape, shard_id, offset, size, crc32c}
    <prefix>.data-00000-of-00001    raw little-endian tensor bytes, back to back

Table format (tensorflow/core/lib/io/table_builder.cc == leveldb): data blocks of
prefix-compressed entries with a restart array, each block followed by a 1-byte compression
type and a masked CRC-32C; a meta-index block; an index block; a 48-byte footer ending in
the magic 0xdb4775248b80fb57.  No TF install exists on this box to cross-validate against, so
the tests pin golden bytes derived from the format and round-trip through the independent
reader below.
"""
from __future__ import annotations

import os
import struct
from typing import Dict, List, Tuple

import numpy as np

from . import proto as pb
from .crc32c import masked_crc32c

TABLE_MAGIC = 0xDB4775248B80FB57
BLOCK_SIZE = 262144
RESTART_INTERVAL = 16

# tensorflow DataType enum values
DT = {"float32": 1, "float64": 2, "int32": 3, "uint8": 4, "int16": 5, "int8": 6, "int64": 9, "bool": 10,
      "uint16": 17, "float16": 19, "uint32": 22, "uint64": 23}
DT_BFLOAT16 = 14
DT_INV = {v: k for k, v in DT.items()}


# ------------------------------------------------------------------------------ table writer
class _BlockBuilder:
    def __init__(self):
        self.buf = bytearray()
        self.restarts = [0]
        self.counter = 0
        self.last_key = b""

    def add(self, key: bytes, value: bytes):
        shared = 0
        if self.counter < RESTART_INTERVAL:
            m = min(len(self.last_key), len(key))
            while shared < m and self.last_key[shared] == key[shared]:
                shared += 1
        else:
            self.restarts.append(len(self.buf))
            self.counter = 0
        self.buf += pb.varint(shared) + pb.varint(len(key) - shared) + pb.varint(len(value))
        self.buf += key[shared:] + value
        self.last_key = key
        self.counter += 1

    def size_estimate(self) -> int:
        return len(self.buf) + 4 * len(self.restarts) + 4

    def finish(self) -> bytes:
        out = bytes(self.buf) + b"".join(struct.pack("<I", r) for r in self.restarts)
        return out + struct.pack("<I", len(self.restarts))

    def empty(self) -> bool:
        return not self.buf


def _shortest_separator(start: bytes, limit: bytes) -> bytes:
    m = min(len(start), len(limit))
    d = 0
    while d < m and start[d] == limit[d]:
        d += 1
    if d < m:
        b = start[d]
        if b < 0xFF and b + 1 < limit[d]:
            return start[:d] + bytes([b + 1])
    return start


def _short_successor(key: bytes) -> bytes:
    for i, b in enumerate(key):
        if b != 0xFF:
            return key[:i] + bytes([b + 1])
    return key


def _handle(offset: int, size: int) -> bytes:
    return pb.varint(offset) + pb.varint(size)


def build_table(items: List[Tuple[bytes, bytes]]) -> bytes:
    """Serialise sorted (key, value) pairs as an uncompressed leveldb/TF table."""
    items = sorted(items, key=lambda kv: kv[0])
    out = bytearray()
    index = _BlockBuilder()
    block = _BlockBuilder()
    pending = None  # (last_key_of_flushed_block, handle)

    def write_block(b: _BlockBuilder) -> bytes:
        raw = b.finish()
        h = _handle(len(out), len(raw))
        out.extend(raw)
        trailer = b"\x00"
        out.extend(trailer + struct.pack("<I", masked_crc32c(raw + trailer)))
        return h

    last_key = b""
    for k, v in items:
        if pending is not None:
            index.add(_shortest_separator(pending[0], k), pending[1])
            pending = None
        block.add(k, v)
        last_key = k
        if block.size_estimate() >= BLOCK_SIZE:
            pending = (last_key, write_block(block))
            block = _BlockBuilder()
    if not block.empty():
        pending = (last_key, write_block(block))
    meta_h = write_block(_BlockBuilder())
    if pending is not None:
        index.add(_short_successor(pending[0]), pending[1])
    index_h = write_block(index)
    footer = meta_h + index_h
    footer += b"\x00" * (40 - len(footer))
    footer += struct.pack("<Q", TABLE_MAGIC)
    out.extend(footer)
    return bytes(out)


# ------------------------------------------------------------------------------ table reader
def _read_handle(buf: bytes, pos: int):
    off, pos = pb.read_varint(buf, pos)
    size, pos = pb.read_varint(buf, pos)
    return off, size, pos


def _read_block(data: bytes, off: int, size: int, verify: bool = True) -> bytes:
    raw = data[off: off + size]
    trailer = data[off + size: off + size + 5]
    if len(raw) != size or len(trailer) != 5:
        raise ValueError("truncated table block")
    if trailer[0] != 0:
        raise ValueError("compressed table blocks are not supported")
    if verify and struct.unpack("<I", trailer[1:])[0] != masked_crc32c(raw + trailer[:1]):
        raise ValueError("table block checksum mismatch")
    return raw


def _block_entries(raw: bytes) -> List[Tuple[bytes, bytes]]:
    n_restarts = struct.unpack_from("<I", raw, len(raw) - 4)[0]
    limit = len(raw) - 4 - 4 * n_restarts
    pos, key, out = 0, b"", []
    while pos < limit:
        shared, pos = pb.read_varint(raw, pos)
        non_shared, pos = pb.read_varint(raw, pos)
        vlen, pos = pb.read_varint(raw, pos)
        key = key[:shared] + raw[pos: pos + non_shared]
        pos += non_shared
        out.append((key, raw[pos: pos + vlen]))
        pos += vlen
    return out


def read_table(data: bytes) -> List[Tuple[bytes, bytes]]:
    if len(data) < 48 or struct.unpack("<Q", data[-8:])[0] != TABLE_MAGIC:
        raise ValueError("not a TF/leveldb table (bad magic)")
    footer = data[-48:]
    _mo, _ms, p = _read_handle(footer, 0)
    io, isz, _ = _read_handle(footer, p)
    out = []
    for _k, h in _block_entries(_read_block(data, io, isz)):
        off, size, _ = _read_handle(h, 0)
        out.extend(_block_entries(_read_block(data, off, size)))
    return out


# ------------------------------------------------------------------------------ bundle protos
def _header_proto(num_shards: int = 1) -> bytes:
    version = pb.f_varint(1, 1)                      # VersionDef.producer = kTensorBundleVersion
    return pb.f_varint(1, num_shards) + pb.f_msg(3, version)   # endianness LITTLE (0) is the default


def _shape_proto(shape) -> bytes:
    return b"".join(pb.f_msg(2, pb.f_varint(1, int(d), keep_zero=False)) for d in shape)


def _entry_proto(dtype: int, shape, offset: int, size: int, crc: int) -> bytes:
    return (pb.f_varint(1, dtype) + pb.f_msg(2, _shape_proto(shape)) + pb.f_varint(4, offset) +
            pb.f_varint(5, size) + pb.f_fixed32(6, crc))


def _np_dtype_enum(a: np.ndarray) -> int:
    name = a.dtype.name
    if name not in DT:
        raise TypeError(f"dtype {name} is not checkpointable")
    return DT[name]


def _to_numpy(t) -> np.ndarray:
    if isinstance(t, np.ndarray):
        return t
    try:
        import torch
        if isinstance(t, torch.Tensor):
            return t.detach().cpu().contiguous().numpy()
    except ImportError:
        pass
    return np.asarray(t)


def _native():
    """libb200ps.so when it has been built (csrc/io/bundle.cpp), else None; B200PS_NATIVE_BUNDLE=0 forces Python."""
    if os.environ.get("B200PS_NATIVE_BUNDLE", "1") == "0":
        return None
    try:
        from ..ops import native
        return native if native.available() else None
    except Exception:
        return None


def _write_bundle_native(native, prefix: str, tensors: Dict[str, object]) -> None:
    import ctypes as C
    lib = native.load()
    w = C.c_void_p()
    native.check(lib.b200ps_bundle_writer_create(C.byref(w), prefix.encode()), "bundle_writer_create")
    try:
        for name in sorted(tensors):
            a = _to_numpy(tensors[name])
            if not a.flags.c_contiguous:
                a = a.copy(order="C")
            if a.dtype.byteorder == ">":
                a = a.byteswap().newbyteorder()
            shape = (C.c_int64 * max(1, a.ndim))(*a.shape)
            native.check(lib.b200ps_bundle_writer_add(w, name.encode("utf-8"), _np_dtype_enum(a), shape, a.ndim,
                                                      a.ctypes.data, a.nbytes), "bundle_writer_add")
    except Exception:
        lib.b200ps_bundle_writer_abort(w)
        raise
    native.check(lib.b200ps_bundle_writer_finish(w), "bundle_writer_finish")


def write_bundle(prefix: str, tensors: Dict[str, object]) -> None:
    """Write ``tensors`` (name -> array) as ``<prefix>.index`` + ``<prefix>.data-00000-of-00001``.
    Files are written to a temporary name and renamed, so a reader never sees a torn checkpoint.
    With the native library present the data streams from the arrays to disk in C++ (one CRC pass, no
    ``tobytes`` copy); the pure-Python writer below produces the same bytes."""
    d = os.path.dirname(prefix)
    if d:
        os.makedirs(d, exist_ok=True)
    nat = _native()
    if nat is not None:
        return _write_bundle_native(nat, prefix, tensors)
    items = [(b"", _header_proto(1))]
    data_path = prefix + ".data-00000-of-00001"
    off = 0
    with open(data_path + ".tmp", "wb") as f:
        for name in sorted(tensors):
            a = _to_numpy(tensors[name])
            if not a.flags.c_contiguous:          # (ascontiguousarray would turn scalars into [1])
                a = a.copy(order="C")
            if a.dtype.byteorder == ">":
                a = a.byteswap().newbyteorder()
            raw = a.tobytes()
            f.write(raw)
            items.append((name.encode("utf-8"),
                          _entry_proto(_np_dtype_enum(a), a.shape, off, len(raw), masked_crc32c(raw))))
            off += len(raw)
    with open(prefix + ".index.tmp", "wb") as f:
        f.write(build_table(items))
    os.replace(data_path + ".tmp", data_path)
    os.replace(prefix + ".index.tmp", prefix + ".index")


def read_bundle_index(prefix: str):
    """name -> dict(dtype, shape, shard_id, offset, size, crc32c); plus the header."""
    with open(prefix + ".index", "rb") as f:
        entries = read_table(f.read())
    header, out = None, {}
    for k, v in entries:
        m = pb.decode(v)
        if k == b"":
            ver = pb.decode(m.get(3, [b""])[0])
            header = {"num_shards": m.get(1, [0])[0], "endianness": m.get(2, [0])[0],
                      "version_producer": ver.get(1, [0])[0]}
            continue
        shape = []
        for dim in pb.decode(m.get(2, [b""])[0]).get(2, []):
            shape.append(pb.decode(dim).get(1, [0])[0])
        out[k.decode("utf-8")] = {"dtype": m.get(1, [0])[0], "shape": tuple(shape),
                                  "shard_id": m.get(3, [0])[0], "offset": m.get(4, [0])[0],
                                  "size": m.get(5, [0])[0], "crc32c": m.get(6, [0])[0]}
    return header, out


def _read_bundle_native(native, prefix: str, verify: bool) -> Dict[str, np.ndarray]:
    import ctypes as C
    lib = native.load()
    r = C.c_void_p()
    rc = lib.b200ps_bundle_reader_open(C.byref(r), prefix.encode())
    if rc:
        raise ValueError(lib.b200ps_last_error().decode("utf-8", "replace"))
    out = {}
    try:
        n = C.c_int32()
        native.check(lib.b200ps_bundle_reader_count(r, C.byref(n)), "bundle_reader_count")
        name = C.create_string_buffer(4096)
        dt, nd, nb, shape = C.c_int32(), C.c_int32(), C.c_uint64(), (C.c_int64 * 8)()
        for i in range(n.value):
            native.check(lib.b200ps_bundle_reader_entry(r, i, name, 4096, C.byref(dt), shape, C.byref(nd), C.byref(nb)),
                         "bundle_reader_entry")
            key = name.value.decode("utf-8")
            if dt.value not in DT_INV:
                raise TypeError(f"{key}: unsupported dtype enum {dt.value}")
            a = np.empty(tuple(shape[k] for k in range(nd.value)), dtype=DT_INV[dt.value])
            if a.nbytes != nb.value:
                raise ValueError(f"{key}: entry size {nb.value} does not match shape {a.shape}")
            if lib.b200ps_bundle_reader_read(r, i, a.ctypes.data, a.nbytes, int(verify)):
                raise ValueError(lib.b200ps_last_error().decode("utf-8", "replace"))
            out[key] = a
    finally:
        lib.b200ps_bundle_reader_close(r)
    return out


def read_bundle(prefix: str, verify: bool = True) -> Dict[str, np.ndarray]:
    nat = _native()
    if nat is not None:
        return _read_bundle_native(nat, prefix, verify)
    header, index = read_bundle_index(prefix)
    if header is None:
        raise ValueError("bundle has no header entry")
    n = header["num_shards"]
    out = {}
    files = {}
    for name, e in index.items():
        sid = e["shard_id"]
        if sid not in files:
            files[sid] = open(f"{prefix}.data-{sid:05d}-of-{n:05d}", "rb")
        f = files[sid]
        f.seek(e["offset"])
        raw = f.read(e["size"])
        if len(raw) != e["size"]:
            raise ValueError(f"{name}: truncated data file")
        if verify and masked_crc32c(raw) != e["crc32c"]:
            raise ValueError(f"{name}: data checksum mismatch")
        if e["dtype"] not in DT_INV:
            raise TypeError(f"{name}: unsupported dtype enum {e['dtype']}")
        out[name] = np.frombuffer(raw, dtype=DT_INV[e["dtype"]]).reshape(e["shape"]).copy()
    for f in files.values():
        f.close()
    return out
