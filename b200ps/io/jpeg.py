"""JPEG decoding for the image input pipeline (``tf.read_file`` + ``tf.image.decode_jpeg(channels=3)``,
input_pipeline.py:68-78).

Two stages: the entropy-coded stream is decoded on the host by the native parser (``csrc/io/jpeg_parse.cpp``:
markers, Huffman, coefficient blocks), everything data parallel runs on the GPU (``csrc/kernels/jpeg.cu``:
dequantisation, inverse DCT, libjpeg-style "fancy" chroma upsampling, YCbCr -> RGB, optional bilinear resize).
On a machine without a CUDA device the second stage runs as the numpy restatement below (same arithmetic), which is
also the CPU oracle of the GPU kernels.  Baseline sequential JPEG only (what PIL / libjpeg write by default).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence, Tuple

import numpy as np
import torch

from ..ops import native


class JpegComp(C.Structure):
    _fields_ = [("h", C.c_int32), ("v", C.c_int32), ("blocks_w", C.c_int32), ("blocks_h", C.c_int32),
                ("coef_offset", C.c_int64), ("plane_offset", C.c_int64), ("quant", C.c_uint16 * 64)]


class JpegDesc(C.Structure):
    """Mirror of ``b200ps_jpeg_desc`` (include/b200ps.h)."""
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("ncomp", C.c_int32), ("hmax", C.c_int32),
                ("vmax", C.c_int32), ("mcus_x", C.c_int32), ("mcus_y", C.c_int32), ("total_blocks", C.c_int32),
                ("total_coefs", C.c_int64), ("comp", JpegComp * 3)]


MAX_PIXELS = int(os.environ.get("B200PS_JPEG_MAX_PIXELS", str(64 << 20)))


def jpeg_info(data: bytes) -> Tuple[int, int, int]:
    """(height, width, components) from the frame header."""
    lib = native.load()
    w, h, n = C.c_int32(), C.c_int32(), C.c_int32()
    native.check(lib.b200ps_jpeg_info(data, len(data), C.byref(w), C.byref(h), C.byref(n)), "jpeg_info")
    return h.value, w.value, n.value


def parse_jpeg(data: bytes):
    """Host stage: (descriptor, int16 coefficient array in natural order)."""
    lib = native.load()
    h, w, n = jpeg_info(data)
    if h * w > MAX_PIXELS:            # a damaged frame header can claim 65535 x 65535: refuse before allocating for it
        raise ValueError(f"jpeg: {w} x {h} pixels exceeds the limit of {MAX_PIXELS} (B200PS_JPEG_MAX_PIXELS)")
    cap = ((w + 15) // 16 * 16) * ((h + 15) // 16 * 16) * max(n, 1) + 64
    coef = np.empty(cap, dtype=np.int16)           # the parser zero-fills the total_coefs it reports
    desc = JpegDesc()
    native.check(lib.b200ps_jpeg_parse(data, len(data), coef.ctypes.data_as(C.c_void_p), cap, C.byref(desc)), "jpeg_parse")
    return desc, coef[: desc.total_coefs]


# ---- numpy restatement of the device stage (CPU mode + oracle of the kernels): libjpeg's integer arithmetic
def _islow_1d(x: np.ndarray, shift: int) -> np.ndarray:
    """libjpeg's 8-point "islow" inverse DCT along the last axis (int64 in, int64 out)."""
    i0, i1, i2, i3, i4, i5, i6, i7 = (x[..., k] for k in range(8))
    z1 = (i2 + i6) * 4433
    t2, t3 = z1 + i6 * (-15137), z1 + i2 * 6270
    t0, t1 = (i0 + i4) << 13, (i0 - i4) << 13
    t10, t13, t11, t12 = t0 + t3, t0 - t3, t1 + t2, t1 - t2
    a0, a1, a2, a3 = i7, i5, i3, i1
    z1, z2, z3, z4 = a0 + a3, a1 + a2, a0 + a2, a1 + a3
    z5 = (z3 + z4) * 9633
    a0, a1, a2, a3 = a0 * 2446, a1 * 16819, a2 * 25172, a3 * 12299
    z1, z2, z3, z4 = z1 * -7373, z2 * -20995, z3 * -16069 + z5, z4 * -3196 + z5
    a0, a1, a2, a3 = a0 + z1 + z3, a1 + z2 + z4, a2 + z2 + z3, a3 + z1 + z4
    r = 1 << (shift - 1)
    out = [t10 + a3, t11 + a2, t12 + a1, t13 + a0, t13 - a0, t12 - a1, t11 - a2, t10 - a3]
    return np.stack([(o + r) >> shift for o in out], axis=-1)


def _planes_numpy(desc: JpegDesc, coef: np.ndarray):
    planes = []
    for c in range(desc.ncomp):
        cc = desc.comp[c]
        q = np.array(cc.quant[:], dtype=np.int64).reshape(8, 8)
        nb = cc.blocks_w * cc.blocks_h
        F = coef[cc.coef_offset: cc.coef_offset + nb * 64].astype(np.int64).reshape(nb, 8, 8) * q     # [n, row, col]
        ws = _islow_1d(F.transpose(0, 2, 1), 11).transpose(0, 2, 1)          # pass 1 down the columns
        f = np.clip(_islow_1d(ws, 18) + 128, 0, 255).astype(np.int32)        # pass 2 along the rows
        f = f.reshape(cc.blocks_h, cc.blocks_w, 8, 8)
        planes.append(f.transpose(0, 2, 1, 3).reshape(cc.blocks_h * 8, cc.blocks_w * 8))
    return planes


def _component_numpy(desc: JpegDesc, planes, c: int) -> np.ndarray:
    cc = desc.comp[c]
    H, W = desc.height, desc.width
    hr, vr = desc.hmax // cc.h, desc.vmax // cc.v
    p = planes[c]
    if hr == 1 and vr == 1:
        return p[:H, :W]
    dw, dh = -(-W * cc.h // desc.hmax), -(-H * cc.v // desc.vmax)
    ys, xs = np.arange(H), np.arange(W)
    sy, sx = ys // vr, xs // hr
    if hr == 2 and vr == 1:
        s = p[sy][:, sx]
        nb = np.where(xs & 1, np.minimum(sx + 1, dw - 1), np.maximum(sx - 1, 0))
        o = p[sy][:, nb]
        return (3 * s + o + np.where(xs & 1, 2, 1)[None, :]) >> 2
    if hr == 2 and vr == 2:
        fy = np.where(ys & 1, np.minimum(sy + 1, dh - 1), np.maximum(sy - 1, 0))
        xo = np.where(xs & 1, np.minimum(sx + 1, dw - 1), np.maximum(sx - 1, 0))
        t_this = 3 * p[sy][:, sx] + p[fy][:, sx]
        t_other = 3 * p[sy][:, xo] + p[fy][:, xo]
        return (3 * t_this + t_other + np.where(xs & 1, 7, 8)[None, :]) >> 4
    return p[np.minimum(sy, dh - 1)][:, np.minimum(sx, dw - 1)]


def parse_jpeg_batch(datas: Sequence[bytes], threads: Optional[int] = None):
    """Host stage of many files at once: the entropy decoder is serial per file and releases the GIL (a ctypes
    call), so files are decoded on a thread pool — the stage that bounds an image pipeline (~120 Mpx/s per core).
    Returns [(descriptor, coefficients), ...] in input order."""
    n = len(datas)
    if n <= 1 or threads == 1:
        return [parse_jpeg(d) for d in datas]
    from concurrent.futures import ThreadPoolExecutor
    native.load()                                   # load the library once, outside the pool
    with ThreadPoolExecutor(max_workers=min(n, threads or min(8, os.cpu_count() or 1))) as pool:
        return list(pool.map(parse_jpeg, datas))


def decode_jpeg_numpy(data: bytes, _parsed=None) -> np.ndarray:
    """uint8 [H, W, 3] — the whole decoder on the host (numpy arithmetic identical to the kernels')."""
    desc, coef = parse_jpeg(data) if _parsed is None else _parsed
    planes = _planes_numpy(desc, coef)
    Y = _component_numpy(desc, planes, 0).astype(np.int64)
    if desc.ncomp == 1:
        return np.repeat(Y[:, :, None], 3, axis=2).astype(np.uint8)
    cb = _component_numpy(desc, planes, 1).astype(np.int64) - 128
    cr = _component_numpy(desc, planes, 2).astype(np.int64) - 128
    rgb = np.stack([Y + ((91881 * cr + 32768) >> 16), Y + ((-22554 * cb + 32768 - 46802 * cr) >> 16),
                    Y + ((116130 * cb + 32768) >> 16)], axis=2)             # libjpeg's fixed-point conversion
    return np.clip(rgb, 0, 255).astype(np.uint8)


def decode_jpeg(data: bytes, device="cuda", size: Optional[Tuple[int, int]] = None, stream=None,
                _parsed=None) -> torch.Tensor:
    """``tf.image.decode_jpeg(data, channels=3)``: uint8 tensor [H, W, 3] on ``device`` (``size=(h, w)`` adds a
    bilinear resize on the GPU).  Host Huffman decode, GPU inverse DCT + colour conversion."""
    dev = torch.device(device)
    if dev.type != "cuda":
        img = torch.from_numpy(decode_jpeg_numpy(data, _parsed))
        if size is not None and tuple(size) != tuple(img.shape[:2]):
            img = torch.nn.functional.interpolate(img.permute(2, 0, 1)[None].float(), size=tuple(size), mode="bilinear",
                                                  align_corners=False)[0].permute(1, 2, 0).round().clamp(0, 255).to(torch.uint8)
        return img
    lib = native.load()
    desc, coef = parse_jpeg(data) if _parsed is None else _parsed
    out_h, out_w = (desc.height, desc.width) if size is None else (int(size[0]), int(size[1]))
    with torch.cuda.device(dev):
        coef_dev = torch.from_numpy(coef).pin_memory().to(dev, non_blocking=True)
        planes = torch.empty(int(desc.total_coefs), dtype=torch.uint8, device=dev)
        out = torch.empty(out_h, out_w, 3, dtype=torch.uint8, device=dev)
        native.check(lib.b200ps_jpeg_idct_color(coef_dev.data_ptr(), C.byref(desc), planes.data_ptr(), out.data_ptr(),
                                                out_h, out_w, native.stream_ptr(stream)), "jpeg_idct_color")
    return out


def decode_jpeg_batch(datas: Sequence[bytes], device="cuda", size: Optional[Tuple[int, int]] = None) -> torch.Tensor:
    """A batch of equally sized (or resized to ``size``) images -> uint8 [N, H, W, 3].  The host (entropy) stage of
    the files runs on a thread pool; the device stage follows image by image."""
    parsed = parse_jpeg_batch(datas)
    imgs = [decode_jpeg(d, device, size, _parsed=p) for d, p in zip(datas, parsed)]
    return torch.stack(imgs)
