"""Persistence / observability (P14, P25, P26, P28): checkpoints, export, events, logging — CPU only."""
import os
import re
import struct

import numpy as np
import pytest
import torch
from hypothesis import given, settings, strategies as st

from b200ps.io import crc32c, events, exporter, log, proto as pb, saver, tf_bundle


def test_crc32c_known_vectors_native_and_python():
    # RFC 3720 test vectors
    assert crc32c.crc32c(b"123456789") == 0xE3069283
    assert crc32c.crc32c(b"\x00" * 32) == 0x8A9136AA
    assert crc32c.crc32c(b"\xff" * 32) == 0x62A8AB43
    assert crc32c._crc_py(bytes(range(32))) == 0x46DD794E == crc32c.crc32c(bytes(range(32)))
    m = crc32c.masked_crc32c(b"hello")
    assert crc32c.unmask(m) == crc32c.crc32c(b"hello")


def test_bundle_roundtrip_names_shapes_dtypes(tmp_path):
    rng = np.random.RandomState(0)
    t = {"global_step": np.float32(12), "weights/Variable": rng.randn(784, 100).astype(np.float32),
         "weights/Variable_1": rng.randn(100, 10).astype(np.float32), "biases/Variable": np.zeros(100, np.float32),
         "biases/Variable_1": np.zeros(10, np.float32), "counter": np.arange(5, dtype=np.int64)}
    prefix = str(tmp_path / "model.ckpt-12")
    tf_bundle.write_bundle(prefix, t)
    assert sorted(os.listdir(tmp_path)) == ["model.ckpt-12.data-00000-of-00001", "model.ckpt-12.index"]
    header, index = tf_bundle.read_bundle_index(prefix)
    assert header == {"num_shards": 1, "endianness": 0, "version_producer": 1}
    assert index["global_step"]["shape"] == () and index["global_step"]["dtype"] == 1          # DT_FLOAT scalar
    assert index["weights/Variable"]["shape"] == (784, 100) and index["counter"]["dtype"] == 9  # DT_INT64
    # entries are laid out back to back in key order
    off = 0
    for name in sorted(t):
        assert index[name]["offset"] == off
        off += index[name]["size"]
    assert off == os.path.getsize(prefix + ".data-00000-of-00001")
    back = tf_bundle.read_bundle(prefix)
    assert all(np.array_equal(back[k], t[k]) and back[k].dtype == np.asarray(t[k]).dtype for k in t)


def test_bundle_golden_bytes(tmp_path):
    """Pin the on-disk format: header proto, entry proto, block trailer, footer magic."""
    prefix = str(tmp_path / "g")
    tf_bundle.write_bundle(prefix, {"b": np.array([1.0, 2.0], np.float32)})
    raw = open(prefix + ".index", "rb").read()
    assert raw[-8:] == struct.pack("<Q", 0xDB4775248B80FB57)
    data = struct.pack("<ff", 1.0, 2.0)
    assert open(prefix + ".data-00000-of-00001", "rb").read() == data
    header = bytes([0x08, 0x01, 0x1A, 0x02, 0x08, 0x01])            # num_shards=1, version{producer=1}
    entry = (bytes([0x08, 0x01, 0x12, 0x04, 0x12, 0x02, 0x08, 0x02, 0x28, 0x08, 0x35]) +
             struct.pack("<I", crc32c.masked_crc32c(data)))         # dtype, shape{dim{size=2}}, size=8, crc
    block = (bytes([0, 0, len(header)]) + header + bytes([0, 1, len(entry)]) + b"b" + entry +
             struct.pack("<II", 0, 1))                               # restart array [0], count 1
    assert raw.startswith(block + b"\x00" + struct.pack("<I", crc32c.masked_crc32c(block + b"\x00")))
    # index block key is the short successor of the last key ("b" -> "c")
    entries = dict(tf_bundle.read_table(raw))
    assert set(entries) == {b"", b"b"} and entries[b""] == header and entries[b"b"] == entry


def test_bundle_detects_corruption(tmp_path):
    prefix = str(tmp_path / "c")
    tf_bundle.write_bundle(prefix, {"w": np.ones(64, np.float32)})
    p = prefix + ".data-00000-of-00001"
    raw = bytearray(open(p, "rb").read()); raw[10] ^= 0xFF; open(p, "wb").write(raw)
    with pytest.raises(ValueError, match="checksum"):
        tf_bundle.read_bundle(prefix)
    idx = bytearray(open(prefix + ".index", "rb").read()); idx[3] ^= 0x01; open(prefix + ".index", "wb").write(idx)
    with pytest.raises(ValueError):
        tf_bundle.read_bundle_index(prefix)


@settings(max_examples=25, deadline=None)
@given(st.dictionaries(st.text(alphabet="abcdefghij/_0123456789", min_size=1, max_size=40),
                       st.lists(st.floats(-1e6, 1e6, width=32), max_size=50), min_size=1, max_size=40))
def test_table_fuzz_roundtrip(d):
    items = [(k.encode(), struct.pack(f"<{len(v)}f", *v)) for k, v in d.items()]
    back = tf_bundle.read_table(tf_bundle.build_table(items))
    assert back == sorted(items)


def test_table_many_entries_span_restarts_and_blocks():
    items = [(f"var/{i:05d}".encode(), bytes([i % 251]) * 700) for i in range(900)]   # > 262144 B => several blocks
    back = tf_bundle.read_table(tf_bundle.build_table(items))
    assert back == items


def test_saver_state_file_and_max_to_keep(tmp_path):
    s = saver.Saver(max_to_keep=2)
    for step in (10, 20, 30):
        s.save({"w": np.full((2, 2), step, np.float32), "global_step": np.float32(step)}, str(tmp_path / "model.ckpt"), step)
    txt = open(tmp_path / "checkpoint").read()
    assert txt.splitlines()[0] == 'model_checkpoint_path: "model.ckpt-30"'
    assert re.findall(r'all_model_checkpoint_paths: "(.*)"', txt) == ["model.ckpt-20", "model.ckpt-30"]
    assert not os.path.exists(tmp_path / "model.ckpt-10.index")
    latest = saver.latest_checkpoint(str(tmp_path))
    assert latest.endswith("model.ckpt-30") and saver.Saver.restore(latest)["global_step"] == 30
    assert saver.latest_checkpoint(str(tmp_path / "nothing")) is None


def test_events_roundtrip_and_tensorboard_reads_it(tmp_path):
    w = events.SummaryWriter(str(tmp_path))
    for step in range(5):
        w.add_summary({"cost": 2.0 / (step + 1), "accuracy": 0.1 * step}, step)
    w.close()
    f = events.find_event_files(str(tmp_path))
    assert len(f) == 1 and os.path.basename(f[0]).startswith("events.out.tfevents.")
    recs = list(events.read_events(f[0]))
    assert recs[0][3] == "brain.Event:2" and [r[1] for r in recs[1:]] == list(range(5))
    assert recs[3][2]["cost"] == pytest.approx(2.0 / 3) and recs[3][2]["accuracy"] == pytest.approx(0.2)
    tb = pytest.importorskip("tensorboard.backend.event_processing.event_file_loader")
    seen = {}
    for ev in tb.EventFileLoader(f[0]).Load():
        for v in ev.summary.value:
            seen.setdefault(v.tag, []).append(ev.step)
    assert seen == {"cost": list(range(5)), "accuracy": list(range(5))}


def test_exporter_layout_and_signatures(tmp_path):
    e = exporter.Exporter(saver.Saver())
    with pytest.raises(RuntimeError):
        e.export(str(tmp_path), "0003", {})
    e.init(None, named_graph_signatures={"inputs": exporter.generic_signature({"x": "Placeholder:0"}),
                                         "outputs": exporter.generic_signature({"y": "test/add:0"})})
    out = e.export(str(tmp_path / "model"), "0003", {"test/weights": np.ones((1, 1), np.float32),
                                                     "test/bias": np.zeros(1, np.float32)})
    assert os.path.basename(out) == "00000003"                              # '0003' -> int 3 -> %08d
    assert sorted(os.listdir(out)) == ["assets", "export.data-00000-of-00001", "export.index", "export.meta", "signature.json"]
    assert exporter.parse_signatures_from_meta(os.path.join(out, "export.meta")) == {
        "inputs": {"x": "Placeholder:0"}, "outputs": {"y": "test/add:0"}}
    sigs, variables = exporter.load_export(out)
    assert sigs["outputs"] == {"y": "test/add:0"} and variables["test/weights"].shape == (1, 1)


def test_log_prefix_format():
    log.set_role("worker", 3)
    msg = log.format_message("hello", now=0)
    assert re.fullmatch(r" \[\d{4}-\d\d-\d\d \d\d:\d\d:\d\d\] \[worker:3\] hello", msg)


def test_proto_codec_roundtrip():
    msg = pb.f_varint(1, 300) + pb.f_bytes(2, "abc") + pb.f_fixed32(6, 0xDEADBEEF) + pb.f_msg(3, pb.f_varint(1, -1))
    d = pb.decode(msg)
    assert d[1] == [300] and d[2] == [b"abc"] and d[6] == [0xDEADBEEF]
    assert pb.decode(d[3][0])[1] == [(1 << 64) - 1]
    with pytest.raises(ValueError):
        pb.decode(b"\x0a\x05ab")


def test_native_bundle_writer_reader_match_python_twin(tmp_path, monkeypatch):
    """csrc/io/bundle.cpp against b200ps/io/tf_bundle.py: identical bytes for both files (300 entries span
    several restart groups; scalars, empty and 3-D tensors, every dtype), and each reader reads the other's files."""
    import numpy as np
    from b200ps.io import tf_bundle
    from b200ps.ops import native
    if not native.available():
        pytest.skip("libb200ps.so not built")
    rng = np.random.default_rng(0)
    tensors = {"global_step": np.float32(17.0), "empty": np.zeros((0, 4), np.float32),
               "cube": rng.standard_normal((3, 4, 5)).astype(np.float64),
               "ids": rng.integers(-2**40, 2**40, size=37), "flags": rng.integers(0, 2, size=9).astype(bool)}
    for i in range(300):
        tensors["layer_%03d/weights/Variable_%d" % (i // 7, i)] = rng.standard_normal((i % 5 + 1, 3)).astype(np.float32)
    for dt in ("int32", "uint8", "int16", "int8", "uint16", "float16", "uint32", "uint64"):
        tensors["dtypes/" + dt] = (rng.standard_normal(11) * 10).astype(dt)
    monkeypatch.setenv("B200PS_NATIVE_BUNDLE", "1")
    tf_bundle.write_bundle(str(tmp_path / "nat" / "model.ckpt-5"), tensors)
    monkeypatch.setenv("B200PS_NATIVE_BUNDLE", "0")
    tf_bundle.write_bundle(str(tmp_path / "py" / "model.ckpt-5"), tensors)
    for suffix in (".index", ".data-00000-of-00001"):
        a = open(str(tmp_path / "nat" / "model.ckpt-5") + suffix, "rb").read()
        b = open(str(tmp_path / "py" / "model.ckpt-5") + suffix, "rb").read()
        assert a == b, suffix
    assert not [f for f in (tmp_path / "nat").iterdir() if f.name.endswith(".tmp")]
    from_nat_by_py = tf_bundle.read_bundle(str(tmp_path / "nat" / "model.ckpt-5"))
    monkeypatch.setenv("B200PS_NATIVE_BUNDLE", "1")
    from_py_by_nat = tf_bundle.read_bundle(str(tmp_path / "py" / "model.ckpt-5"))
    assert set(from_nat_by_py) == set(from_py_by_nat) == set(tensors)
    for k, v in tensors.items():
        v = np.asarray(v)
        for got in (from_nat_by_py[k], from_py_by_nat[k]):
            assert got.dtype == v.dtype and got.shape == v.shape and np.array_equal(got, v), k
    # corruption is caught natively too: a flipped data byte, a flipped index byte
    p = str(tmp_path / "py" / "model.ckpt-5")
    raw = bytearray(open(p + ".data-00000-of-00001", "rb").read()); raw[100] ^= 0xFF
    open(p + ".data-00000-of-00001", "wb").write(bytes(raw))
    with pytest.raises(ValueError, match="checksum"):
        tf_bundle.read_bundle(p)
    idx = bytearray(open(p + ".index", "rb").read()); idx[40] ^= 0x01
    open(p + ".index", "wb").write(bytes(idx))
    with pytest.raises(ValueError):
        tf_bundle.read_bundle(p)


@pytest.mark.parametrize("use_native", ["1", "0"], ids=["native", "python"])
def test_bundle_readers_refuse_damaged_files(tmp_path, monkeypatch, use_native):
    """Truncation, flipped bytes and inserted garbage in either file of a checkpoint: both readers answer with a
    ValueError (or, when the damage is a no-op, the right tensors) — never a crash, never silently wrong values."""
    import numpy as np
    from b200ps.io import tf_bundle
    from b200ps.ops import native
    if use_native == "1" and not native.available():
        pytest.skip("libb200ps.so not built")
    monkeypatch.setenv("B200PS_NATIVE_BUNDLE", use_native)
    rng = np.random.default_rng(3)
    tensors = {"global_step": np.float32(17.0), "w/Variable": rng.standard_normal((20, 7)).astype(np.float32),
               "ids": rng.integers(-2**40, 2**40, size=37)}
    for i in range(40):
        tensors["layer_%03d/weights" % i] = rng.standard_normal((i % 5 + 1, 3)).astype(np.float32)
    p = str(tmp_path / "model.ckpt-1")
    tf_bundle.write_bundle(p, tensors)
    files = {s: open(p + s, "rb").read() for s in (".index", ".data-00000-of-00001")}
    refused = 0
    for it in range(240):
        cur = {k: bytearray(v) for k, v in files.items()}
        tgt = cur[".index" if it % 2 == 0 else ".data-00000-of-00001"]
        mode = (it // 2) % 3
        if mode == 0:
            del tgt[int(rng.integers(0, len(tgt))):]
        elif mode == 1:
            for _ in range(int(rng.integers(1, 4))):
                tgt[int(rng.integers(0, len(tgt)))] = int(rng.integers(0, 256))
        else:
            i = int(rng.integers(0, len(tgt)))
            tgt[i:i] = bytes(rng.integers(0, 256, size=int(rng.integers(1, 30)), dtype=np.uint8))
        for k, v in cur.items():
            open(p + k, "wb").write(bytes(v))
        try:
            got = tf_bundle.read_bundle(p)
        except ValueError:
            refused += 1
            continue
        assert set(got) == set(tensors) and all(np.array_equal(got[k], np.asarray(v)) for k, v in tensors.items())
    assert refused > 200


def test_native_events_writer_matches_python_twin(tmp_path, monkeypatch):
    """csrc/io/events.cpp vs b200ps/io/events.py: byte-identical scalar records for the same inputs."""
    from b200ps.io import events
    from b200ps.ops import native
    if not native.available():
        pytest.skip("libb200ps.so not built")
    paths = {}
    for mode in ("1", "0"):
        monkeypatch.setenv("B200PS_NATIVE_EVENTS", mode)
        w = events.SummaryWriter(str(tmp_path / mode), flush_secs=0.05)
        assert (w._native is not None) == (mode == "1")
        for step in range(0, 300, 3):
            w.add_summary({"cost": 1.0 / (step + 1), "accuracy": step / 300.0, "zero": 0.0}, step, wall_time=1000.5 + step)
        w.flush()
        assert len(list(events.read_events(w.path))) == 101          # visible after flush, before close
        w.close()
        paths[mode] = w.path
    a, b = open(paths["1"], "rb").read(), open(paths["0"], "rb").read()
    first = 16 + len(b"\x09" + b"\x00" * 8 + b"\x1a\x0d" + b"brain.Event:2")   # the file_version record (wall time differs)
    assert len(a) == len(b) and a[first:] == b[first:]
    got = list(events.read_events(paths["1"]))
    assert got[0][3] == "brain.Event:2" and got[1][1] == 0 and got[2][1] == 3 and abs(got[2][2]["cost"] - 0.25) < 1e-7


def _jpeg_cases():
    import io
    from PIL import Image
    rng = np.random.default_rng(0)
    out = []
    for (h, w), sub, q, restart in [((28, 28), 0, 90, 0), ((28, 28), 2, 75, 0), ((33, 47), 1, 85, 0), ((64, 48), 2, 95, 0),
                                    ((17, 23), 0, 50, 0), ((120, 97), 2, 30, 0), ((28, 28), None, 90, 0)]:
        yy, xx = np.mgrid[0:h, 0:w]
        img = np.stack([128 + 100 * np.sin(xx / 5.0 + yy / 7.0), 128 + 100 * np.cos(xx / 3.0), (xx * 255 / w + yy * 255 / h) / 2], 2)
        img = np.clip(img + rng.normal(0, 8, img.shape), 0, 255).astype(np.uint8)
        buf = io.BytesIO()
        if sub is None:
            Image.fromarray(img[:, :, 0]).save(buf, format="JPEG", quality=q)          # greyscale
        else:
            Image.fromarray(img).save(buf, format="JPEG", quality=q, subsampling=sub)
        data = buf.getvalue()
        ref = np.array(Image.open(io.BytesIO(data)).convert("RGB"))
        out.append((data, ref))
    return out


def test_jpeg_decoder_host_path_matches_pil(native_lib):
    """Native Huffman front end + the numpy restatement of the device stage (libjpeg's integer IDCT, fancy chroma
    upsampling, fixed-point colour conversion) against PIL: within +-1 LSB everywhere (in fact equal), for
    4:4:4 / 4:2:2 / 4:2:0 / greyscale files of odd sizes."""
    from b200ps.io.jpeg import decode_jpeg, decode_jpeg_numpy, jpeg_info
    for data, ref in _jpeg_cases():
        assert jpeg_info(data)[:2] == ref.shape[:2]
        got = decode_jpeg_numpy(data)
        assert got.shape == ref.shape and np.abs(got.astype(int) - ref.astype(int)).max() <= 1
        assert torch.equal(decode_jpeg(data, device="cpu"), torch.from_numpy(got))
    with pytest.raises(Exception):
        jpeg_info(b"not a jpeg at all")
    # a batch: the host stage of the files on a thread pool, same pixels as one by one
    from b200ps.io.jpeg import decode_jpeg_batch, parse_jpeg_batch
    same = [d for d, r in _jpeg_cases() if r.shape[:2] == (28, 28)] * 3
    got = decode_jpeg_batch(same, device="cpu")
    assert torch.equal(got, torch.stack([torch.from_numpy(decode_jpeg_numpy(d)) for d in same]))
    for (d1, c1), (d2, c2) in zip(parse_jpeg_batch(same, threads=4), parse_jpeg_batch(same, threads=1)):
        assert np.array_equal(c1, c2) and d1.total_coefs == d2.total_coefs


def test_jpeg_decoder_survives_damaged_files(native_lib):
    """Truncated files, flipped bytes and inserted garbage: the entropy decoder either refuses the file with an
    error or returns an image of the announced shape — it never reads outside its buffers (a crash here would take
    the test process down) and never allocates for a frame header that claims gigapixels."""
    from b200ps.io import jpeg
    from b200ps.ops import native
    rng = np.random.default_rng(1)
    cases = [d for d, _ in _jpeg_cases()]
    decoded = refused = 0
    for it in range(900):
        d = bytearray(cases[it % len(cases)])
        mode = it % 3
        if mode == 0:
            d = d[: rng.integers(0, len(d))]
        elif mode == 1:
            for _ in range(rng.integers(1, 6)):
                d[rng.integers(0, len(d))] = rng.integers(0, 256)
        else:
            i = int(rng.integers(0, len(d)))
            d[i:i] = bytes(rng.integers(0, 256, size=rng.integers(1, 40), dtype=np.uint8))
        try:
            out = jpeg.decode_jpeg_numpy(bytes(d))
            assert out.ndim == 3 and out.shape[2] == 3 and out.dtype == np.uint8
            decoded += 1
        except (ValueError, native.NativeError):
            refused += 1
    assert decoded > 100 and refused > 100
    big = bytearray(cases[0])
    sof = big.index(b"\xff\xc0")
    big[sof + 5:sof + 9] = b"\xff\xff\xff\xff"                # 65535 x 65535
    with pytest.raises(ValueError, match="exceeds the limit"):
        jpeg.parse_jpeg(bytes(big))


@pytest.mark.gpu
def test_jpeg_decoder_gpu_matches_pil(native_lib):
    """The CUDA stage (inverse DCT, chroma upsampling, YCbCr->RGB) against PIL, +-1 LSB (decode_jpeg(channels=3),
    input_pipeline.py:70); the fused bilinear resize against torch's on the decoded image."""
    from b200ps.io.jpeg import decode_jpeg
    for data, ref in _jpeg_cases():
        got = decode_jpeg(data, device="cuda")
        assert got.dtype == torch.uint8 and tuple(got.shape) == ref.shape
        assert (got.cpu().int() - torch.from_numpy(ref).int()).abs().max().item() <= 1
    data, ref = _jpeg_cases()[3]
    small = decode_jpeg(data, device="cuda", size=(28, 28))
    want = torch.nn.functional.interpolate(torch.from_numpy(ref).permute(2, 0, 1)[None].float(), size=(28, 28),
                                           mode="bilinear", align_corners=False)[0].permute(1, 2, 0)
    assert tuple(small.shape) == (28, 28, 3) and (small.cpu().float() - want).abs().max().item() <= 1.0
