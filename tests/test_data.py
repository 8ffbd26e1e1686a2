"""Data layer (P18-P24): parser, batch formatter, sharding, loaders, queues — CPU only."""
import threading
import time

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from b200ps.data import libsvm, provider, synthetic
from b200ps.data import queue as Q


def test_parse_line_space_and_tab():
    s = libsvm.Sample().parse_line_libsvm("1 3:0.5\t7:2 11:1e-3\n")
    assert s.label == 1 and s.indices == [3, 7, 11] and s.values == [0.5, 2.0, 1e-3]
    s2 = libsvm.Sample().parse_line_libsvm("0")
    assert s2.label == 0 and s2.indices == []


def test_format_samples_sparse_matches_reference_layout_and_csr():
    lines = ["1 5:1.5 9:2", "0 2:0.25", "1 9:1 5:3 100:0.5"]
    samples = [libsvm.Sample().parse_line_libsvm(l) for l in lines]
    labels, fids, fvals, sp_indices, size = libsvm.Sample.format_samples_sparse(samples)
    assert labels.shape == (3, 1) and size == 3
    assert fids == [5, 9, 2, 9, 5, 100] and fvals == [1.5, 2.0, 0.25, 1.0, 3.0, 0.5]
    assert sp_indices == [[0, 5], [0, 9], [1, 2], [2, 9], [2, 5], [2, 100]]     # [row, fid] (lr2.py:82)
    rp, fid, fv, y = libsvm.Sample.format_samples_csr(samples)
    rp2, fid2, fv2, y2 = libsvm.csr_from_sparse_triple(sp_indices, fids, fvals, labels)
    assert rp.tolist() == [0, 2, 3, 6] == rp2.tolist()
    assert np.array_equal(fid, fid2) and np.allclose(fv, fv2) and np.array_equal(y, y2)


_line = st.tuples(st.integers(0, 1),
                  st.lists(st.tuples(st.integers(0, 10 ** 9), st.floats(-100, 100, allow_nan=False, width=32)),
                           max_size=12))


@settings(max_examples=60, deadline=None)
@given(st.lists(_line, min_size=1, max_size=20), st.booleans())
def test_native_parser_equals_python_parser(rows, tabs):
    sep = "\t" if tabs else " "
    text = "".join(str(lab) + "".join(f"{sep}{i}:{v!r}" for i, v in feats) + "\n" for lab, feats in rows)
    rp, fid, fv, y = libsvm.parse_buffer(text.encode())
    samples = [libsvm.Sample().parse_line_libsvm(l) for l in text.splitlines() if len(l) >= 2]
    rp2, fid2, fv2, y2 = libsvm.Sample.format_samples_csr(samples)
    assert np.array_equal(rp, rp2) and np.array_equal(fid, fid2) and np.array_equal(y, y2)
    assert np.allclose(fv, fv2, rtol=1e-6, atol=1e-30)


def _float32_of_decimal(text: str) -> np.float32:
    """The correctly rounded (nearest, ties to even) float32 of a decimal string, by exact rational arithmetic."""
    from fractions import Fraction
    x = Fraction(text)
    f0 = np.float32(float(text))
    with np.errstate(over="ignore"):
        cands = {f0, np.nextafter(f0, np.float32(np.inf)), np.nextafter(f0, np.float32(-np.inf))}
    best = None
    for c in cands:
        if not np.isfinite(c):
            continue
        err = abs(Fraction(float(c)) - x)
        even = (int(np.float32(c).view(np.uint32)) & 1) == 0
        key = (err, 0 if even else 1)
        if best is None or key < best[0]:
            best = (key, c)
    return np.float32(best[1])


def test_native_parser_decimals_are_correctly_rounded():
    """csrc/data/fast_parse.h converts plain decimals itself (m / 10^k in double, then to float) and must hand
    anything it cannot convert exactly to strtof: every value is compared BIT FOR BIT with the correctly rounded
    float32 of the decimal — random short and long digit strings, float rounding boundaries and their neighbours."""
    rng = np.random.RandomState(7)
    texts = ["0", "-0", "0.0", "5.", ".5", "000.250", "16777217", "16777217.0000001", "16777216.9999999",
             "0.1", "0.2", "0.3", "1.00000005960464", "1.0000000596046448", "33554434", "33554435", "8388608.5",
             "8388609.5", "0.000000000000000000001", "123456789012345", "1234567890123456", "340282346638528859811704183484516925440",
             "0.00000000000000000000000000000000000001", "3.4028234e38", "1e-40", "1.17549435e-38", "4.5"]
    for _ in range(6000):
        nd, nf = rng.randint(0, 9), rng.randint(0, 12)
        if nd + nf == 0:
            nd = 1
        ip = "".join(str(d) for d in rng.randint(0, 10, size=nd))
        fp = "".join(str(d) for d in rng.randint(0, 10, size=nf))
        texts.append(("-" if rng.rand() < 0.3 else "") + ip + ("." + fp if nf or rng.rand() < 0.1 else ""))
    for _ in range(2000):                       # float32 halfway points (exactly, and one decimal digit to either side)
        f = np.float32(rng.rand() * 10 ** rng.randint(-3, 6))
        from fractions import Fraction
        mid = (Fraction(float(f)) + Fraction(float(np.nextafter(f, np.float32(np.inf))))) / 2
        digits = 12
        scaled = mid * 10 ** digits
        base = scaled.numerator // scaled.denominator
        for delta in (-1, 0, 1):
            v = base + delta
            texts.append("%d.%0*d" % (v // 10 ** digits, digits, v % 10 ** digits))
    texts = [t for t in texts if t not in ("", "-")]
    buf = "".join("1 %d:%s\n" % (i, t) for i, t in enumerate(texts)).encode()
    rp, fid, fv, y = libsvm.parse_buffer(buf)
    assert len(fv) == len(texts) and fid.tolist() == list(range(len(texts)))
    want = np.asarray([_float32_of_decimal(t) for t in texts], dtype=np.float32)
    bad = np.nonzero(fv.view(np.uint32) != want.view(np.uint32))[0]
    assert bad.size == 0, [(texts[i], float(fv[i]), float(want[i])) for i in bad[:5]]


def test_native_parser_reports_bad_line():
    with pytest.raises(ValueError, match="line 2"):
        libsvm.parse_buffer(b"1 3:1\n0 oops\n")


def test_file_sharding_rules():
    files = ",".join(f"f{i}" for i in range(7)) + ","
    gl = provider.DataProvider.GetFileList
    assert gl(files, 3, 0) == ["f0", "f3", "f6"] and gl(files, 3, 1) == ["f1", "f4"] and gl(files, 3, 2) == ["f2", "f5"]
    dp = provider.DataProvider(1, 0, 3, "all")
    threads = dp.InitThreads([f"f{i}" for i in range(7)], 7, 3)
    assert [t.files for t in threads] == [["f0", "f3", "f6"], ["f1", "f4"], ["f2", "f5"]]   # files[tid::thread_num]


def test_provider_all_mode_and_sampling(tmp_path):
    tr = synthetic.write_libsvm_files(str(tmp_path / "train"), 4, 250, features=300)
    te = synthetic.write_libsvm_files(str(tmp_path / "test"), 2, 100, features=300)
    dp = provider.DataProvider(2, 1, 2, "all", ",".join(tr), ",".join(te), batch_size=64, seed=3)
    with pytest.raises(RuntimeError):
        dp.LoadData()                      # the stale reference test calls this without init()
    dp.init(); dp.LoadData()
    assert len(dp.train_samples) == 500 and len(dp.test_samples) == 100      # files[1::2]
    sizes = [len(b) for b in dp.NextBatch("train")]
    assert sizes == [64] * 7 + [52]
    before = [s.indices for s in dp.train_samples[:20]]
    dp.Shuffle()
    assert sorted(map(tuple, before)) != [] and len(dp.train_samples) == 500
    labels, fids, fvals, sp, n = dp.GetTestSamplesSampled(0.1, 1000)
    assert n == 10 and labels.shape == (10, 1)
    half = provider.DataProvider(1, 0, 2, "all", ",".join(tr), ",".join(te), train_sampling_rate=0.5, seed=7)
    half.init(); half.LoadData()
    assert 380 < len(half.train_samples) < 620 and len(half.test_samples) == 200


def test_provider_queue_mode(tmp_path):
    tr = synthetic.write_libsvm_files(str(tmp_path / "train"), 2, 120, features=300)
    te = synthetic.write_libsvm_files(str(tmp_path / "test"), 1, 80, features=300, seed=9)
    dp = provider.DataProvider(1, 0, 2, "queue", ",".join(tr), ",".join(te), batch_size=50,
                               train_queue_capacity=200, test_queue_capacity=200, seed=1)
    dp.init(); dp.LoadData()
    for _ in range(6):                      # the loaders loop over the files forever
        batches = list(dp.NextBatch("train"))
        assert len(batches) == 1 and len(batches[0]) == 50          # one batch per "epoch"
    test_b = dp.NextBatchQueue("test")
    train_ids = {tuple(s.indices) for f in tr for s in [libsvm.Sample().parse_line_libsvm(l) for l in open(f)]}
    assert all(tuple(s.indices) not in train_ids for s in test_b)   # is_train honoured: test queue gets test rows
    t0 = time.time(); dp.close(); assert time.time() - t0 < 10


def test_fifo_queue_semantics():
    q = Q.FIFOQueue(3)
    q.enqueue_many([1, 2, 3])
    with pytest.raises(Q.DeadlineExceededError):
        q.enqueue(4, timeout=0.05)
    assert q.dequeue() == 1 and q.size() == 2
    assert q.dequeue_many(2) == [2, 3]
    with pytest.raises(Q.DeadlineExceededError):
        q.dequeue(timeout=0.05)
    t = threading.Thread(target=lambda: (time.sleep(0.05), q.enqueue(9)))
    t.start()
    assert q.dequeue(timeout=2) == 9
    t.join()
    q.enqueue(5)
    q.close()
    assert q.dequeue() == 5                     # closed queues still drain
    with pytest.raises(Q.QueueClosedError):
        q.dequeue()
    with pytest.raises(Q.QueueClosedError):
        q.enqueue(1)


def test_batcher_and_slice_producer_pipeline():
    data = list(range(10)); labels = [i % 3 for i in range(10)]
    src = Q.slice_input_producer([data, labels], shuffle=False, capacity=4, num_epochs=2)
    b = Q.Batcher(src, 5, capacity=5)
    coord = Q.Coordinator()
    threads = Q.start_queue_runners(coord)
    got = [b.next(timeout=5) for _ in range(4)]
    assert [x for batch in got for x, _ in batch] == data + data            # order kept, two epochs
    with pytest.raises(Q.QueueClosedError):
        b.next(timeout=5)                                                     # end of input propagates
    coord.request_stop(); coord.join(threads)


def test_coordinator_surfaces_producer_errors():
    q = Q.FIFOQueue(2)
    def boom():
        raise ValueError("bad producer")
    coord = Q.Coordinator()
    threads = Q.QueueRunner(q, [boom]).create_threads(coord)
    coord.wait_for_stop(2)
    with pytest.raises(ValueError, match="bad producer"):
        coord.join(threads)


def test_synthetic_mnist_next_batch_epochs():
    ds = synthetic.SyntheticMNIST(250, seed=1)
    x, y = ds.next_batch(100)
    assert x.shape == (100, 784) and y.shape == (100,) and 0 <= float(x.min()) and float(x.max()) <= 1
    ds.next_batch(100)
    assert ds.epochs_completed == 0
    ds.next_batch(100)                          # wraps: reshuffle + new epoch
    assert ds.epochs_completed == 1
    oh = synthetic.SyntheticMNIST(50, one_hot=True)
    assert oh.next_batch(10)[1].shape == (10, 10) and oh.labels.shape == (50, 10)
    sets = synthetic.read_data_sets("/nonexistent")
    assert sets.train.num_examples == 55000 and sets.test.num_examples == 10000


def test_gfile_hdfs_mapping(tmp_path, monkeypatch):
    (tmp_path / "user").mkdir(); (tmp_path / "user" / "part-0").write_text("1 1:1\n0 2:1\n")
    with pytest.raises(IOError):
        libsvm.GFile("hdfs://localhost:9000/user/part-0")
    monkeypatch.setenv("B200PS_HDFS_ROOT", str(tmp_path))
    assert libsvm.count_lines("hdfs://localhost:9000/user/part-0") == 2
    assert libsvm.list_files(str(tmp_path / "user")) == [str(tmp_path / "user" / "part-0")]


def test_native_loader_packed_batches_match_python_parser(tmp_path):
    """csrc/data/loader.cpp: every kept line lands in exactly one packed batch with the values the Python
    parser gives; thread sharding files[t::threads]; incomplete final batches are dropped and counted."""
    import numpy as np
    from b200ps.data.libsvm import Sample
    from b200ps.data.native_loader import NativeLibsvmLoader, PackedLayout
    from b200ps.data.synthetic import write_libsvm_files
    files = write_libsvm_files(str(tmp_path / "train"), 3, 130, nnz_mean=12, spread=4, features=5000, seed=3)
    # a few awkward lines: tabs, CRLF, exponent notation, a malformed one, a too-short one, no trailing newline
    with open(files[0], "a") as f:
        f.write("1\t7:0.5\t9:1e-3\r\n0 3:2 4:.25\nbroken line\n\n1 11:3.125")
    want = {}
    for path in files:
        for line in open(path):
            if len(line.strip()) < 2 or line.startswith("broken"):
                continue
            smp = Sample().parse_line_libsvm(line)
            want[(smp.label, tuple(smp.indices))] = np.asarray(smp.values, dtype=np.float32)
    B = 50
    loader = NativeLibsvmLoader(files, batch_size=B, threads=2, max_nnz_per_row=40, ring=4, epochs=1, pinned=False)
    got, seqs, threads_seen = 0, [], set()
    for batch in loader:
        rp, fid, fval, y = (a.numpy() for a in batch.arrays())
        assert batch.rows == B and rp[0] == 0 and rp[-1] == batch.nnz == fid.size
        seqs.append(batch.seq); threads_seen.add(batch.thread)
        for i in range(B):
            key = (int(y[i]), tuple(int(v) for v in fid[rp[i]:rp[i + 1]]))
            np.testing.assert_allclose(fval[rp[i]:rp[i + 1]], want[key], rtol=1e-6)
            got += 1
    st = loader.stats()
    assert st["malformed"] == 1 and st["kept"] == len(want) == 3 * 130 + 3
    assert got == st["batches"] * B and got + st["dropped_incomplete_batch_rows"] == st["kept"]
    assert sorted(seqs) == list(range(st["batches"])) and threads_seen == {0, 1}
    loader.close()
    # sampling: roughly half of the lines survive; epochs=2 doubles the data; the nnz cap drops rows
    loader = NativeLibsvmLoader(files, batch_size=10, threads=3, epochs=2, sampling_rate=0.5, seed=7, pinned=False)
    n = sum(b.rows for b in loader)
    st = loader.stats()
    assert st["lines"] == 2 * (3 * 130 + 4) and abs(st["kept"] / st["lines"] - 0.5) < 0.08 and n <= st["kept"]
    loader.close()
    loader = NativeLibsvmLoader(files[:1], batch_size=10, threads=1, cap_nnz=60, pinned=False)
    assert sum(1 for _ in loader) == 0               # 10 rows of ~12 non-zeros never fit 60: nothing is published,
    st = loader.stats()                              # the rows are counted, and the loader still terminates
    assert st["dropped_nnz_cap"] > 0 and st["kept"] + st["dropped_nnz_cap"] + st["malformed"] == st["lines"]
    loader.close()
    # pack(): the same layout from arrays (what the benchmark and the tests use to feed HostFeeder)
    lay = PackedLayout.make(4, 16)
    buf = lay.pack(np.array([0, 2, 2, 5, 6]), np.arange(6) * 10, np.arange(6) / 4, np.array([1, 0, 0, 1]), pin=False)
    rp, fid, fval, y = lay.views(buf)
    assert rp.tolist() == [0, 2, 2, 5, 6] and fid[:6].tolist() == [0, 10, 20, 30, 40, 50] and y.tolist() == [1, 0, 0, 1]
    assert lay.off_fid % 16 == 0 and lay.off_fval % 16 == 0 and lay.off_labels % 16 == 0 and lay.bytes % 16 == 0


_float_text = st.one_of(
    st.floats(min_value=-1e6, max_value=1e6, allow_nan=False, width=32).map(lambda v: repr(float(v))),
    st.floats(min_value=-1e3, max_value=1e3, allow_nan=False).map(lambda v: "%.9f" % v),
    st.floats(min_value=1e-30, max_value=1e30, allow_nan=False).map(lambda v: "%.6e" % v),
    st.integers(min_value=-10**9, max_value=10**9).map(str),
    st.sampled_from(["0", "-0", "+1.5", ".5", "5.", "1e3", "1E-3", "-2.5e+2", "000.250", "123456789012345678.5"]))


@settings(max_examples=60, deadline=None)
@given(st.lists(st.tuples(st.integers(min_value=-1, max_value=1),
                          st.lists(st.tuples(st.integers(min_value=0, max_value=2**40), _float_text), min_size=1, max_size=6),
                          st.sampled_from([" ", "\t", "  ", " \t"]), st.sampled_from(["\n", "\r\n"])),
                min_size=1, max_size=12))
def test_native_loader_number_parsing_fuzz(tmp_path_factory, rows):
    """The loader's in-place number parser (fast decimal path + strtof fallback) against Python's int()/float()."""
    import numpy as np
    from b200ps.data.native_loader import NativeLibsvmLoader
    d = tmp_path_factory.mktemp("fuzz")
    text = "".join(str(lab) + "".join(sep + "%d:%s" % (i, v) for i, v in feats) + eol for lab, feats, sep, eol in rows)
    (d / "part-0").write_text(text + "0 1:1\n")                    # a short-line-proof terminator row
    loader = NativeLibsvmLoader([str(d / "part-0")], batch_size=len(rows) + 1, threads=1, max_nnz_per_row=8, pinned=False)
    batch = loader.next()
    assert batch is not None and batch.rows == len(rows) + 1 and loader.stats()["malformed"] == 0
    rp, fid, fval, y = (a.numpy() for a in batch.arrays())
    for r, (lab, feats, _, _) in enumerate(rows):
        assert y[r] == lab and fid[rp[r]:rp[r + 1]].tolist() == [i for i, _ in feats]
        want = np.asarray([float(v) for _, v in feats], dtype=np.float64).astype(np.float32)
        np.testing.assert_allclose(fval[rp[r]:rp[r + 1]], want, rtol=2e-7, atol=0)
    batch.release(); loader.close()
