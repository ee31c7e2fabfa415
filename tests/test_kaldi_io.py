"""Kaldi ark/scp I/O, CPU: the native library (lib/libxvkaldi.so through its C-ABI and the
reference-named Python mirror) and the numpy restatement against golden vectors produced by the reference's own
dataset/kaldi_io.py (tests/golden/make_kaldi_io_golden.py).  Bit-exact everywhere."""
import ctypes as C
import gzip
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="module")
def K():
    import xvector_b200.kaldi_io as kaldi_io
    return kaldi_io


@pytest.fixture(scope="module")
def gold():
    return dict(np.load(os.path.join(GOLDEN, "kaldi_io_golden.npz")))


def _same(a, b):
    a, b = np.asarray(a), np.asarray(b)
    return a.dtype == b.dtype and a.shape == b.shape and np.array_equal(a.view(np.uint8), b.view(np.uint8))


def test_library_exports_every_declared_symbol(K):
    header = open(os.path.join(ROOT, "include", "xvkaldi.h")).read()
    declared = set(re.findall(r"\b(xvk_[a-z0-9_]+)\s*\(", header))
    assert declared == set(K.PROTOTYPES), declared ^ set(K.PROTOTYPES)
    lib = K.load()
    for name in declared:
        assert hasattr(lib, name), name
    assert b"xvkaldi" in lib.xvk_version()


@pytest.mark.parametrize("ark,prefix", [("kaldi_io_mats.ark", "mats/"), ("kaldi_io_cm.ark", "cm/"), ("kaldi_io_text.ark", "text/")])
def test_read_mat_ark_matches_reference(K, gold, ark, prefix):
    got = list(K.read_mat_ark(os.path.join(GOLDEN, ark)))
    want = [(k[len(prefix):], v) for k, v in gold.items() if k.startswith(prefix)]
    assert [k for k, _ in got] == [k for k, _ in want]
    for (k, m), (_, ref) in zip(got, want):
        assert _same(m, ref), k


def test_read_vec_flt_ark_matches_reference(K, gold):
    got = list(K.read_vec_flt_ark("ark:" + os.path.join(GOLDEN, "kaldi_io_vecs.ark")))
    want = [(k[5:], v) for k, v in gold.items() if k.startswith("vecs/")]
    assert [k for k, _ in got] == [k for k, _ in want]
    for (k, v), (_, ref) in zip(got, want):
        assert _same(v, ref), k


def test_oracle_restatement_matches_reference(gold):
    from oracle import kaldi_io_ref as O
    for ark, prefix in (("kaldi_io_mats.ark", "mats/"), ("kaldi_io_cm.ark", "cm/"), ("kaldi_io_vecs.ark", "vecs/")):
        parsed = O.parse_ark(open(os.path.join(GOLDEN, ark), "rb").read())
        want = [(k[len(prefix):], v) for k, v in gold.items() if k.startswith(prefix)]
        assert [k for k, _ in parsed] == [k for k, _ in want]
        for (k, m), (_, ref) in zip(parsed, want):
            assert _same(np.ascontiguousarray(m), ref), k


def test_writers_reproduce_the_reference_bytes(K, gold, tmp_path):
    """write_mat / write_vec_flt: the bytes the reference's writers produced for the same arrays."""
    for ark, prefix, write in (("kaldi_io_mats.ark", "mats/", K.write_mat), ("kaldi_io_vecs.ark", "vecs/", K.write_vec_flt)):
        out = str(tmp_path / ark)
        with K.open_or_fd(out, "wb") as f:
            for k, v in gold.items():
                if k.startswith(prefix):
                    write(f, v, key=k[len(prefix):])
        assert open(out, "rb").read() == open(os.path.join(GOLDEN, ark), "rb").read()
    from oracle import kaldi_io_ref as O
    blob = b"".join(O.write_mat_bytes(v, k[5:]) for k, v in gold.items() if k.startswith("mats/"))
    assert blob == open(os.path.join(GOLDEN, "kaldi_io_mats.ark"), "rb").read()
    with pytest.raises(K.UnsupportedDataType):
        K.write_mat(str(tmp_path / "bad.ark"), np.zeros((2, 2), np.int32), key="x")


def test_submatrix_reads_match_reference(K, gold):
    for name, ref in gold.items():
        if not name.startswith("sub/"):
            continue
        _, key, start, length = name.split("/")
        ark = "kaldi_io_cm.ark" if key.startswith("cm") else "kaldi_io_mats.ark"
        fd = K.open_or_fd(os.path.join(GOLDEN, ark))
        k = K.read_key(fd)
        while k != key:
            k = K.read_key(fd)                 # an unread body is skipped
        _, kind, rows, cols = fd._head
        got = fd.read_rows(int(start), int(length), cols)
        fd._head = None
        assert _same(got, ref), name
        nxt = K.read_key(fd)                   # the stream is positioned on the next object
        assert nxt is None or isinstance(nxt, str)
        fd.close()
    fd = K.open_or_fd(os.path.join(GOLDEN, "kaldi_io_mats.ark"))
    K.read_key(fd)
    with pytest.raises(K.KaldiIOError, match="not enough for length"):
        fd.read_rows(0, 5, 30)                 # fm_0 has one row (dataset/kaldi_io.py:1046)
    fd.close()


def test_cm_decode_every_byte_value(K, tmp_path):
    """All 256 byte values in every column, extreme percentiles: native table == restated float32 evaluation."""
    from oracle import kaldi_io_ref as O
    import struct
    rng = np.random.default_rng(3)
    cols, rows = 7, 256
    heads = np.sort(rng.integers(0, 65536, (cols, 4)).astype(np.uint16), axis=1)
    heads[0] = (0, 1, 65534, 65535)
    data = np.stack([rng.permutation(256).astype(np.uint8) for _ in range(cols)])
    body = struct.pack("<ffii", -13.25, 77.125, rows, cols) + heads.tobytes() + data.tobytes()
    path = str(tmp_path / "cm.ark")
    open(path, "wb").write(b"k \0BCM " + body)
    (key, got), = list(K.read_mat_ark(path))
    ref, used = O.decode_cm(body)
    assert key == "k" and used == len(body) and _same(got, np.ascontiguousarray(ref))


def test_specifiers_pipes_gzip_offsets_and_scp(K, gold, tmp_path):
    mats = os.path.join(GOLDEN, "kaldi_io_mats.ark")
    want = [(k[5:], v) for k, v in gold.items() if k.startswith("mats/")]
    # input pipe and the ark: prefix with options
    got = list(K.read_mat_ark("ark,s,cs:cat %s |" % mats))
    assert [k for k, _ in got] == [k for k, _ in want] and all(_same(a, b) for (_, a), (_, b) in zip(got, want))
    # gzip in, gzip out
    gz = str(tmp_path / "m.ark.gz")
    with K.open_or_fd(gz, "wb") as f:
        for k, v in want:
            K.write_mat(f, v, key=k)
    assert gzip.open(gz, "rb").read() == open(mats, "rb").read()
    assert all(_same(a, b) for (_, a), (_, b) in zip(K.read_mat_ark(gz), want))
    # output pipe
    piped = str(tmp_path / "piped.ark")
    with K.open_or_fd("ark:| cat > %s" % piped, "wb") as f:
        for k, v in want:
            K.write_mat(f, v, key=k)
    assert open(piped, "rb").read() == open(mats, "rb").read()
    # ark,scp writer -> scp random access (single, threaded, header-only)
    ark, scp = str(tmp_path / "x.ark"), str(tmp_path / "x.scp")
    with K.open_or_fd("ark,scp:%s,%s" % (ark, scp), "wb") as f:
        for k, v in want:
            K.write_mat(f, v, key=k)
    assert open(ark, "rb").read() == open(mats, "rb").read()
    got = list(K.read_mat_scp(scp))
    assert [k for k, _ in got] == [k for k, _ in want] and all(_same(a, b) for (_, a), (_, b) in zip(got, want))
    table = K.ScpTable("scp:" + scp)
    f32 = [i for i, (k, v) in enumerate(want) if v.dtype == np.float32 and v.shape[1] == 30]
    rows, cols = table.dims(f32, threads=3)
    assert list(rows) == [want[i][1].shape[0] for i in f32] and set(cols) == {30}
    packed, lens = table.read_many(f32, threads=4)
    assert _same(packed, np.concatenate([want[i][1] for i in f32]))
    one = K.read_mat(table.rxfile(2))          # "path:offset"
    assert _same(one, want[2][1])
    table.close()


def test_batch_reader_refuses_a_dimension_change(K):
    br = K.BatchReader(os.path.join(GOLDEN, "kaldi_io_cm.ark"), 30, max_rows=400)
    with pytest.raises(K.KaldiIOError, match="feature dimension 24, expected 30"):
        list(br)
    with pytest.raises(K.KaldiIOError, match="expected a feature matrix, found a vector"):
        list(K.BatchReader(os.path.join(GOLDEN, "kaldi_io_vecs.ark"), 512))


@pytest.mark.parametrize("threads", [1, 4])
def test_batch_reader_shapes(K, gold, tmp_path, threads):
    mats = [(k[5:], v) for k, v in gold.items() if k.startswith("mats/") and v.shape[1] == 30 and v.dtype == np.float32]
    path = str(tmp_path / "b.ark")
    with K.open_or_fd(path, "wb") as f:
        for rep in range(3):
            for k, v in mats:
                K.write_mat(f, v, key="%s_%d" % (k, rep))
    allk, allf, alll = [], [], []
    for keys, feats, lens in K.BatchReader(path, 30, max_rows=230, max_utts=3, threads=threads):
        assert len(keys) <= 3 and feats.shape[0] == lens.sum() <= 230
        allk += keys
        allf.append(feats.copy())
        alll += list(lens)
    assert allk == ["%s_%d" % (k, rep) for rep in range(3) for k, _ in mats]
    assert _same(np.concatenate(allf), np.concatenate([v for _ in range(3) for _, v in mats]))
    assert alll == [v.shape[0] for _ in range(3) for _, v in mats]
    with pytest.raises(K.KaldiIOError, match="does not fit an empty batch"):
        list(K.BatchReader(path, 30, max_rows=100, threads=threads))
    # compressed and text objects through the same batch path
    mixed = str(tmp_path / "mixed.ark")
    from oracle import kaldi_io_ref as O
    with open(mixed, "wb") as f:
        for i, (k, v) in enumerate(mats):
            f.write(O.write_cm_bytes(v, k) if i % 2 == 0 and v.shape[0] > 1 else O.write_mat_bytes(v, k))
    want = dict(O.parse_ark(open(mixed, "rb").read()))
    got = {}
    for keys, feats, lens in K.BatchReader(mixed, 30, max_rows=300, threads=threads):
        offs = np.concatenate([[0], np.cumsum(lens)])
        for i, k in enumerate(keys):
            got[k] = feats[offs[i]:offs[i + 1]].copy()
    assert list(got) == list(want) and all(_same(got[k], np.ascontiguousarray(want[k])) for k in want)


def test_error_behaviour(K, tmp_path):
    bad = str(tmp_path / "bad.ark")
    open(bad, "wb").write(b"k \0BXM \4\1\0\0\0\4\1\0\0\0")
    with pytest.raises(K.UnknownMatrixHeader, match="The header contained 'XM '"):
        list(K.read_mat_ark(bad))
    open(bad, "wb").write(b"k \0BCM2 ")
    with pytest.raises(K.KaldiIOError, match="CM2, CM3 are not supported"):
        list(K.read_mat_ark(bad))
    open(bad, "wb").write(b"k \0BFM \4\5\0\0\0\4\3\0\0\0abc")
    with pytest.raises(K.KaldiIOError, match="unexpected end of stream"):
        list(K.read_mat_ark(bad))
    with pytest.raises(K.KaldiIOError, match="cannot open"):
        K.open_or_fd(str(tmp_path / "missing.ark"))
    with pytest.raises(K.SubprocessFailed):
        list(K.read_mat_ark("ark:exit 3 |"))
    assert list(K.read_mat_ark("ark:true |")) == []
    with pytest.raises(ValueError):
        K.open_or_fd(bad, "x")


def test_batch_reader_soft_batch_cut(tmp_path):
    """batch_rows cuts the batches below the buffer capacity; an utterance longer than a batch still gets through (its batch is
    filled to the capacity), and the keys of every batch come back without the padding of the key buffer."""
    from xvector_b200 import kaldi_io as K
    rng = np.random.default_rng(0)
    mats = [rng.standard_normal((n, 5)).astype(np.float32) for n in (10, 300, 20, 7, 120, 500, 3)]
    path = str(tmp_path / "soft.ark")
    with K.open_or_fd(path, "wb") as f:
        for i, m in enumerate(mats):
            K.write_mat(f, m, key="u%d" % i)
    batches = []
    for keys, feats, lens in K.BatchReader(path, 5, max_rows=600, max_utts=16, threads=2, batch_rows=128):
        batches.append(keys)
        assert len(keys) == len(lens) and feats.shape[0] == int(lens.sum())
        off = 0
        for k, n in zip(keys, lens):
            assert np.array_equal(feats[off:off + n], mats[int(k[1:])])
            off += n
    assert batches == [["u0"], ["u1", "u2", "u3", "u4"], ["u5", "u6"]]
    # without the soft cut: capacity only
    assert [k for k, _, _ in K.BatchReader(path, 5, max_rows=600, max_utts=16, threads=2)] == [["u0", "u1", "u2", "u3", "u4"], ["u5", "u6"]]
    # an utterance longer than the capacity is an error, not a truncation
    with pytest.raises(K.KaldiIOError):
        list(K.BatchReader(path, 5, max_rows=400, max_utts=16, threads=2, batch_rows=128))


def test_worker_pool_concurrent_readers_and_fork(tmp_path):
    """The native reader's parked worker threads: two Python threads reading at once (one gets the pool, the other starts its
    own threads), and a forked child that reads after the parent has used the pool (the parent's threads do not exist there)."""
    import signal
    import threading
    from xvector_b200 import kaldi_io as K
    rng = np.random.default_rng(4)
    mats = [rng.standard_normal((int(n), 6)).astype(np.float32) for n in rng.integers(20, 200, 60)]
    path = str(tmp_path / "pool.ark")
    with K.open_or_fd(path, "wb") as f:
        for i, m in enumerate(mats):
            K.write_mat(f, m, key="u%02d" % i)

    def read_all():
        total = 0
        for keys, feats, lens in K.BatchReader(path, 6, max_rows=1000, max_utts=16, threads=4):
            off = 0
            for k, n in zip(keys, lens):
                assert np.array_equal(feats[off:off + n], mats[int(k[1:])])
                off += n
            total += len(keys)
        assert total == len(mats)

    read_all()                                             # the pool exists from here on
    errors = []

    def guarded():
        try:
            for _ in range(5):
                read_all()
        except BaseException as e:                         # noqa: BLE001 -- reported in the main thread
            errors.append(e)

    threads = [threading.Thread(target=guarded) for _ in range(3)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(60)
    assert not errors and not any(t.is_alive() for t in threads)
    pid = os.fork()
    if pid == 0:                                           # child: must not wait for threads that were never copied
        signal.alarm(30)
        try:
            read_all()
            os._exit(0)
        except BaseException:                              # noqa: BLE001
            os._exit(1)
    _, status = os.waitpid(pid, 0)
    assert os.WIFEXITED(status) and os.WEXITSTATUS(status) == 0, status
    read_all()                                             # and the parent's pool is still fine


def test_scp_read_segments(tmp_path):
    """xvk_scp_read_segments_f32 (FeatureReader.read_segment over an scp, dataset/kaldi_io.py:1034-1053): every segment is
    the slice of its matrix, for float and compressed matrices, mapped and unmapped files, and a segment that does not fit
    is an error that names the key."""
    from oracle import kaldi_io_ref
    from xvector_b200 import kaldi_io as K
    rng = np.random.default_rng(6)
    mats = [rng.standard_normal((int(n), 5)).astype(np.float32) for n in rng.integers(30, 90, 12)]
    ark, scp = str(tmp_path / "seg.ark"), str(tmp_path / "seg.scp")
    with K.open_or_fd("ark,scp:%s,%s" % (ark, scp), "wb") as f:
        for i, m in enumerate(mats):
            K.write_mat(f, m, key="u%02d" % i)
    table = K.ScpTable(scp)
    idx = rng.integers(0, len(mats), 40)
    length = 17
    starts = np.array([rng.integers(0, mats[i].shape[0] - length + 1) for i in idx], np.int32)
    for threads in (1, 4):
        got = table.read_segments(idx, starts, length, threads=threads)
        assert got.shape == (40, length, 5) and got.dtype == np.float32
        for k, (i, s) in enumerate(zip(idx, starts)):
            assert np.array_equal(got[k], mats[i][s:s + length])
    with pytest.raises(K.KaldiIOError, match="not enough"):
        table.read_segments([0], [mats[0].shape[0] - length + 1], length)
    table.close()
    # compressed matrices: the rows of a segment decode to what the whole-matrix decode gives
    cm_ark, cm_scp = str(tmp_path / "cm.ark"), str(tmp_path / "cm.scp")
    with open(cm_ark, "wb") as f, open(cm_scp, "w") as g:
        for i, m in enumerate(mats[:4]):
            rec = kaldi_io_ref.write_cm_bytes(m, "c%d" % i)
            g.write("c%d %s:%d\n" % (i, cm_ark, f.tell() + len("c%d " % i)))
            f.write(rec)
    t2 = K.ScpTable(cm_scp)
    whole, lens = t2.read_many(np.arange(4))
    off = np.concatenate([[0], np.cumsum(lens)])
    seg = t2.read_segments([2, 0, 3], [5, 0, 11], 9)
    for k, (i, s) in enumerate(((2, 5), (0, 0), (3, 11))):
        assert np.array_equal(seg[k], whole[off[i] + s:off[i] + s + 9])
    t2.close()


def test_feature_reader_interface(tmp_path):
    """FeatureReader (dataset/kaldi_io.py:40-150): get_dim, read / read_segment by "utt path:offset" strings with the
    reference's length / shuffle / start rules."""
    from xvector_b200 import kaldi_io as K
    rng = np.random.default_rng(12)
    data = tmp_path / "data"
    data.mkdir()
    mats = {}
    with K.open_or_fd("ark,scp:%s,%s" % (data / "feats.ark", data / "feats.scp"), "wb") as f:
        for i in range(5):
            m = rng.standard_normal((int(rng.integers(30, 60)), 9)).astype(np.float32)
            mats["utt%d" % i] = m
            K.write_mat(f, m, key="utt%d" % i)
    (data / "utt2num_frames").write_text("".join("%s %d\n" % (k, m.shape[0]) for k, m in mats.items()))
    lines = [l.strip() for l in open(data / "feats.scp")]
    fr = K.FeatureReader(str(data))
    assert fr.get_dim() == 9 and fr.dim == 9 and fr.utt2num_frames["utt3"] == mats["utt3"].shape[0]
    for line in lines:
        utt = line.split(" ")[0]
        m, start = fr.read(line)
        assert start is None and np.array_equal(m, mats[utt])
        seg, start = fr.read_segment(line, 20)                       # no shuffle: from frame 0
        assert start == 0 and np.array_equal(seg, mats[utt][:20])
        seg, start = fr.read_segment(line, 20, shuffle=True)
        assert 0 <= start <= mats[utt].shape[0] - 20 and np.array_equal(seg, mats[utt][start:start + 20])
        seg, start = fr.read_segment(line, 10, start=7)
        assert start == 7 and np.array_equal(seg, mats[utt][7:17])
        whole, start = fr.read_segment(line, 1000)                   # longer than the utterance: cut to it
        assert start == 0 and np.array_equal(whole, mats[utt])
        part, start = fr.read(line, length=15, shuffle=True)
        assert np.array_equal(part, mats[utt][start:start + 15])
    with pytest.raises(AssertionError):
        fr.read_segment(lines[0], 10, shuffle=True, start=3)
    with pytest.raises(IOError):
        fr.read("nokey /nonexistent:0")
    fr.close()
