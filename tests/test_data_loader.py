"""CPU: the training-batch loader keeps the reference's batch contract (dataset/data_loader.py:322-349)."""
import os

import numpy as np


def _make_data(tmp_path, n_spk=6, utts_per_spk=3, dim=7):
    from xvector_b200 import kaldi_io
    rng = np.random.default_rng(0)
    d = tmp_path / "data"
    d.mkdir()
    mats = {}
    with kaldi_io.open_or_fd("ark,scp:%s,%s" % (d / "feats.ark", d / "feats.scp"), "wb") as fd:
        for s in range(n_spk):
            for u in range(utts_per_spk):
                key = "spk%d-utt%d" % (s, u)
                rows = int(rng.integers(40, 120)) if not (s == 5) else 30        # speaker 5: only short utterances
                m = rng.standard_normal((rows, dim)).astype(np.float32) + 100 * s
                mats[key] = m
                kaldi_io.write_mat(fd, m, key)
    with open(d / "spk2utt", "w") as f:
        for s in range(n_spk):
            f.write("spk%d %s\n" % (s, " ".join("spk%d-utt%d" % (s, u) for u in range(utts_per_spk))))
    with open(d / "utt2num_frames", "w") as f:
        for k, m in mats.items():
            f.write("%s %d\n" % (k, m.shape[0]))
    with open(tmp_path / "spklist", "w") as f:
        for s in range(n_spk):
            f.write("spk%d %d\n" % (s, s))
    return str(d), str(tmp_path / "spklist"), mats


def test_random_segment_loader_batch_contract(tmp_path):
    from xvector_b200.data_loader import RandomSegmentLoader
    data, spklist, mats = _make_data(tmp_path)
    ld = RandomSegmentLoader(data, spklist, num_speakers=4, num_segments=2, min_len=32, max_len=38, seed=3)
    seen_len = set()
    for _ in range(6):
        x, y = ld.fetch()
        assert x.dtype == np.float32 and y.dtype == np.int32
        b, length, dim = x.shape
        assert b == 8 and dim == 7 and 32 <= length <= 38 and y.shape == (8,)
        seen_len.add(length)
        assert (y.reshape(4, 2)[:, 0] == y.reshape(4, 2)[:, 1]).all() and len(set(y[::2])) == 4   # distinct speakers, num_segments each
        assert 5 not in y                                      # its utterances are never longer than the batch length
        for j in range(b):
            # every segment is a contiguous slice of one utterance of that speaker
            cands = [m for k, m in mats.items() if k.startswith("spk%d-" % y[j])]
            hit = False
            for m in cands:
                for s in range(m.shape[0] - length + 1):
                    if np.array_equal(m[s:s + length], x[j]):
                        hit = True
                        break
                if hit:
                    break
            assert hit
    assert len(seen_len) > 1
    # dim slicing (model/trainer.py:626-627 does it on the trainer side; the loader can do it at the source)
    ld2 = RandomSegmentLoader(data, spklist, num_speakers=2, num_segments=1, min_len=32, max_len=32, dim=5, seed=1)
    x, _ = ld2.fetch()
    assert x.shape == (2, 32, 5)
    # without utt2num_frames the lengths come from the matrix headers
    os.remove(os.path.join(data, "utt2num_frames"))
    ld3 = RandomSegmentLoader(data, spklist, num_speakers=3, num_segments=1, min_len=35, max_len=35, seed=2)
    assert ld3.fetch()[0].shape == (3, 35, 7)


def test_sequential_segment_loader_contract(tmp_path):
    """KaldiDataSeqQueue / batch_sequence (dataset/data_loader.py:498-549): len // batch + 1 full batches in spk2utt order, the
    list wrapping around, one length per batch cut to the shortest utterance, segments from frame 0, DataOutOfRange at the end."""
    import pytest
    from xvector_b200.data_loader import DataOutOfRange, SequentialSegmentLoader
    data, spklist, mats = _make_data(tmp_path)
    keys = list(mats)                                           # spk2utt order: speaker by speaker
    ld = SequentialSegmentLoader(data, spklist, batch_size=4, min_len=35, max_len=35)
    assert ld.num_total_speakers == 6 and ld.num_batches == len(keys) // 4 + 1 and ld.batch_size == 4
    seen = []
    for i in range(ld.num_batches):
        x, y = ld.fetch()
        want = [keys[(4 * i + j) % len(keys)] for j in range(4)]
        length = min(35, min(mats[k].shape[0] for k in want))   # speaker 5's utterances have 30 frames
        assert x.shape == (4, length, 7) and y.dtype == np.int32
        for j, k in enumerate(want):
            assert np.array_equal(x[j], mats[k][:length]) and y[j] == int(k[3])
        seen += want
    assert set(seen) == set(keys)
    with pytest.raises(DataOutOfRange):
        ld.fetch()
    # fewer utterances than a batch: one batch of all of them (min(batch_size, len)), twice over the list (num_batches = 1)
    ld2 = SequentialSegmentLoader(data, spklist, batch_size=64, min_len=20, max_len=20, dim=5)
    x, y = ld2.fetch()
    assert x.shape == (len(keys), 20, 5) and ld2.num_batches == 1
    with pytest.raises(DataOutOfRange):
        ld2.fetch()


def test_reference_queue_names(tmp_path):
    """KaldiDataRandomQueue / KaldiDataSeqQueue by the reference's names and signatures (dataset/data_loader.py:369, :554):
    num_total_speakers before start(), set_batch / set_length, start / fetch / stop, DataOutOfRange at the end of a pass."""
    import pytest
    from xvector_b200.data_loader import DataOutOfRange, KaldiDataRandomQueue, KaldiDataSeqQueue
    data, spklist, mats = _make_data(tmp_path)
    q = KaldiDataRandomQueue(data, spklist, num_parallel=8, max_qsize=10, num_speakers=3, num_segments=2, min_len=32, max_len=36)
    assert q.num_total_speakers == 6
    q.set_batch(2, 2)
    q.set_length(33, 33)
    q.start()
    x, y = q.fetch()
    assert x.shape == (4, 33, 7) and y.shape == (4,) and y[0] == y[1] and y[2] == y[3] and y[0] != y[2]
    q.stop()
    s = KaldiDataSeqQueue(data, spklist, num_parallel=2, max_qsize=20, batch_size=5, min_len=35, max_len=35, shuffle=False)
    s.start()
    n = 0
    while True:
        try:
            x, y = s.fetch()
        except DataOutOfRange:
            break
        assert x.shape[0] == 5 and x.shape[2] == 7
        n += 1
    assert n == len(mats) // 5 + 1
    s.stop()
