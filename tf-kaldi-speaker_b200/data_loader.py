"""Training batches from a Kaldi data directory: the sampling rule of the reference's KaldiDataRandomQueue
(dataset/data_loader.py:249-360, :369-495) restated over the native scp reader (csrc/xv_kaldi_io.cpp).

The reference fills a multiprocessing queue from `num_parallel_datasets` producer processes because its Python reader is
slow (13 M frames/s here); the native reader decodes a 64 x 400-frame batch in well under a millisecond per thread, so
one in-process loader feeds xv_train_step and the queue machinery is not reproduced.  What is kept is the batch contract
(dataset/data_loader.py:322-349): features [num_speakers*num_segments, L, dim] float32 with ONE random length
L in [min_len, max_len] per batch, labels = the speaker index of the spklist, speakers drawn without replacement
(optionally with probability proportional to their frame count), per speaker `num_segments` utterances longer than L
(the utterance list is repeated when it is short; a speaker with no such utterance is replaced), each segment cut at
a random offset.

Files read from the data directory: feats.scp, spk2utt, utt2num_frames (optional: lengths are otherwise taken from the
matrix headers); spklist = "<speaker> <index>" lines (egs/voxceleb/v1/nnet/lib/train.py:83-84)."""
import os

import numpy as np

from . import kaldi_io


class DataOutOfRange(Exception):
    """Kept for call-site compatibility (dataset/data_loader.py:14): the random loader never runs out."""


def read_spklist(path):
    spk2index = {}
    with open(path) as f:
        for line in f:
            parts = line.split()
            if len(parts) >= 2:
                spk2index[parts[0]] = int(parts[1])
    return spk2index


class RandomSegmentLoader(object):
    def __init__(self, data, spklist, num_speakers=64, num_segments=1, min_len=200, max_len=400, dim=None, seed=0,
                 sample_with_prob=False, shuffle=True, threads=8):
        self.table = kaldi_io.ScpTable(os.path.join(data, "feats.scp"))
        keys = self.table.keys()
        index_of = {k: i for i, k in enumerate(keys)}
        n2f = os.path.join(data, "utt2num_frames")
        if os.path.isfile(n2f):
            frames = np.zeros(len(keys), np.int32)
            with open(n2f) as f:
                for line in f:
                    u, n = line.split()
                    if u in index_of:
                        frames[index_of[u]] = int(n)
            cols = self.table.dims(np.arange(1, dtype=np.int64))[1]
        else:
            frames, cols = self.table.dims(threads=threads)
        self.feat_dim = int(cols[0])
        self.dim = int(dim) if dim else self.feat_dim
        self.frames = frames
        spk2index = read_spklist(spklist)
        self.num_total_speakers = len(spk2index)         # dataset/data_loader.py:425 (train.py:86 sizes the softmax with it)
        self.spk_utts = {}                               # speaker index -> scp entry indices
        with open(os.path.join(data, "spk2utt")) as f:
            for line in f:
                parts = line.split()
                if not parts or parts[0] not in spk2index:
                    continue
                utts = [index_of[u] for u in parts[1:] if u in index_of]
                if utts:
                    self.spk_utts[spk2index[parts[0]]] = np.asarray(utts, np.int64)
        self.speakers = sorted(self.spk_utts)
        if not self.speakers:
            raise ValueError("no speaker of %s has features in %s" % (spklist, data))
        self.num_speakers, self.num_segments = int(num_speakers), int(num_segments)
        self.min_len, self.max_len = int(min_len), int(max_len)
        self.shuffle = shuffle
        self.threads = threads
        self.sample_with_prob = sample_with_prob
        self.rng = np.random.default_rng(seed)
        self.spk_frames = np.array([int(self.frames[self.spk_utts[s]].sum()) for s in self.speakers], np.float64)
        # [speaker, slot] tables of scp entry indices and frame counts (-1 / 0 beyond a speaker's utterances): a batch is
        # drawn with a handful of array operations instead of a Python loop over its speakers
        width = max(len(self.spk_utts[s]) for s in self.speakers)
        self.utt_table = np.full((len(self.speakers), width), -1, np.int64)
        self.len_table = np.zeros((len(self.speakers), width), np.int64)
        for row, spk in enumerate(self.speakers):
            u = self.spk_utts[spk]
            self.utt_table[row, :len(u)] = u
            self.len_table[row, :len(u)] = self.frames[u]
        self.spk_labels = np.asarray(self.speakers, np.int32)
        self.max_frames = self.len_table.max(axis=1)
        self.pool = np.arange(len(self.speakers))
        if len(self.pool) < self.num_speakers:            # dataset/data_loader.py:293-297: duplicate speakers
            self.pool = np.tile(self.pool, self.num_speakers // len(self.pool) + 1)

    def _draw_speakers(self, length):
        """Row indices (into self.speakers) of the batch's speakers: drawn without replacement (optionally with probability
        proportional to their frame count), speakers without an utterance longer than `length` replaced by others that have
        one (dataset/data_loader.py:322-333)."""
        n = self.num_speakers
        if self.sample_with_prob and len(self.speakers) >= n:
            rows = self.rng.choice(len(self.speakers), n, replace=False, p=self.spk_frames / self.spk_frames.sum())
        else:
            rows = self.rng.permutation(self.pool)[:n]
        bad = self.max_frames[rows] <= length
        if bad.any():
            usable = np.nonzero(self.max_frames > length)[0]
            if usable.size == 0:
                raise DataOutOfRange("no utterance longer than %d frames" % length)
            fresh = np.setdiff1d(usable, rows[~bad])
            cand = fresh if fresh.size >= bad.sum() else usable
            rows = rows.copy()
            rows[bad] = self.rng.choice(cand, int(bad.sum()), replace=cand.size < bad.sum())
        return rows

    def fetch(self):
        """-> (features [B, L, dim] float32, labels [B] int32), B = num_speakers * num_segments."""
        length = int(self.rng.integers(self.min_len, self.max_len + 1))
        rows = self._draw_speakers(length)
        m = self.num_segments
        # per speaker: m utterances longer than L, without replacement while there are enough of them (the reference repeats
        # the list when it is short): random keys, invalid slots last, the first m slots of the order (cycled over the valid ones)
        ok = self.len_table[rows] > length
        keys = self.rng.random(ok.shape)
        keys[~ok] = 2.0
        order = np.argsort(keys, axis=1)
        n_ok = ok.sum(axis=1)
        slot = order[np.arange(len(rows))[:, None], np.arange(m)[None, :] % n_ok[:, None]]
        idx = self.utt_table[rows[:, None], slot].reshape(-1)
        frames = self.len_table[rows[:, None], slot].reshape(-1)
        labels = np.repeat(self.spk_labels[rows], m)
        b = idx.size
        starts = (self.rng.random(b) * (frames - length + 1)).astype(np.int32) if self.shuffle else np.zeros(b, np.int32)
        if self.dim == self.feat_dim:
            features = self.table.read_segments(idx, starts, length, threads=self.threads,
                                                out=np.empty((b, length, self.dim), np.float32))
        else:
            full = self.table.read_segments(idx, starts, length, threads=self.threads,
                                            out=np.empty((b, length, self.feat_dim), np.float32))
            features = np.ascontiguousarray(full[:, :, :self.dim])
        return features, labels

    def stop(self):
        self.table.close()


class SequentialSegmentLoader(object):
    """Validation batches: the rule of the reference's KaldiDataSeqQueue / batch_sequence (dataset/data_loader.py:498-549,
    :554-660) over the native scp reader.  Every utterance of the speakers in `spklist` appears once per pass, speaker by
    speaker in spk2utt order; there are len // batch_size + 1 batches of min(batch_size, len) utterances, the list wrapping
    around to fill the last one (:528-532); a batch takes ONE length in [min_len, max_len], shortened to its shortest
    utterance (:535-541), and with shuffle=False every segment starts at frame 0.  fetch() raises DataOutOfRange after the
    last batch (:642-657).  (The reference splits the list over `num_parallel` processes whose batches interleave in queue
    order; here one stream yields them in list order.)"""

    def __init__(self, data, spklist, batch_size=128, min_len=200, max_len=400, dim=None, shuffle=False, seed=0, threads=8):
        self.table = kaldi_io.ScpTable(os.path.join(data, "feats.scp"))
        keys = self.table.keys()
        index_of = {k: i for i, k in enumerate(keys)}
        n2f = os.path.join(data, "utt2num_frames")
        if os.path.isfile(n2f):
            frames = np.zeros(len(keys), np.int32)
            with open(n2f) as f:
                for line in f:
                    u, n = line.split()
                    if u in index_of:
                        frames[index_of[u]] = int(n)
            cols = self.table.dims(np.arange(1, dtype=np.int64))[1]
        else:
            frames, cols = self.table.dims(threads=threads)
        self.feat_dim = int(cols[0])
        self.dim = int(dim) if dim else self.feat_dim
        spk2index = read_spklist(spklist)
        self.num_total_speakers = len(spk2index)
        order, labels = [], []
        with open(os.path.join(data, "spk2utt")) as f:
            for line in f:
                parts = line.split()
                if not parts or parts[0] not in spk2index:
                    continue
                for u in parts[1:]:
                    if u in index_of:
                        order.append(index_of[u])
                        labels.append(spk2index[parts[0]])
        if not order:
            raise ValueError("no speaker of %s has features in %s" % (spklist, data))
        self.rng = np.random.default_rng(seed)
        order, labels = np.asarray(order, np.int64), np.asarray(labels, np.int32)
        if shuffle:
            perm = self.rng.permutation(len(order))
            order, labels = order[perm], labels[perm]
        self.num_batches = len(order) // int(batch_size) + 1
        self.batch_size = min(int(batch_size), len(order))
        need = self.num_batches * self.batch_size
        reps = -(-need // len(order))
        self.order, self.labels = np.tile(order, reps)[:need], np.tile(labels, reps)[:need]
        self.frames = frames
        self.min_len, self.max_len = int(min_len), int(max_len)
        self.shuffle = shuffle
        self.threads = threads
        self.next_batch = 0

    def fetch(self):
        """-> (features [batch_size, L, dim] float32, labels [batch_size] int32); DataOutOfRange when the pass is over."""
        if self.next_batch >= self.num_batches:
            raise DataOutOfRange("the sequential loader has delivered its %d batches" % self.num_batches)
        lo = self.next_batch * self.batch_size
        self.next_batch += 1
        idx = self.order[lo:lo + self.batch_size]
        length = int(self.rng.integers(self.min_len, self.max_len + 1))
        length = min(length, int(self.frames[idx].min()))
        if self.shuffle:
            starts = (self.rng.random(len(idx)) * (self.frames[idx] - length + 1)).astype(np.int32)
        else:
            starts = np.zeros(len(idx), np.int32)
        full = self.table.read_segments(idx, starts, length, threads=self.threads,
                                        out=np.empty((len(idx), length, self.feat_dim), np.float32))
        features = full if self.dim == self.feat_dim else np.ascontiguousarray(full[:, :, :self.dim])
        return features, self.labels[lo:lo + self.batch_size].copy()

    def stop(self):
        self.table.close()


class KaldiDataRandomQueue(object):
    """The reference's class by name and signature (dataset/data_loader.py:369-495) over RandomSegmentLoader: the queue
    parameters (num_parallel, max_qsize) are accepted and unused -- batches are produced on demand by fetch(), fast enough to
    stay ahead of the training step -- and start() / stop() keep their places in the caller's loop.  The batch shape may be
    set after construction (set_batch / set_length) as in the reference; the loader is built at start()."""

    def __init__(self, data_dir, spklist, num_parallel=1, max_qsize=20, num_speakers=None, num_segments=None, min_len=None,
                 max_len=None, shuffle=True, sample_with_prob=False):
        self.data, self.spklist = data_dir, spklist
        self.num_parallel_datasets = num_parallel
        self.num_speakers, self.num_segments = num_speakers, num_segments
        self.min_len, self.max_len = min_len, max_len
        self.shuffle, self.sample_with_prob = shuffle, sample_with_prob
        self.num_total_speakers = len(read_spklist(spklist))
        self._loader = None

    def set_batch(self, num_speakers, num_segments):
        self.num_speakers, self.num_segments = num_speakers, num_segments

    def set_length(self, min_len, max_len):
        self.min_len, self.max_len = min_len, max_len

    def start(self):
        self._loader = RandomSegmentLoader(self.data, self.spklist, num_speakers=self.num_speakers, num_segments=self.num_segments,
                                           min_len=self.min_len, max_len=self.max_len, shuffle=self.shuffle,
                                           sample_with_prob=self.sample_with_prob, seed=int.from_bytes(os.urandom(4), "little"))

    def fetch(self):
        if self._loader is None:
            self.start()
        return self._loader.fetch()

    def stop(self):
        if self._loader is not None:
            self._loader.stop()
            self._loader = None


class KaldiDataSeqQueue(object):
    """dataset/data_loader.py:554-660 by name and signature over SequentialSegmentLoader (see KaldiDataRandomQueue)."""

    def __init__(self, data_dir, spklist, num_parallel=1, max_qsize=20, batch_size=128, min_len=None, max_len=None, shuffle=True):
        self.data, self.spklist = data_dir, spklist
        self.num_parallel_datasets = num_parallel
        self.batch_size, self.min_len, self.max_len, self.shuffle = batch_size, min_len, max_len, shuffle
        self.num_total_speakers = len(read_spklist(spklist))
        self._loader = None

    def set_batch(self, batch_size):
        self.batch_size = batch_size

    def set_length(self, min_len, max_len):
        self.min_len, self.max_len = min_len, max_len

    def start(self):
        self._loader = SequentialSegmentLoader(self.data, self.spklist, batch_size=self.batch_size, min_len=self.min_len,
                                               max_len=self.max_len, shuffle=self.shuffle,
                                               seed=int.from_bytes(os.urandom(4), "little"))

    def fetch(self):
        if self._loader is None:
            self.start()
        return self._loader.fetch()

    def stop(self):
        if self._loader is not None:
            self._loader.stop()
            self._loader = None
