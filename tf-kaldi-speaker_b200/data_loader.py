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
import random

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
        self.rd = random.Random(seed)
        self.spk_frames = np.array([int(self.frames[self.spk_utts[s]].sum()) for s in self.speakers], np.float64)
        self.pool = list(self.speakers)
        if len(self.pool) < self.num_speakers:            # dataset/data_loader.py:293-297: duplicate speakers
            self.pool = self.pool * (self.num_speakers // len(self.pool) + 1)

    def _draw_speakers(self):
        if self.sample_with_prob:
            p = self.spk_frames / self.spk_frames.sum()
            rng = np.random.default_rng(self.rd.getrandbits(32))
            if len(self.speakers) >= self.num_speakers:
                return [self.speakers[i] for i in rng.choice(len(self.speakers), self.num_speakers, replace=False, p=p)]
        return self.rd.sample(self.pool, self.num_speakers)

    def fetch(self):
        """-> (features [B, L, dim] float32, labels [B] int32), B = num_speakers * num_segments."""
        batch_speakers = self._draw_speakers()
        length = self.rd.randint(self.min_len, self.max_len)
        b = self.num_speakers * self.num_segments
        labels = np.zeros(b, np.int32)
        picks = []
        for i, spk in enumerate(batch_speakers):
            for _ in range(1000):
                utts = self.spk_utts[spk]
                ok = utts[self.frames[utts] > length]
                if len(ok):
                    break
                # the speaker has no utterance long enough for this batch: resample it (dataset/data_loader.py:330-333)
                rest = list(set(self.speakers) - set(batch_speakers)) or self.speakers
                spk = self.rd.choice(rest)
                batch_speakers[i] = spk
            else:
                raise DataOutOfRange("no utterance longer than %d frames" % length)
            labels[i * self.num_segments:(i + 1) * self.num_segments] = spk
            ok = list(ok)
            if len(ok) < self.num_segments:
                ok = ok * (self.num_segments // len(ok) + 1)
            picks.extend(self.rd.sample(ok, self.num_segments))
        idx = np.asarray(picks, np.int64)
        rows = self.frames[idx]
        packed, _ = self.table.read_many(idx, rows=rows, threads=self.threads,
                                         out=np.empty((int(rows.sum()), self.feat_dim), np.float32))
        off = np.concatenate([[0], np.cumsum(rows, dtype=np.int64)])
        features = np.empty((b, length, self.dim), np.float32)
        for j in range(b):
            start = self.rd.randint(0, int(rows[j]) - length) if self.shuffle else 0
            features[j] = packed[off[j] + start:off[j] + start + length, :self.dim]
        return features, labels

    def stop(self):
        self.table.close()
