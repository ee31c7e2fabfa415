"""Post-extraction vector operations restated in numpy fp64.  TEST INFRASTRUCTURE: only tests/
import this module.

parity unpinned for the two Kaldi steps: `ivector-normalize-length` and `ivector-mean` are Kaldi binaries
(kaldi-asr/kaldi, src/ivectorbin/) that run_extract_embeddings.sh:80-103 calls and that are absent from
/root/reference; their published algorithm is restated here (norm = |x|; ratio = norm / sqrt(dim), or norm with
--scaleup=false; x *= 1/ratio unless ratio == 0 -- ivector-mean: per speaker of spk2utt, the sum of the utterance
vectors that exist divided by their count, speakers without any vector dropped).  The cosine scores follow the
reference's own code (misc/utils.py:318-335) and are pinned by running that arithmetic directly in the test."""
import numpy as np


def normalize_length(x, scaleup=False):
    x = np.asarray(x, np.float64)
    norm = np.sqrt(np.sum(x * x, axis=1, keepdims=True))
    ratio = norm / np.sqrt(x.shape[1]) if scaleup else norm
    return np.where(ratio == 0.0, x, x / np.where(ratio == 0.0, 1.0, ratio))


def speaker_means(keys, vectors, spk2utt, normalize=False):
    vectors = np.asarray(vectors, np.float64)
    if normalize:
        vectors = normalize_length(vectors, False)
    row = {k: i for i, k in enumerate(keys)}
    spk, means, counts = [], [], []
    for s, utts in spk2utt:
        have = [row[u] for u in utts if u in row]
        if not have:
            continue
        spk.append(s)
        means.append(vectors[have].sum(axis=0) / len(have))
        counts.append(len(have))
    means = np.array(means)
    if normalize and len(means):
        means = normalize_length(means, False)
    return spk, means, np.array(counts)


def pairwise_trials(embeddings, labels, max_num_embeddings=1000):
    """misc/utils.py:318-335 (integer division restored for Python 3)."""
    embeddings = np.array(embeddings, np.float64)
    labels = np.asarray(labels)
    embeddings /= np.sqrt(np.sum(embeddings ** 2, axis=1, keepdims=True) + 1e-12)
    n = embeddings.shape[0]
    if n > max_num_embeddings:
        step = n // max_num_embeddings
        embeddings = embeddings[range(0, n, step), :]
        labels = labels[range(0, n, step)]
        n = embeddings.shape[0]
    score_mat = np.dot(embeddings, np.transpose(embeddings))
    scores = np.zeros(n * (n - 1) // 2)
    keys = np.zeros(n * (n - 1) // 2)
    index = 0
    for i in range(n - 1):
        for j in range(i + 1, n):
            scores[index] = score_mat[i, j]
            keys[index] = 1 if labels[i] == labels[j] else 0
            index += 1
    return scores, keys
