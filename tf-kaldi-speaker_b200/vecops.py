"""Post-extraction steps of the recipe on the device: egs/voxceleb/v1/nnet/run_extract_embeddings.sh:80-103
(length normalisation, per-speaker means) and misc/utils.py:307-346 (pairwise cosine EER), over libxvector.so's
xv_vec_normalize_length / xv_vec_group_mean / xv_cosine_scores.  Kaldi's ivector-normalize-length / ivector-mean are
third-party binaries that are not in the reference tree; their published behaviour is restated in oracle/vecops_ref.py."""
import numpy as np


def read_spk2utt(path):
    """Kaldi spk2utt: one line per speaker, "spk utt1 utt2 ..." -> [(spk, [utts])] in file order."""
    out = []
    with open(path, "r") as f:
        for line in f:
            parts = line.split()
            if parts:
                out.append((parts[0], parts[1:]))
    return out


def speaker_means(engine, keys, vectors, spk2utt, normalize=False):
    """Stage 2 of run_extract_embeddings.sh (:80-95).  keys/vectors: the utterance x-vectors; spk2utt: [(spk, [utts])].
    normalize=True is the `$normalize` branch: length-normalise the utterance vectors, average, length-normalise the
    means (all with --scaleup=false).  Utterances missing from `keys` are left out (ivector-mean counts them as
    errors and goes on); a speaker with no vector at all is dropped.  -> (spk_keys, means [n_spk, dim], num_utts)."""
    vectors = np.ascontiguousarray(vectors, np.float32)
    row = {k: i for i, k in enumerate(keys)}
    group = np.full(len(keys), -1, np.int32)
    order = []                                     # ivector-mean sums in spk2utt order: rows are re-ordered to match
    for g, (spk, utts) in enumerate(spk2utt):
        for u in utts:
            if u in row:
                order.append((row[u], g))
    if normalize:
        vectors = engine.normalize_length(vectors, scaleup=False)
    idx = np.array([i for i, _ in order], np.int64)
    grp = np.array([g for _, g in order], np.int32)
    means, counts = engine.group_mean(vectors[idx] if len(idx) else vectors[:0], grp, len(spk2utt))
    keep = counts > 0
    means = means[keep]
    if normalize and len(means):
        means = engine.normalize_length(means, scaleup=False)
    spk_keys = [s for (s, _), k in zip(spk2utt, keep) if k]
    return spk_keys, means, counts[keep]


def speaker_means_ark(engine, spk2utt_path, rspecifier, wspecifier, num_utts_path=None, normalize=False):
    """The same stage ark to ark: x-vector ark/scp in, spk_xvector ark out, num_utts as a text ark ("spk n" lines)."""
    from . import kaldi_io
    reader = kaldi_io.read_vec_flt_scp if rspecifier.split(":")[0].startswith("scp") or rspecifier.endswith(".scp") else kaldi_io.read_vec_flt_ark
    keys, vecs = [], []
    for k, v in reader(rspecifier):
        keys.append(k)
        vecs.append(v)
    spk_keys, means, counts = speaker_means(engine, keys, np.stack(vecs), read_spk2utt(spk2utt_path), normalize)
    kaldi_io.write_vecs_flt(wspecifier, spk_keys, means)
    if num_utts_path:
        with open(num_utts_path, "w") as f:
            for s, n in zip(spk_keys, counts):
                f.write("%s %d\n" % (s, n))
    return spk_keys, means, counts


def pairwise_trials(engine, embeddings, labels, max_num_embeddings=1000):
    """Scores and target flags of compute_cos_pairwise_eer (misc/utils.py:318-335): cosine of every pair i < j after
    the strided down-sampling to at most max_num_embeddings vectors."""
    embeddings = np.asarray(embeddings, np.float32)
    labels = np.asarray(labels)
    n = embeddings.shape[0]
    if n > max_num_embeddings:
        step = n // max_num_embeddings                          # Python-2 integer division in the reference (:322)
        embeddings = embeddings[range(0, n, step), :]
        labels = labels[range(0, n, step)]
        n = embeddings.shape[0]
    score_mat = engine.cosine_scores(embeddings, None, eps=1e-12)
    iu = np.triu_indices(n, 1)                                  # row-major order of the reference's double loop
    return score_mat[iu].astype(np.float64), (labels[iu[0]] == labels[iu[1]]).astype(np.float64)


def compute_cos_pairwise_eer(engine, embeddings, labels, max_num_embeddings=1000):
    """misc/utils.py:307-346 with the score matrix computed on the device.  The reference estimates the EER with
    sklearn's roc_curve + interp1d + brentq ("not the exact value"); the same estimator is used here."""
    from scipy.interpolate import interp1d
    from scipy.optimize import brentq
    from sklearn import metrics
    scores, keys = pairwise_trials(engine, embeddings, labels, max_num_embeddings)
    fpr, tpr, _ = metrics.roc_curve(keys, scores, pos_label=1)
    return brentq(lambda x: 1. - x - interp1d(fpr, tpr)(x), 0., 1.)
