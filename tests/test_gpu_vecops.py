"""GPU: post-extraction vector operations through the C-ABI against the fp64 restatement."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = 2e-6          # fp32 kernels against fp64 numpy, relative to the result's scale


@pytest.fixture(scope="module")
def eng():
    import xvector_b200
    e = xvector_b200.Engine({"network_type": "tdnn", "pooling_type": "statistics_pooling"}, 8, precision="fp32")
    yield e
    e.close()


@pytest.mark.parametrize("n,dim", [(1, 512), (257, 512), (33, 200), (5, 7)])
def test_normalize_length(eng, n, dim):
    from oracle import vecops_ref
    rng = np.random.default_rng(n + dim)
    x = (rng.standard_normal((n, dim)) * rng.uniform(0.01, 30, (n, 1))).astype(np.float32)
    if n > 2:
        x[2] = 0                                         # "Zero iVector": written unchanged
    for scaleup in (False, True):
        out = eng.normalize_length(x, scaleup)
        ref = vecops_ref.normalize_length(x, scaleup)
        assert np.abs(out - ref).max() <= TOL * max(1.0, np.sqrt(dim) if scaleup else 1.0)
    unit = eng.normalize_length(x)
    norms = np.linalg.norm(unit.astype(np.float64), axis=1)
    ok = np.ones(n, bool)
    if n > 2:
        ok[2] = False
        assert not unit[2].any()
    assert np.abs(norms[ok] - 1).max() < 1e-6


def test_speaker_means_recipe_stage(eng, tmp_path):
    from oracle import vecops_ref
    from xvector_b200 import kaldi_io, vecops
    rng = np.random.default_rng(4)
    keys = ["spk%02d-utt%03d" % (i % 11, i) for i in range(300)]
    vecs = (rng.standard_normal((300, 512)) + rng.standard_normal((11, 512))[np.arange(300) % 11] * 2).astype(np.float32)
    spk2utt = [("spk%02d" % s, [k for k in keys if k.startswith("spk%02d-" % s)] + ["missing-%d" % s]) for s in range(11)]
    spk2utt.append(("nobody", ["ghost-1", "ghost-2"]))                   # no vector at all: dropped
    for normalize in (False, True):
        spk, means, counts = vecops.speaker_means(eng, keys, vecs, spk2utt, normalize)
        rspk, rmeans, rcounts = vecops_ref.speaker_means(keys, vecs, spk2utt, normalize)
        assert spk == rspk and list(counts) == list(rcounts)
        assert np.abs(means - rmeans).max() <= TOL * np.abs(rmeans).max() * 4
    # ark to ark
    xv, s2u, out, nu = (str(tmp_path / n) for n in ("xvector.ark", "spk2utt", "spk_xvector.ark", "num_utts.ark"))
    kaldi_io.write_vecs_flt(xv, keys, vecs)
    with open(s2u, "w") as f:
        for s, utts in spk2utt:
            f.write("%s %s\n" % (s, " ".join(utts)))
    vecops.speaker_means_ark(eng, s2u, "ark:" + xv, "ark:" + out, nu, normalize=True)
    got = dict(kaldi_io.read_vec_flt_ark(out))
    assert list(got) == rspk and np.abs(np.stack(list(got.values())) - rmeans).max() <= 1e-5
    assert open(nu).read().split("\n")[0] == "spk00 %d" % rcounts[0]


def test_cosine_scores_and_pairwise_eer(eng):
    from oracle import vecops_ref
    from xvector_b200 import vecops
    rng = np.random.default_rng(9)
    a = rng.standard_normal((130, 512)).astype(np.float32)
    b = rng.standard_normal((77, 512)).astype(np.float32)
    s = eng.cosine_scores(a, b)
    an = a.astype(np.float64) / np.sqrt((a.astype(np.float64) ** 2).sum(1, keepdims=True) + 1e-12)
    bn = b.astype(np.float64) / np.sqrt((b.astype(np.float64) ** 2).sum(1, keepdims=True) + 1e-12)
    assert np.abs(s - an @ bn.T).max() <= 2e-6
    # compute_cos_pairwise_eer: 2300 embeddings of 23 speakers, down-sampled by the reference's stride to 1150
    labels = np.arange(2300) % 23
    emb = (rng.standard_normal((2300, 64)) + 1.5 * rng.standard_normal((23, 64))[labels]).astype(np.float32)
    scores, keys = vecops.pairwise_trials(eng, emb, labels, max_num_embeddings=1000)
    rscores, rkeys = vecops_ref.pairwise_trials(emb, labels, max_num_embeddings=1000)
    assert scores.shape == rscores.shape == (1150 * 1149 // 2,) and np.array_equal(keys, rkeys)
    assert np.abs(scores - rscores).max() <= 2e-6
    eer = vecops.compute_cos_pairwise_eer(eng, emb, labels)
    from scipy.interpolate import interp1d
    from scipy.optimize import brentq
    from sklearn import metrics
    fpr, tpr, _ = metrics.roc_curve(rkeys, rscores, pos_label=1)
    ref = brentq(lambda x: 1. - x - interp1d(fpr, tpr)(x), 0., 1.)
    assert 0.0 < eer < 0.5 and abs(eer - ref) <= 1e-4
