"""CPU: the oracle against the reference-derived golden vectors, and the two restatements against each other."""
import os

import numpy as np
import pytest

from oracle import numpy_ref, synth, torch_ref

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def test_attention_core_matches_reference_numpy_statement():
    """Pinned: outputs of the reference's own compute_self_attention (model/test_utils.py:321-392)."""
    g = np.load(os.path.join(GOLDEN, "attention_golden.npz"))
    names = sorted({k.split("/")[0] for k in g.files})
    assert len(names) == 5
    for n in names:
        sv, sk, sc = [bool(v) for v in g[n + "/flags"]]
        att, w = numpy_ref.attention_core(g[n + "/value"], g[n + "/key"], g[n + "/query"], sv, sk, sc)
        np.testing.assert_allclose(att, g[n + "/att"], rtol=1e-12, atol=1e-14)
        np.testing.assert_allclose(w.sum(axis=2), 1.0, rtol=1e-12)


def test_stat_pooling_matches_reference_numpy_statement():
    """Pinned: p1 = [np.mean | np.std] of the reference's compute_phone_pooling (model/test_utils.py:1110-1112), and the
    same statistic from its explicit loops (:1114-1128, +1e-16 under the root).  np.std has no variance floor; where the
    variance is exactly 0 the oracle must give the floor of model/pooling.py:46-47 (sqrt(1e-12))."""
    g = np.load(os.path.join(GOLDEN, "statpool_golden.npz"))
    names = sorted({k.split("/")[0] for k in g.files})
    assert names == ["offset", "relu_like", "small"]
    for n in names:
        x, p1, loops = g[n + "/x"], g[n + "/p1"], g[n + "/loops"]
        c = x.shape[2]
        out = numpy_ref.statistics_pooling(x)
        np.testing.assert_allclose(out[:, :c], p1[:, :c], rtol=1e-12, atol=1e-12)
        live = p1[:, c:] > 1e-5
        assert (~live).any() == (n != "small")
        np.testing.assert_allclose(out[:, c:][live], p1[:, c:][live], rtol=1e-9)
        np.testing.assert_allclose(out[:, c:][~live], 1e-6, rtol=1e-9)          # VAR2STD_EPSILON floor
        np.testing.assert_allclose(out[:, :c], loops[:, :c], rtol=1e-10, atol=1e-10)
        np.testing.assert_allclose(out[:, c:][live], loops[:, c:][live], rtol=1e-6)


def test_oracle_regression_fixtures():
    g = np.load(os.path.join(GOLDEN, "tdnn_small_golden.npz"))
    for tag, cfg in (("stat", synth.TDNN_STAT_CONFIG), ("att", dict(synth.TDNN_ATT_CONFIG, att_num_heads=3))):
        cfg = dict(cfg, num_nodes_pooling_layer=1500)
        w = synth.make_tdnn_weights(23, cfg, seed=99)
        _, ep = numpy_ref.entire_network(g[tag + "/x"], numpy_ref.OracleParams(**cfg), w)
        for node in ("tdnn1_relu", "tdnn3_relu", "tdnn5_relu", "pooling", "tdnn6_dense", "output"):
            v = ep[node]
            v = v[:, 0] if v.ndim == 4 else v            # conv2d endpoints: [B, 1, T', 512] (model/tdnn.py:38-91)
            v = v if v.ndim < 3 else v[:, ::7, ::37]
            np.testing.assert_allclose(v, g[tag + "/" + node], rtol=1e-10, atol=1e-12)


@pytest.mark.parametrize("cfg", [
    synth.TDNN_STAT_CONFIG,
    dict(synth.TDNN_STAT_CONFIG, network_relu_type="prelu"),
    dict(synth.TDNN_STAT_CONFIG, network_relu_type="lrelu"),
    synth.TDNN_ATT_CONFIG,
    dict(synth.TDNN_ATT_CONFIG, att_num_heads=4),
    dict(synth.TDNN_ATT_CONFIG, att_num_heads=5, att_split_key=False, att_key_network_type=3),
    dict(synth.TDNN_ATT_CONFIG, att_num_heads=2, att_split_value=False, att_key_network_type=2, att_apply_nonlinear=True),
], ids=["stat", "prelu", "lrelu", "att1", "att4", "att5_sharedkey_tanh", "att2_nosplit_post"])
def test_numpy_and_torch_restatements_agree(cfg):
    """Two independent implementations (fp64 numpy, fp32 torch kernels) of the same graph agree to <= 1e-5."""
    w = synth.make_tdnn_weights(30, cfg, seed=4)
    x = synth.config1_features(2, 80, 30, seed=3)
    e64 = numpy_ref.predict(x, numpy_ref.OracleParams(**cfg), w, 30)
    e32 = torch_ref.TorchTdnn(w, cfg).embed(x)
    assert e64.shape == (2, 512)
    assert np.abs(e64 - e32).max() / np.abs(e64).max() < 1e-5


def test_stat_pooling_semantics():
    rng = np.random.default_rng(0)
    x = rng.standard_normal((3, 17, 5))
    x[:, :, 2] = 1.25                                     # constant channel -> variance floor
    p = numpy_ref.statistics_pooling(x)
    np.testing.assert_allclose(p[:, :5], x.mean(axis=1))
    np.testing.assert_allclose(p[:, [5, 6, 8, 9]], x.std(axis=1)[:, [0, 1, 3, 4]])    # population std
    np.testing.assert_allclose(p[:, 7], 1e-6)            # sqrt(1e-12)


def test_tdnn_shapes_context_and_defaults():
    cfg = dict(synth.TDNN_STAT_CONFIG)
    w = synth.make_tdnn_weights(30, cfg)
    p = numpy_ref.OracleParams(**cfg)
    x = synth.config1_features(1, 40, 30)
    out, ep = numpy_ref.entire_network(x, p, w)
    assert ep["tdnn1_relu"].shape == (1, 1, 36, 512) and ep["tdnn2_relu"].shape == (1, 1, 32, 512)    # model/tdnn.py:38-91
    assert ep["tdnn3_bn"].shape == (1, 1, 26, 512)
    assert ep["tdnn3_relu"].shape == (1, 26, 512) and ep["tdnn5_relu"].shape == (1, 26, 1500)
    assert ep["pooling"].shape == (1, 3000) and ep["tdnn6_dense"].shape == (1, 512) and out.shape == (1, 512)
    # lazily injected defaults are visible on params afterwards (model/tdnn.py:114-116,152-154)
    assert p.num_nodes_pooling_layer == 1500 and p.num_nodes_last_layer == 512
    assert p.last_layer_no_bn is False and p.last_layer_linear is False
    with pytest.raises(ValueError):
        numpy_ref.entire_network(x[:, :14], p, w)        # shorter than the 15-frame receptive field


def test_predict_rank_and_dim_slicing():
    cfg = dict(synth.TDNN_STAT_CONFIG)
    w = synth.make_tdnn_weights(20, cfg)
    p = numpy_ref.OracleParams(**cfg)
    x = synth.config1_features(2, 50, 24)
    e3 = numpy_ref.predict(x, p, w, 20)
    e2 = numpy_ref.predict(x[0], p, w, 20)
    assert e3.shape == (2, 512) and e2.shape == (512,)
    np.testing.assert_allclose(e2, e3[0], rtol=1e-12)
    np.testing.assert_allclose(numpy_ref.predict(x[:, :, :20], p, w, 20), e3, rtol=1e-12)


def test_chunk_plan_follows_extractor():
    # extract.py:68-78 with Python-2 integer division
    assert numpy_ref.chunk_plan(25000, 10000) == [(0, 10000), (5000, 10000), (10000, 10000), (15000, 10000)]
    assert numpy_ref.chunk_plan(10001, 10000) == [(0, 10000), (5000, 5001)]
    assert numpy_ref.chunk_plan(2501, 1001) == [(0, 1001), (500, 1001), (1000, 1001), (1500, 1001)]
    plan = numpy_ref.chunk_plan(12345, 2000)
    assert plan[0] == (0, 2000) and all(z == 2000 for _, z in plan[:-1]) and 1000 < plan[-1][1] <= 2000
    assert plan[-1][0] + plan[-1][1] == 12345


def test_extract_embedding_chunk_average():
    calls = []

    def fake_predict(x):
        x = np.asarray(x)
        calls.append(x.shape)
        return x.mean(axis=-2)[..., :4].astype(np.float64)

    feat = np.random.default_rng(1).standard_normal((250, 6))
    assert numpy_ref.extract_embedding(feat[:10], fake_predict, 100, 25) is None
    e = numpy_ref.extract_embedding(feat, fake_predict, chunk_size=100, min_chunk_size=25)
    assert calls == [(3, 100, 6), (100, 6)]
    plan = numpy_ref.chunk_plan(250, 100)
    want = sum(z * feat[s:s + z].mean(axis=0)[:4] for s, z in plan) / sum(z for _, z in plan)
    np.testing.assert_allclose(e, want, rtol=1e-5, atol=1e-7)   # chunks are cast to float32 (extract.py:81)
    e_n = numpy_ref.extract_embedding(feat, fake_predict, 100, 25, normalize=True)
    np.testing.assert_allclose(np.linalg.norm(e_n), 1.0)


def test_flops_model():
    assert abs(numpy_ref.tdnn_flops(300, 30) / 1e9 - 2.4529) < 1e-3
    assert abs(numpy_ref.tdnn_flops(200, 30) / 1e9 - 1.6023) < 1e-3


@pytest.mark.parametrize("dim,ts,mp,relu,frames", [
    (40, False, False, None, 33),            # the shipped resnet18_softmax_1e-2.json
    (16, True, True, "prelu", 37),           # time stride (odd T -> asymmetric SAME padding) + max pooling
    (24, True, False, "lrelu", 30),
], ids=["v3_config", "tstride_maxpool_prelu", "tstride_lrelu"])
def test_resnet18_restatements_agree(dim, ts, mp, relu, frames):
    cfg = dict(synth.RESNET_CONFIG, resnet_time_stride=ts, resnet_maxpooling=mp)
    if relu:
        cfg["network_relu_type"] = relu
    w = synth.make_resnet_weights(dim, cfg, seed=11)
    x = synth.fbank_like(2 * frames, dim, seed=5).reshape(2, frames, dim)
    e64 = numpy_ref.predict(x, numpy_ref.OracleParams(**cfg), w, dim)
    e32 = torch_ref.make_net(w, cfg).embed(x)
    assert e64.shape == (2, 512)
    assert np.abs(e64 - e32).max() / np.abs(e64).max() < 1e-5


def test_same_padding_is_tensorflow_asymmetric():
    # 40 -> 20 with k=3, s=2: TF pads (0, 1), not the symmetric (1, 1) of other frameworks
    assert numpy_ref.same_pads(40, 3, 2) == (20, 0, 1)
    assert numpy_ref.same_pads(5, 3, 2) == (3, 1, 1)
    assert numpy_ref.same_pads(37, 3, 1) == (37, 1, 1)
    assert numpy_ref.same_pads(37, 1, 2) == (19, 0, 0)
    x = np.arange(2 * 6 * 4, dtype=np.float64).reshape(1, 6, 4, 2)
    k = np.zeros((3, 3, 2, 1))
    k[0, 0, 0, 0] = 1.0                       # picks input (t-pt0+0, f-pf0+0), channel 0
    y = numpy_ref.conv2d_same(x, k, (1, 2))
    # time: pad (1,1) -> row t-1; freq: pad (0,1) with stride 2 -> column 2*fo
    assert y.shape == (1, 6, 2, 1)
    np.testing.assert_array_equal(y[0, 1:, :, 0], x[0, :-1, ::2, 0])
    np.testing.assert_array_equal(y[0, 0], 0.0)


def test_etdnn_restatements_agree():
    cfg = dict(synth.ETDNN_STAT_CONFIG)
    w = synth.make_etdnn_weights(30, cfg, seed=4)
    x = synth.config1_features(2, 90, 30, seed=3)
    e64 = numpy_ref.predict(x, numpy_ref.OracleParams(**cfg), w, 30)
    e32 = torch_ref.make_net(w, cfg).embed(x)
    assert e64.shape == (2, 512)
    assert np.abs(e64 - e32).max() / np.abs(e64).max() < 1e-5
    _, ep = numpy_ref.entire_network(x, numpy_ref.OracleParams(**cfg), w)
    assert ep["tdnn10_relu"].shape == (2, 90 - 22, 1500)          # context 4+4+6+8
