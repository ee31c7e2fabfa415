"""GPU: end-to-end parity of the CUDA path (through the C-ABI) with the fp64 oracle on seeded inputs.
Tolerances: BASELINE.json north_star / SURVEY.md 8(c): per utterance cos >= 0.9999 and
max_i|e - e_ref| / max_i|e_ref| <= 1e-2 (bf16) / 1e-4 (fp32 validation mode)."""
import numpy as np
import pytest

from util import MIN_COS, TOL, max_rel, min_cos, oracle_embeddings, tensor_rel

pytestmark = pytest.mark.gpu


def _engine(cfg, dim, precision, weights, **kw):
    import xvector_b200
    e = xvector_b200.Engine(cfg, dim, precision=precision, **kw)
    e.set_weights(weights)
    return e


def _check(out, ref, precision):
    assert np.isfinite(out).all()
    assert max_rel(out, ref) <= TOL[precision], "max_rel %.3e" % max_rel(out, ref)
    assert min_cos(out, ref) >= MIN_COS


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_config1_subset_stat_pooling(precision):
    """cfg 1: standard x-vector TDNN, 30-dim MFCC x 300 frames (a 24-utterance subset keeps the oracle fast)."""
    from oracle import synth
    cfg = dict(synth.TDNN_STAT_CONFIG)
    w = synth.make_tdnn_weights(30, cfg, seed=1234)
    x = synth.config1_features(24, 300, 30, seed=101)
    lengths = np.full(24, 300, np.int32)
    ref = np.stack(oracle_embeddings(x.reshape(-1, 30), lengths, cfg, w, 30))
    eng = _engine(cfg, 30, precision, w)
    out = eng.forward(x.reshape(-1, 30), lengths)
    _check(out, ref, precision)
    assert eng.launch_count() > 0
    eng.close()


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_ragged_lengths_and_small_batches(precision):
    """cfg 2 shape: variable lengths incl. the minimum (15 frames = one valid output frame), utterances that
    straddle row tiles, and a frame budget small enough to force several batches."""
    from oracle import synth
    cfg = dict(synth.TDNN_STAT_CONFIG)
    w = synth.make_tdnn_weights(30, cfg, seed=1234)
    lengths = np.array([15, 16, 127, 128, 129, 200, 1000, 33, 2500, 256, 15, 641], np.int32)
    feats = synth.packed_features(lengths, 30, seed=202)
    ref = np.stack(oracle_embeddings(feats, lengths, cfg, w, 30))
    eng = _engine(cfg, 30, precision, w, batch_frames=1024)
    out = eng.forward(feats, lengths)
    _check(out, ref, precision)
    # permutation invariance: utterances are independent
    perm = np.random.default_rng(0).permutation(len(lengths))
    offs = np.concatenate([[0], np.cumsum(lengths)])
    feats_p = np.concatenate([feats[offs[i]:offs[i + 1]] for i in perm])
    out_p = eng.forward(feats_p, lengths[perm])
    assert max_rel(out_p, out[perm]) < (1e-5 if precision == "fp32" else 2e-3)
    eng.close()


def test_error_behaviour():
    import xvector_b200
    from oracle import synth
    cfg = dict(synth.TDNN_STAT_CONFIG)
    w = synth.make_tdnn_weights(30, cfg)
    eng = xvector_b200.Engine(cfg, 30, precision="fp32")
    x = np.zeros((40, 30), np.float32)
    with pytest.raises(xvector_b200.XVectorError) as ei:
        eng.forward(x, np.array([40], np.int32))
    assert ei.value.code == -4                       # XV_ESTATE: weights not finalized
    bad = dict(w)
    del bad["tdnn/tdnn3_bn/gamma"]
    with pytest.raises(xvector_b200.XVectorError, match="tdnn/tdnn3_bn/gamma"):
        eng.set_weights(bad)
    eng.set_weights(w)
    with pytest.raises(xvector_b200.XVectorError, match="context"):
        eng.forward(x[:14], np.array([14], np.int32))    # 14 frames <= receptive field - 1
    with pytest.raises(xvector_b200.XVectorError, match="unknown node"):
        eng.forward(x, np.array([40], np.int32), node="tdnn9_dense")
    assert eng.forward(np.zeros((0, 30), np.float32), np.zeros(0, np.int32)).shape == (0, 512)
    eng.close()
    with pytest.raises(xvector_b200.XVectorError, match="Not implement ftdnn network"):
        xvector_b200.Engine(dict(cfg, network_type="ftdnn"), 30)
    with pytest.raises(xvector_b200.XVectorError, match="Not implement lde pooling"):
        xvector_b200.Engine(dict(cfg, pooling_type="lde"), 30)


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("relu", [None, "prelu", "lrelu"])
def test_every_endpoint(precision, relu):
    """All endpoints of model/tdnn.py:48-179, frame level and utterance level, incl. l2_scaling of `output`."""
    from oracle import numpy_ref, synth
    cfg = dict(synth.TDNN_STAT_CONFIG, feature_norm=True, feature_scaling_factor=7.5, num_nodes_last_layer=400)
    if relu:
        cfg["network_relu_type"] = relu
    w = synth.make_tdnn_weights(24, cfg, seed=77)
    lengths = np.array([90, 41], np.int32)
    feats = synth.packed_features(lengths, 24, seed=8)
    offs = np.concatenate([[0], np.cumsum(lengths)])
    eps = [numpy_ref.entire_network(feats[None, offs[i]:offs[i + 1]], numpy_ref.OracleParams(**cfg), w)[1] for i in range(2)]
    eng = _engine(cfg, 24, precision, w)
    tol = TOL[precision]
    for node in eps[0].keys():
        width, frame_level, ctx = eng.node_info(node)
        if frame_level:
            got = eng.forward_frames(feats, lengths, node)
            for u in range(2):
                ref = eps[u][node][0]
                if ref.ndim == 3:          # tdnn()'s conv2d endpoints are [B, 1, T', 512]; the C-ABI hands out [T', 512] blocks
                    ref = ref[0]
                assert got[u].shape == ref.shape, node
                assert tensor_rel(got[u], ref) <= tol, "%s: %.3e" % (node, tensor_rel(got[u], ref))
        else:
            got = eng.forward(feats, lengths, node)
            ref = np.stack([eps[u][node][0] for u in range(2)])
            assert got.shape == ref.shape, node
            assert max_rel(got, ref) <= tol, "%s: %.3e" % (node, max_rel(got, ref))
    eng.close()


ATT_VARIANTS = {
    "shipped_h1": {},
    "h4_split": {"att_num_heads": 4},
    "h5_shared_key_tanh": {"att_num_heads": 5, "att_split_key": False, "att_key_network_type": 3},
    "h2_nosplit_value_bn_post": {"att_num_heads": 2, "att_split_value": False, "att_key_network_type": 2,
                                 "att_apply_nonlinear": True},
    "value_net_key_linear": {"att_num_heads": 3, "att_value_num_nodes": [600], "att_value_network_type": 1,
                             "att_key_network_type": 0, "att_key_num_nodes": [300], "att_use_scale": False,
                             "att_key_input": "tdnn5_relu"},
}


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("variant", sorted(ATT_VARIANTS))
def test_attention_configs(precision, variant):
    """cfg 3: self-attention pooling (egs/voxceleb/v1/nnet_conf/tdnn_softmax_1e-2_tdnn4_att_pretrain.json and
    multi-head / shared-key / non-split variants), 64-dim fbank, ragged."""
    from oracle import numpy_ref, synth
    cfg = dict(synth.TDNN_ATT_CONFIG, **ATT_VARIANTS[variant])
    w = synth.make_tdnn_weights(64, cfg, seed=303)
    lengths = np.array([300, 77, 15, 450], np.int32)
    rng = np.random.default_rng(303)
    feats = rng.standard_normal((int(lengths.sum()), 64)).astype(np.float32)
    ref = np.stack(oracle_embeddings(feats, lengths, cfg, w, 64))
    eng = _engine(cfg, 64, precision, w)
    out = eng.forward(feats, lengths)
    _check(out, ref, precision)
    # attention_weights endpoint (extract_attention.py reads it)
    offs = np.concatenate([[0], np.cumsum(lengths)])
    wts = eng.forward_frames(feats, lengths, "attention_weights")
    for u in range(len(lengths)):
        _, ep = numpy_ref.entire_network(feats[None, offs[u]:offs[u + 1]], numpy_ref.OracleParams(**cfg), w)
        assert wts[u].shape == ep["attention_weights"][0].shape
        assert np.abs(wts[u] - ep["attention_weights"][0]).max() <= (2e-2 if precision == "bf16" else 1e-4) * ep["attention_weights"][0].max()
    pooled = eng.forward(feats, lengths, "pooling")
    assert pooled.shape[1] == eng.node_info("pooling")[0]
    eng.close()


def test_staged_api_matches_forward_and_pinned_input():
    import xvector_b200
    from oracle import synth
    cfg = dict(synth.TDNN_STAT_CONFIG)
    w = synth.make_tdnn_weights(30, cfg)
    lengths = synth.config2_lengths(40, seed=5)
    feats = synth.packed_features(lengths, 30, seed=5)
    eng = _engine(cfg, 30, "bf16", w, batch_frames=8192)
    a = eng.forward(feats, lengths)
    eng.stage(feats, lengths)
    eng.event_record(0)
    eng.run()
    eng.event_record(1)
    b = eng.fetch()
    assert eng.elapsed_ms(0, 1) > 0
    pin = xvector_b200.PinnedArray(feats.shape)
    pin.array[...] = feats
    c = eng.forward(pin.array, lengths)
    # atomics make the pooled sums order-dependent at the fp32 rounding level only
    assert max_rel(b, a) < 1e-4 and max_rel(c, a) < 1e-4
    pin.close()
    eng.close()


def test_large_job_properties():
    """BASELINE-size property check (no oracle at this size): a 4096-utterance cfg-2 set gives finite embeddings,
    and a random sample of utterances re-run alone gives the same rows (utterance independence)."""
    from oracle import synth
    cfg = dict(synth.TDNN_STAT_CONFIG)
    w = synth.make_tdnn_weights(30, cfg)
    lengths = synth.config2_lengths(4096)
    feats = synth.packed_features(lengths, 30)
    eng = _engine(cfg, 30, "bf16", w)
    out = eng.forward(feats, lengths)
    assert out.shape == (4096, 512) and np.isfinite(out).all()
    offs = np.concatenate([[0], np.cumsum(lengths)])
    pick = np.random.default_rng(1).choice(4096, 16, replace=False)
    sub = eng.forward(np.concatenate([feats[offs[i]:offs[i + 1]] for i in pick]), lengths[pick])
    assert max_rel(sub, out[pick]) < 2e-3
    ref = np.stack(oracle_embeddings(np.concatenate([feats[offs[i]:offs[i + 1]] for i in pick[:4]]), lengths[pick[:4]], cfg, w, 30))
    _check(sub[:4], ref, "bf16")
    eng.close()


def test_storing_layers_are_bit_reproducible_over_200_runs():
    """Stress of the hand-rolled cluster barrier protocol (cta_group::2 pairs: remote CTA-scope mbarrier arrives, multicast
    commits, TMEM hand-back) on the layers that STORE their output: 200 runs of the same ragged batch must give the same
    bits for tdnn1..tdnn4 (a lost or early barrier shows up as a stale or half-written tile sooner or later).  The batch
    has an odd number of row tiles (phantom tile of the last pair) and several batches (ring wrap, both TMEM stages)."""
    from oracle import synth
    cfg = dict(synth.TDNN_STAT_CONFIG)
    w = synth.make_tdnn_weights(30, cfg, seed=3)
    lengths = np.array([700, 333, 1500, 61, 2222, 129], np.int32)          # 4945 rows: 39 row tiles at batch_frames 4096 -> 2 batches
    feats = synth.packed_features(lengths, 30, seed=5)
    eng = _engine(cfg, 30, "bf16", w, batch_frames=4096)
    try:
        first = {}
        for it in range(200):
            node = ("tdnn1_relu", "tdnn2_relu", "tdnn3_relu", "tdnn4_relu")[it % 4]
            got = np.concatenate([g.reshape(-1) for g in eng.forward_frames(feats, lengths, node)])
            if node not in first:
                first[node] = got.copy()
                assert np.isfinite(got).all()
            else:
                assert np.array_equal(got.view(np.uint32), first[node].view(np.uint32)), "run %d of %s differs" % (it, node)
        # the pooled path (fp64 atomics: order-independent at fp32 precision) stays reproducible to the last bit as well
        e0 = eng.forward(feats, lengths)
        for _ in range(20):
            assert np.array_equal(eng.forward(feats, lengths).view(np.uint32), e0.view(np.uint32))
    finally:
        eng.close()
