"""GPU: resnet_18 (implicit-GEMM tcgen05 convolutions) and etdnn parity against the fp64 oracle, through the C-ABI."""
import numpy as np
import pytest

from util import MIN_COS, TOL, bf16_round, max_rel, min_cos, oracle_embeddings, tensor_rel

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    import xvector_b200
    e = xvector_b200.Engine({"network_type": "tdnn", "pooling_type": "statistics_pooling"}, 8, precision="bf16")
    yield e
    e.close()


def _bf16_bits_to_f32(raw):
    return (raw.astype(np.uint32) << 16).view(np.float32)


def _run_conv(eng, x, kernel, st, sf, scale, shift, slope, residual):
    """x [nb,t,f,cin] fp32 (bf16 representable), kernel TF layout [kh,kw,cin,cout]; returns fp32 [nb,to,fo,cout]."""
    nb, t, f, cin = x.shape
    kh, kw, _, cout = kernel.shape
    to, fo = -(-t // st), -(-f // sf)
    wt = np.ascontiguousarray(kernel.reshape(kh * kw * cin, cout).T)              # [cout][K], K = (dt, df, ci)
    bufs = []

    def dev(nbytes):
        p = eng.device_alloc(nbytes)
        bufs.append(p)
        return p
    try:
        d_x, d_w, d_o = dev(x.size * 2), dev(wt.size * 2), dev(nb * to * fo * cout * 2)
        d_s, d_h, d_l = dev(cout * 4), dev(cout * 4), dev(cout * 4)
        eng.upload_bf16(d_x, x)
        eng.upload_bf16(d_w, wt)
        eng.h2d(d_s, scale.astype(np.float32))
        eng.h2d(d_h, shift.astype(np.float32))
        eng.h2d(d_l, slope.astype(np.float32))
        d_r = None
        if residual is not None:
            d_r = dev(residual.size * 2)
            eng.upload_bf16(d_r, residual)
        eng._check(eng.lib.xv_conv2d_bf16_async(eng._ctx, d_x, nb, t, f, cin, d_w, kh, kw, st, sf, cout, d_s, d_h, d_l, d_r, d_o))
        eng.sync()
        raw = np.empty((nb, to, fo, cout), np.uint16)
        eng.d2h(raw, d_o)
    finally:
        for p in bufs:
            eng.device_free(p)
    return _bf16_bits_to_f32(raw)


@pytest.mark.parametrize("nb,t,f,cin,cout,k,st,sf,res", [
    (2, 37, 40, 64, 64, 3, 1, 1, True),        # stage-1 shape, odd T, residual add, 64-wide tile
    (3, 50, 40, 64, 128, 3, 1, 2, False),      # conv2a_conv0: frequency stride 2 (TF pads (0,1))
    (2, 50, 40, 64, 128, 1, 1, 2, False),      # conv2a_conv_short: 1x1 projection with stride
    (2, 40, 10, 256, 512, 3, 2, 2, False),     # time + frequency stride, 256-wide tile, two N tiles
    (1, 400, 5, 512, 512, 3, 1, 1, True),      # stage-4 shape (tile = 5 bins x 25 steps = 125 rows)
    (5, 33, 8, 128, 256, 3, 1, 1, True),       # F = 8 (64-dim fbank, stage 4 of F=64)
    (40, 64, 20, 128, 128, 3, 1, 1, False),    # more tiles than SMs
    (2, 30, 64, 64, 64, 3, 1, 1, True),        # halo tile 16 bins x 7 steps (F = 64 stage 1), ragged last time tile
    (3, 23, 32, 128, 128, 3, 1, 1, True),      # halo tile, two channel blocks per tap (F = 64 stage 2)
    (1, 5, 16, 64, 128, 3, 1, 1, False),       # halo tile taller than the utterance
    (150, 21, 40, 64, 64, 3, 1, 1, True),      # halo tile 40 bins x 3 steps, several tiles per CTA (ring wrap-around)
])
def test_conv2d_implicit_gemm_matches_fp64(eng, nb, t, f, cin, cout, k, st, sf, res):
    from oracle import numpy_ref
    rng = np.random.default_rng(nb * 1000 + t + f + cin + cout + k)
    x = bf16_round(rng.standard_normal((nb, t, f, cin)).astype(np.float32))
    kern = bf16_round((rng.standard_normal((k, k, cin, cout)) / np.sqrt(k * k * cin)).astype(np.float32))
    scale = rng.uniform(0.5, 1.5, cout)
    shift = rng.standard_normal(cout) * 0.1
    slope = np.full(cout, 0.2)
    to, fo = -(-t // st), -(-f // sf)
    residual = bf16_round(rng.standard_normal((nb, to, fo, cout)).astype(np.float32)) if res else None
    out = _run_conv(eng, x, kern, st, sf, scale, shift, slope, residual)
    y = numpy_ref.conv2d_same(x, kern, (st, sf)) * scale + shift
    if res:
        y = y + residual
    ref = np.where(y >= 0, y, 0.2 * y)
    err = np.abs(out - ref)
    assert (err <= 2.0 ** -8 * np.abs(ref) + 1e-3).all(), "max abs err %.4g (rel %.3g)" % (err.max(), tensor_rel(out, ref))


RESNET_CASES = [
    # (dim, time_stride, maxpool, relu, lengths)
    (40, False, False, None, [50, 50, 50]),           # the shipped resnet18_softmax_1e-2.json, equal lengths
    (40, False, False, None, [33, 61, 33, 40]),       # ragged: equal-length runs become separate batches
    (16, True, True, "prelu", [37, 37]),              # time stride with odd T (asymmetric SAME padding) + max pooling
    (64, False, False, "lrelu", [30]),                # BASELINE config 4's 64-dim fbank
]


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("dim,ts,mp,relu,lengths", RESNET_CASES, ids=["v3", "ragged", "tstride_pool_prelu", "f64_lrelu"])
def test_resnet18_embeddings_match_oracle(precision, dim, ts, mp, relu, lengths):
    import xvector_b200
    from oracle import synth
    cfg = dict(synth.RESNET_CONFIG, resnet_time_stride=ts, resnet_maxpooling=mp)
    if relu:
        cfg["network_relu_type"] = relu
    w = synth.make_resnet_weights(dim, cfg, seed=21)
    lengths = np.asarray(lengths, np.int32)
    feats = synth.fbank_like(int(lengths.sum()), dim, seed=8)
    ref = np.stack(oracle_embeddings(feats, lengths, cfg, w, dim))
    eng = xvector_b200.Engine(cfg, dim, precision=precision)
    try:
        eng.set_weights(w)
        out = eng.forward(feats, lengths)
        pooled = eng.forward(feats, lengths, "pooling")
    finally:
        eng.close()
    ref_pool = np.stack(oracle_embeddings(feats, lengths, cfg, w, dim, node="pooling"))
    assert max_rel(pooled, ref_pool) <= TOL[precision] * 2, "pooling max_rel %.3e" % max_rel(pooled, ref_pool)
    assert max_rel(out, ref) <= TOL[precision], "max_rel %.3e" % max_rel(out, ref)
    assert min_cos(out, ref) >= MIN_COS


def test_resnet18_trainer_predict_dropin():
    """Trainer.predict (model/trainer.py:886-913) with network_type resnet_18: rank-2 and rank-3 inputs."""
    from oracle import numpy_ref, synth
    from xvector_b200.misc.utils import ParamsPlain
    from xvector_b200.model.trainer import Trainer
    cfg = dict(synth.RESNET_CONFIG)
    w = synth.make_resnet_weights(40, cfg, seed=3)
    params = ParamsPlain()
    params.dict.update(cfg)
    tr = Trainer(params, "/nonexistent", 40, single_cpu=True)
    tr.build("predict")
    tr.set_weights(w)
    x = synth.fbank_like(2 * 44, 40, seed=1).reshape(2, 44, 40)
    e3 = tr.predict(x)
    e2 = tr.predict(x[1])
    tr.close()
    ref = numpy_ref.predict(x, numpy_ref.OracleParams(**cfg), w, 40)
    assert e3.shape == (2, 512) and e2.shape == (512,)
    assert max_rel(e3, ref) <= TOL["bf16"]
    np.testing.assert_allclose(e2, e3[1], rtol=0, atol=1e-6 + 1e-3 * np.abs(e3[1]).max())


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_etdnn_embeddings_and_endpoints_match_oracle(precision):
    import xvector_b200
    from oracle import numpy_ref, synth
    cfg = dict(synth.ETDNN_STAT_CONFIG)
    w = synth.make_etdnn_weights(30, cfg, seed=6)
    lengths = np.array([120, 47, 233], np.int32)
    feats = synth.packed_features(lengths, 30, seed=12)
    ref = np.stack(oracle_embeddings(feats, lengths, cfg, w, 30))
    eng = xvector_b200.Engine(cfg, 30, precision=precision)
    try:
        eng.set_weights(w)
        out = eng.forward(feats, lengths)
        assert eng.node_info("tdnn7_relu") == (512, True, 22)
        fr = eng.forward_frames(feats, lengths, "tdnn7_relu")
        out13 = eng.forward(feats, lengths, "tdnn13_relu")
    finally:
        eng.close()
    assert out.shape == (3, 512)
    assert max_rel(out, ref) <= TOL[precision], "max_rel %.3e" % max_rel(out, ref)
    assert min_cos(out, ref) >= MIN_COS
    offs = np.concatenate([[0], np.cumsum(lengths)])
    for u in range(3):
        _, ep = numpy_ref.entire_network(feats[None, offs[u]:offs[u + 1]], numpy_ref.OracleParams(**cfg), w)
        assert fr[u].shape == (lengths[u] - 22, 512)
        assert tensor_rel(fr[u], ep["tdnn7_relu"][0]) <= TOL[precision] * 2
        assert tensor_rel(out13[u], ep["tdnn13_relu"][0]) <= TOL[precision] * 2
