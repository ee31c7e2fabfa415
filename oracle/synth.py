"""Seeded synthetic weights / features / configs.

TEST INFRASTRUCTURE.  No pretrained weights or feature files ship with the
reference (README.md:101-119), so every parity and bench run uses these
generators.  Weights are keyed by TF variable name (model/tdnn.py:36 scope
"tdnn"; model/pooling.py:96 scope "attention"; model/common.py:132-142).
Pure numpy; importable anywhere (bench.py uses it to make inputs, which is not
"executing the oracle": nothing here computes the path).
"""
from __future__ import annotations

import numpy as np

# nnet_conf JSONs restated as dicts (egs/voxceleb/v1/nnet_conf/tdnn_softmax_1e-2.json and
# tdnn_softmax_1e-2_tdnn4_att_pretrain.json: only the keys the forward path reads).
TDNN_STAT_CONFIG = {
    "network_type": "tdnn",
    "last_layer_linear": False,
    "loss_func": "softmax",
    "pooling_type": "statistics_pooling",
    "embedding_node": "tdnn6_dense",
    "weight_l2_regularizer": 1e-2,
    "batchnorm_momentum": 0.99,
}

TDNN_ATT_CONFIG = {
    "network_type": "tdnn",
    "last_layer_linear": False,
    "loss_func": "softmax",
    "num_nodes_pooling_layer": 1500,
    "pooling_type": "self_attention",
    "embedding_node": "tdnn6_dense",
    "att_key_input": "tdnn4_relu",
    "att_key_num_nodes": [1500, 1500],
    "att_key_network_type": 1,
    "att_value_input": "tdnn5_relu",
    "att_value_num_nodes": [],
    "att_value_network_type": 0,
    "att_apply_nonlinear": False,
    "att_use_scale": True,
    "att_num_heads": 1,
    "att_split_value": True,
    "att_split_key": True,
    "att_penalty_term": 0,
    "weight_l2_regularizer": 1e-2,
    "batchnorm_momentum": 0.99,
}


def _glorot(rng, shape, fan_in, fan_out):
    lim = np.sqrt(6.0 / (fan_in + fan_out))
    return rng.uniform(-lim, lim, size=shape).astype(np.float32)


def _bn(rng, w, scope, n):
    w[scope + "/gamma"] = rng.uniform(0.5, 1.5, n).astype(np.float32)
    w[scope + "/beta"] = (0.1 * rng.standard_normal(n)).astype(np.float32)
    w[scope + "/moving_mean"] = (0.1 * rng.standard_normal(n)).astype(np.float32)
    w[scope + "/moving_variance"] = rng.uniform(0.5, 1.5, n).astype(np.float32)


def _dense(rng, w, scope, n_in, n_out):
    w[scope + "/kernel"] = _glorot(rng, (n_in, n_out), n_in, n_out)
    w[scope + "/bias"] = (0.1 * rng.standard_normal(n_out)).astype(np.float32)


def make_tdnn_weights(dim, config=None, seed=1234, hidden=512):
    """All variables tdnn() creates for `config` (a dict of nnet_conf keys)."""
    cfg = dict(TDNN_STAT_CONFIG if config is None else config)
    rng = np.random.default_rng(seed)
    w = {}
    pool_nodes = cfg.get("num_nodes_pooling_layer", 1500)
    last_nodes = cfg.get("num_nodes_last_layer", 512)
    prelu = cfg.get("network_relu_type", None) == "prelu"
    cin = dim
    for i, k in ((1, 5), (2, 5), (3, 7)):
        s = "tdnn/tdnn%d_conv" % i
        w[s + "/kernel"] = _glorot(rng, (1, k, cin, hidden), k * cin, k * hidden)
        w[s + "/bias"] = (0.1 * rng.standard_normal(hidden)).astype(np.float32)
        _bn(rng, w, "tdnn/tdnn%d_bn" % i, hidden)
        cin = hidden
    _dense(rng, w, "tdnn/tdnn4_dense", hidden, hidden)
    _bn(rng, w, "tdnn/tdnn4_bn", hidden)
    _dense(rng, w, "tdnn/tdnn5_dense", hidden, pool_nodes)
    _bn(rng, w, "tdnn/tdnn5_bn", pool_nodes)
    widths = {"tdnn1": hidden, "tdnn2": hidden, "tdnn3": hidden, "tdnn4": hidden, "tdnn5": pool_nodes}
    pooled = 2 * pool_nodes
    if cfg["pooling_type"] == "self_attention":
        def side(prefix, src, nodes, last_type):
            n_in = widths[src.split("_")[0]]
            nodes = list(nodes)
            for j, n in enumerate(nodes):
                name = "%s%d" % (prefix, j)
                base = "tdnn/attention/%s" % name
                _dense(rng, w, base + "/%s_dense" % name, n_in, n)
                last = j == len(nodes) - 1
                if (not last) or last_type == 2:
                    _bn(rng, w, base + "/%s_bn" % name, n)
                if prelu and ((not last) or last_type in (1, 2)):
                    w[base + "/%s_relu/alpha" % name] = rng.uniform(0.0, 0.3, n).astype(np.float32)
                n_in = n
            return n_in
        ck = side("att_key", cfg["att_key_input"], cfg["att_key_num_nodes"], cfg["att_key_network_type"])
        cv = side("att_value", cfg["att_value_input"], cfg["att_value_num_nodes"], cfg["att_value_network_type"])
        heads = cfg["att_num_heads"]
        dk = ck // heads if cfg["att_split_key"] else ck
        # truncated_normal(stddev=0.1): resample beyond 2 sigma (model/pooling.py:182-183)
        q = rng.standard_normal((heads, dk))
        bad = np.abs(q) > 2.0
        while bad.any():
            q[bad] = rng.standard_normal(int(bad.sum()))
            bad = np.abs(q) > 2.0
        w["tdnn/attention/query"] = (0.1 * q).astype(np.float32)
        dv = cv // heads if cfg["att_split_value"] else cv
        pooled = 2 * heads * dv
        if cfg.get("att_apply_nonlinear", False):
            _bn(rng, w, "tdnn/attention/att_post_bn", pooled)
            if prelu:
                w["tdnn/attention/att_post_relu/alpha"] = rng.uniform(0.0, 0.3, pooled).astype(np.float32)
    _dense(rng, w, "tdnn/tdnn6_dense", pooled, 512)
    _bn(rng, w, "tdnn/tdnn6_bn", 512)
    _dense(rng, w, "tdnn/tdnn7_dense", 512, last_nodes)
    _bn(rng, w, "tdnn/tdnn7_bn", last_nodes)
    if prelu:
        for i, n in ((1, hidden), (2, hidden), (3, hidden), (4, hidden), (5, pool_nodes), (6, 512), (7, last_nodes)):
            w["tdnn/tdnn%d_relu/alpha" % i] = rng.uniform(0.0, 0.3, n).astype(np.float32)
    return w


def mfcc_like(rng, n_frames, dim):
    """CMVN'd-MFCC-like rows: zero mean, scale decaying 3.0 -> 0.5 over the cepstral index."""
    scale = np.linspace(3.0, 0.5, dim).astype(np.float32)
    return rng.standard_normal((n_frames, dim), dtype=np.float32) * scale


def config1_features(n_utts=1000, frames=300, dim=30, seed=101):
    """cfg 1: [n_utts, frames, dim] fp32."""
    rng = np.random.default_rng(seed)
    return mfcc_like(rng, n_utts * frames, dim).reshape(n_utts, frames, dim)


def config2_lengths(n_utts=4096, lo=200, hi=6000, seed=202):
    """cfg 2: log-uniform utterance lengths in [lo, hi]."""
    rng = np.random.default_rng(seed)
    return np.rint(np.exp(rng.uniform(np.log(lo), np.log(hi), n_utts))).astype(np.int32)


def packed_features(lengths, dim=30, seed=202):
    """Ragged utterances packed back to back: [sum(lengths), dim] fp32."""
    rng = np.random.default_rng(seed + 1)
    return mfcc_like(rng, int(np.sum(lengths)), dim)


# ----------------------------------------------------------------------------- etdnn / resnet_18
ETDNN_STAT_CONFIG = dict(TDNN_STAT_CONFIG, network_type="extended_tdnn", embedding_node="tdnn12_dense")

# egs/voxceleb/v3/nnet_conf/resnet18_softmax_1e-2.json (keys the forward path reads); v3/run.sh extracts
# --node tdnn6_dense.
RESNET_CONFIG = {
    "network_type": "resnet_18",
    "resnet_time_stride": False,
    "resnet_maxpooling": False,
    "last_layer_linear": False,
    "loss_func": "softmax",
    "pooling_type": "statistics_pooling",
    "embedding_node": "tdnn6_dense",
    "weight_l2_regularizer": 1e-2,
    "batchnorm_momentum": 0.99,
}

ETDNN_LAYERS = (("tdnn1", 5), ("tdnn2", 0), ("tdnn3", 5), ("tdnn4", 0), ("tdnn5", 7), ("tdnn6", 0), ("tdnn7", 9),
                ("tdnn8", 0), ("tdnn9", 0), ("tdnn10", 0))


def make_etdnn_weights(dim, config=None, seed=1234, hidden=512):
    """All variables etdnn() creates (model/tdnn.py:343-591; conv1d kernels are [k, Cin, Cout])."""
    cfg = dict(ETDNN_STAT_CONFIG if config is None else config)
    rng = np.random.default_rng(seed)
    w = {}
    pool_nodes = cfg.get("num_nodes_pooling_layer", 1500)
    last_nodes = cfg.get("num_nodes_last_layer", 512)
    prelu = cfg.get("network_relu_type", None) == "prelu"
    cin = dim
    for name, k in ETDNN_LAYERS:
        cout = pool_nodes if name == "tdnn10" else hidden
        if k:
            w["etdnn/%s_conv/kernel" % name] = _glorot(rng, (k, cin, cout), k * cin, k * cout)
            w["etdnn/%s_conv/bias" % name] = (0.1 * rng.standard_normal(cout)).astype(np.float32)
        else:
            _dense(rng, w, "etdnn/%s_dense" % name, cin, cout)
        _bn(rng, w, "etdnn/%s_bn" % name, cout)
        if prelu:
            w["etdnn/%s_relu/alpha" % name] = rng.uniform(0.0, 0.3, cout).astype(np.float32)
        cin = cout
    _dense(rng, w, "etdnn/tdnn12_dense", 2 * pool_nodes, 512)
    _bn(rng, w, "etdnn/tdnn12_bn", 512)
    _dense(rng, w, "etdnn/tdnn13_dense", 512, last_nodes)
    _bn(rng, w, "etdnn/tdnn13_bn", last_nodes)
    if prelu:
        w["etdnn/tdnn12_relu/alpha"] = rng.uniform(0.0, 0.3, 512).astype(np.float32)
        w["etdnn/tdnn13_relu/alpha"] = rng.uniform(0.0, 0.3, last_nodes).astype(np.float32)
    return w


def make_resnet_weights(dim, config=None, seed=1234):
    """All variables resnet_18() creates (model/resnet.py:209-349): 3x3 kernels [3,3,Cin,Cout] without bias,
    1x1 projection shortcuts, conv5 [1, dim/8, 512, 512] with bias, dense1/dense2, tdnn6/tdnn7."""
    cfg = dict(RESNET_CONFIG if config is None else config)
    rng = np.random.default_rng(seed)
    w = {}
    prelu = cfg.get("network_relu_type", None) == "prelu"
    blocks = cfg.get("resnet_blocks", [2, 2, 2, 2])
    last_nodes = cfg.get("num_nodes_last_layer", 512)

    def conv(name, kh, kw, cin, cout):
        w["resnet_18/%s/kernel" % name] = _glorot(rng, (kh, kw, cin, cout), kh * kw * cin, kh * kw * cout)

    def alpha(name, n):
        if prelu:
            w["resnet_18/%s/alpha" % name] = rng.uniform(0.0, 0.3, n).astype(np.float32)

    conv("conv0_1", 3, 3, 1, 64)
    _bn(rng, w, "resnet_18/conv0_bn", 64)
    alpha("conv0_relu", 64)
    cin = 64
    for stage, filt in enumerate((64, 128, 256, 512)):
        names = ["conv%da" % (stage + 1)] + ["conv%db_%d" % (stage + 1, i) for i in range(blocks[stage] - 1)]
        for bi, name in enumerate(names):
            conv(name + "_conv0", 3, 3, cin, filt)
            _bn(rng, w, "resnet_18/%s_bn0" % name, filt)
            alpha(name + "_relu0", filt)
            conv(name + "_conv1", 3, 3, filt, filt)
            _bn(rng, w, "resnet_18/%s_bn1" % name, filt)
            if bi == 0:
                conv(name + "_conv_short", 1, 1, cin, filt)
                _bn(rng, w, "resnet_18/%s_bn_short" % name, filt)
            alpha(name + "_relu_final", filt)
            cin = filt
    fq = dim // 8
    conv("conv5", 1, fq, 512, 512)
    w["resnet_18/conv5/bias"] = (0.1 * rng.standard_normal(512)).astype(np.float32)
    _bn(rng, w, "resnet_18/conv5_bn", 512)
    alpha("conv5_relu", 512)
    _dense(rng, w, "resnet_18/dense1", 512, 512)
    _bn(rng, w, "resnet_18/dense1_bn", 512)
    alpha("dense1_relu", 512)
    _dense(rng, w, "resnet_18/dense2", 512, 1500)
    _bn(rng, w, "resnet_18/dense2_bn", 1500)
    alpha("dense2_relu", 1500)
    _dense(rng, w, "resnet_18/tdnn6_dense", 3000, 512)
    _bn(rng, w, "resnet_18/tdnn6_bn", 512)
    alpha("tdnn6_relu", 512)
    _dense(rng, w, "resnet_18/tdnn7_dense", 512, last_nodes)
    _bn(rng, w, "resnet_18/tdnn7_bn", last_nodes)
    alpha("tdnn7_relu", last_nodes)
    return w


def fbank_like(n_frames, dim, seed):
    """CMN'd log-mel-like rows: N(0,1) (SURVEY.md 8(d) cfg 3/4)."""
    return np.random.default_rng(seed).standard_normal((n_frames, dim), dtype=np.float32)


# ----------------------------------------------------------------------------- training step (config 5)
# egs/voxceleb/v1/nnet_conf/tdnn_amsoftmax_m0.20_linear_bn_1e-2.json (keys the training step reads)
TDNN_AMSOFTMAX_CONFIG = {
    "network_type": "tdnn",
    "last_layer_no_bn": False,
    "last_layer_linear": True,
    "feature_norm": False,
    "loss_func": "additive_margin_softmax",
    "amsoftmax_m": 0.20,
    "amsoftmax_lambda_min": 0,
    "amsoftmax_lambda_base": 1000,
    "amsoftmax_lambda_gamma": 0.0001,
    "amsoftmax_lambda_power": 5,
    "pooling_type": "statistics_pooling",
    "embedding_node": "tdnn6_dense",
    "learning_rate": 0.01,
    "clip_gradient": False,
    "clip_gradient_norm": 3,
    "weight_l2_regularizer": 1e-2,
    "batchnorm_momentum": 0.99,
    "num_speakers_per_batch": 64,
    "num_segments_per_speaker": 1,
}


def add_softmax_weights(w, num_speakers, embed_dim=512, seed=77, bias=False):
    """softmax/output/kernel [E, S], Xavier-uniform (model/loss.py:236-239) (+ bias for plain softmax :29-34)."""
    rng = np.random.default_rng(seed)
    w = dict(w)
    w["softmax/output/kernel"] = _glorot(rng, (embed_dim, num_speakers), embed_dim, num_speakers)
    if bias:
        w["softmax/output/bias"] = np.zeros(num_speakers, np.float32)
    return w


def config5_batch(batch=64, frames=200, dim=30, num_speakers=5994, seed=505, speaker_signal=0.0):
    """cfg 5: features [batch, frames, dim] fp32 + labels [batch] int32.  speaker_signal > 0 adds a
    per-speaker cepstral offset (so that utterance-level statistics differ between speakers the way real data does;
    with 0 every utterance is the same noise process and the segment-level batch norms amplify rounding noise)."""
    rng = np.random.default_rng(seed)
    x = mfcc_like(rng, batch * frames, dim).reshape(batch, frames, dim)
    y = rng.integers(0, num_speakers, batch).astype(np.int32)
    if speaker_signal:
        spk = np.random.default_rng(seed + 7919).standard_normal((num_speakers, dim)).astype(np.float32)
        x = x + np.float32(speaker_signal) * spk[y][:, None, :] * x.std(axis=(0, 1), keepdims=True)
    return x.astype(np.float32), y
