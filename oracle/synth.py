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
