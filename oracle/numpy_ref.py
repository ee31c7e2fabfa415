"""fp64 numpy restatement of the reference x-vector forward path (the parity oracle).

TEST INFRASTRUCTURE -- see oracle/__init__.py for who may import this.

Every function cites the reference lines it follows (paths relative to
/root/reference).  TensorFlow-1.x library defaults the reference relies on but
never states (tf.layers.batch_normalization epsilon=1e-3, axis=-1; conv2d
padding='valid', NHWC, cross-correlation, kernel [kh,kw,in,out]; dense kernel
[in,out]; leaky_relu alpha=0.2) are restated here as constants.

Parity status: attention pooling is pinned by the reference's own
model/test_utils.py:321-392 (see tests/golden/); tdnn / statistics_pooling /
predict / extract are PARITY UNPINNED (no reference test or fixture exists).
"""
from __future__ import annotations

import json
from collections import OrderedDict

import numpy as np

BN_EPSILON = 1e-3           # tf.layers.batch_normalization default (TF 1.x)
LRELU_ALPHA = 0.2           # tf.nn.leaky_relu default
VAR2STD_EPSILON = 1e-12     # model/pooling.py:6


class OracleParams(object):
    """Attribute/dict view of an nnet_conf JSON (misc/utils.py:13-41)."""

    def __init__(self, json_path=None, **kw):
        if json_path is not None:
            with open(json_path) as f:
                self.__dict__.update(json.load(f))
        self.__dict__.update(kw)

    @property
    def dict(self):
        return self.__dict__


# --------------------------------------------------------------------------- layers
def conv_1xk(x, kernel, bias):
    """tf.layers.conv2d(filters,(1,k)) on [B,1,T,C], VALID, stride 1 (model/tdnn.py:42-47).

    x [B,T,C]; kernel [1,k,C,O] (TF layout, cross-correlation); returns [B,T-k+1,O].
    """
    x = np.asarray(x, np.float64)
    w = np.asarray(kernel, np.float64)[0]            # [k,C,O]
    k = w.shape[0]
    t_out = x.shape[1] - k + 1
    if t_out <= 0:
        raise ValueError("utterance shorter than the layer context")
    y = np.zeros((x.shape[0], t_out, w.shape[2]), np.float64)
    for j in range(k):
        y += x[:, j:j + t_out, :] @ w[j]
    return y + np.asarray(bias, np.float64)


def dense_layer(x, kernel, bias):
    """tf.layers.dense: x . W + b on the last axis (model/tdnn.py:99-103)."""
    return np.asarray(x, np.float64) @ np.asarray(kernel, np.float64) + np.asarray(bias, np.float64)


def batch_norm_infer(x, w, scope):
    """tf.layers.batch_normalization(training=False) (model/tdnn.py:49-52)."""
    g = np.asarray(w[scope + "/gamma"], np.float64)
    b = np.asarray(w[scope + "/beta"], np.float64)
    m = np.asarray(w[scope + "/moving_mean"], np.float64)
    v = np.asarray(w[scope + "/moving_variance"], np.float64)
    return g * (x - m) / np.sqrt(v + BN_EPSILON) + b


def activation(x, params, w, scope):
    """relu / leaky_relu / prelu selection (model/tdnn.py:27-33, model/common.py:27-42)."""
    kind = params.dict.get("network_relu_type", None) if "network_relu_type" in params.dict else None
    if kind == "prelu":
        alpha = np.asarray(w[scope + "/alpha"], np.float64)
        return np.maximum(x, 0.0) + alpha * (x - np.abs(x)) * 0.5
    if kind == "lrelu":
        return np.where(x >= 0.0, x, LRELU_ALPHA * x)
    return np.maximum(x, 0.0)


def l2_scaling(x, scaling_factor, epsilon=1e-12):
    """model/common.py:45-58."""
    sq = np.sum(np.square(x), axis=-1, keepdims=True)
    return x * (1.0 / np.sqrt(np.maximum(sq, epsilon))) * scaling_factor


# --------------------------------------------------------------------------- pooling
def statistics_pooling(features):
    """model/pooling.py:40-50: two-pass mean / population variance over axis 1, floor, sqrt."""
    x = np.asarray(features, np.float64)
    mean = np.mean(x, axis=1, keepdims=True)
    var = np.mean(np.square(x - mean), axis=1, keepdims=True)
    mean = mean[:, 0]
    var = var[:, 0]
    mask = (var <= VAR2STD_EPSILON).astype(np.float64)
    var = (1.0 - mask) * var + mask * VAR2STD_EPSILON
    return np.concatenate([mean, np.sqrt(var)], axis=1)


def _dense_block(x, num_nodes, endpoints, params, w, name, scope, kind):
    """dense_bn_relu / dense / dense_relu / dense_tanh (model/common.py:113-223)."""
    base = "%s/%s" % (scope, name)
    x = dense_layer(x, w[base + "/%s_dense/kernel" % name], w[base + "/%s_dense/bias" % name])
    assert x.shape[-1] == num_nodes
    endpoints["%s_dense" % name] = x
    if kind == "bn_relu":
        x = batch_norm_infer(x, w, base + "/%s_bn" % name)
        endpoints["%s_bn" % name] = x
    if kind in ("bn_relu", "relu"):
        x = activation(x, params, w, base + "/%s_relu" % name)
        endpoints["%s_relu" % name] = x
    if kind == "tanh":
        x = np.tanh(x)
        endpoints["%s_tanh" % name] = x
    return x


_LAST_KIND = {0: "linear", 1: "relu", 2: "bn_relu", 3: "tanh"}


def _side_net(x, nodes, last_type, endpoints, params, w, scope, prefix):
    """key / value sub-network (model/pooling.py:100-136)."""
    nodes = list(nodes)
    if len(nodes) == 0:
        return x
    for i, n in enumerate(nodes[:-1]):
        x = _dense_block(x, n, endpoints, params, w, "%s%d" % (prefix, i), scope, "bn_relu")
    return _dense_block(x, nodes[-1], endpoints, params, w, "%s%d" % (prefix, len(nodes) - 1), scope,
                        _LAST_KIND[last_type])


def softmax_lastaxis(x):
    m = np.max(x, axis=-1, keepdims=True)
    e = np.exp(x - m)
    return e / np.sum(e, axis=-1, keepdims=True)


def attention_core(value, key, query, split_value, split_key, use_scale):
    """model/pooling.py:138-218 on plain arrays.

    value [B,L,Cv], key [B,L,Ck], query [H,dk].  Returns (att [B,2*H*dv], weights [B,H,L]).
    """
    value = np.asarray(value, np.float64)
    key = np.asarray(key, np.float64)
    query = np.asarray(query, np.float64)
    b, l, cv = value.shape
    ck = key.shape[2]
    h = query.shape[0]
    if split_value:
        assert cv % h == 0
        v = value.reshape(b, l, h, cv // h).transpose(0, 2, 1, 3)       # [B,H,L,dv]
    else:
        v = np.broadcast_to(value[:, None], (b, h, l, cv))
    if split_key:
        assert ck % h == 0
        k = key.reshape(b, l, h, ck // h).transpose(0, 2, 1, 3)         # [B,H,L,dk]
        qk = np.einsum("bhld,hd->blh", k, query)
    else:
        k = key[:, None]                                                  # [B,1,L,dk]
        qk = np.einsum("bmld,hd->blh", k, query)
    dk = k.shape[-1]
    assert query.shape[1] == dk
    if use_scale:
        qk = qk * (1.0 / np.sqrt(float(dk)))
    weights = softmax_lastaxis(qk.transpose(0, 2, 1))                     # [B,H,L]
    mean = np.einsum("bhld,bhl->bhd", v, weights)
    var = np.einsum("bhld,bhl->bhd", np.square(v - mean[:, :, None, :]), weights)
    mean = mean.reshape(b, -1)
    var = var.reshape(b, -1)
    mask = (var <= VAR2STD_EPSILON).astype(np.float64)
    var = (1.0 - mask) * var + mask * VAR2STD_EPSILON
    return np.concatenate([mean, np.sqrt(var)], axis=1), weights


def self_attention(features, aux_features, endpoints, params, w, scope="tdnn"):
    """model/pooling.py:55-240 (inference; the penalty term :231-238 is training only)."""
    scope = scope + "/attention"
    value = endpoints[params.att_value_input]
    key = endpoints[params.att_key_input]
    key = _side_net(key, params.att_key_num_nodes, params.att_key_network_type, endpoints, params, w, scope, "att_key")
    value = _side_net(value, params.att_value_num_nodes, params.att_value_network_type, endpoints, params, w, scope,
                      "att_value")
    att, weights = attention_core(value, key, w[scope + "/query"], params.att_split_value, params.att_split_key,
                                  params.att_use_scale)
    endpoints["attention_weights"] = weights
    endpoints["att_output_before_nonlinear"] = att
    if params.att_apply_nonlinear:
        att = batch_norm_infer(att, w, scope + "/att_post_bn")
        endpoints["att_post_bn"] = att
        att = activation(att, params, w, scope + "/att_post_relu")
        endpoints["att_post_relu"] = att
    return att


def general_pooling(features, aux_features, endpoints, params, w, scope="tdnn"):
    """model/pooling.py:8-25."""
    if params.pooling_type == "statistics_pooling":
        features = statistics_pooling(features)
    elif params.pooling_type == "self_attention":
        features = self_attention(features, aux_features, endpoints, params, w, scope)
    else:
        raise NotImplementedError("Not implement %s pooling" % params.pooling_type)
    endpoints["pooling"] = features
    return features


# --------------------------------------------------------------------------- networks
def tdnn(features, params, w):
    """model/tdnn.py:10-181, inference.  features [B,T,D] -> (last layer output, endpoints)."""
    ep = OrderedDict()
    x = np.asarray(features, np.float64)
    # model/tdnn.py:38: features are expanded to [b, 1, l, d] for the three conv2d layers, so their endpoints are rank 4;
    # only tdnn3_relu is overwritten with the squeezed rank-3 tensor (:94-96)
    for i in (1, 2, 3):
        x = conv_1xk(x, w["tdnn/tdnn%d_conv/kernel" % i], w["tdnn/tdnn%d_conv/bias" % i])
        ep["tdnn%d_conv" % i] = x[:, None]
        x = batch_norm_infer(x, w, "tdnn/tdnn%d_bn" % i)
        ep["tdnn%d_bn" % i] = x[:, None]
        x = activation(x, params, w, "tdnn/tdnn%d_relu" % i)
        ep["tdnn%d_relu" % i] = x[:, None] if i < 3 else x
    if "num_nodes_pooling_layer" not in params.dict:
        params.dict["num_nodes_pooling_layer"] = 1500
    for i in (4, 5):
        x = dense_layer(x, w["tdnn/tdnn%d_dense/kernel" % i], w["tdnn/tdnn%d_dense/bias" % i])
        ep["tdnn%d_dense" % i] = x
        x = batch_norm_infer(x, w, "tdnn/tdnn%d_bn" % i)
        ep["tdnn%d_bn" % i] = x
        x = activation(x, params, w, "tdnn/tdnn%d_relu" % i)
        ep["tdnn%d_relu" % i] = x
    x = general_pooling(x, None, ep, params, w, "tdnn")
    x = dense_layer(x, w["tdnn/tdnn6_dense/kernel"], w["tdnn/tdnn6_dense/bias"])
    ep["tdnn6_dense"] = x
    x = batch_norm_infer(x, w, "tdnn/tdnn6_bn")
    ep["tdnn6_bn"] = x
    x = activation(x, params, w, "tdnn/tdnn6_relu")
    ep["tdnn6_relu"] = x
    if "num_nodes_last_layer" not in params.dict:
        params.dict["num_nodes_last_layer"] = 512
    x = dense_layer(x, w["tdnn/tdnn7_dense/kernel"], w["tdnn/tdnn7_dense/bias"])
    ep["tdnn7_dense"] = x
    if "last_layer_no_bn" not in params.dict:
        params.last_layer_no_bn = False
    if not params.last_layer_no_bn:
        x = batch_norm_infer(x, w, "tdnn/tdnn7_bn")
        ep["tdnn7_bn"] = x
    if "last_layer_linear" not in params.dict:
        params.last_layer_linear = False
    if not params.last_layer_linear:
        x = activation(x, params, w, "tdnn/tdnn7_relu")
        ep["tdnn7_relu"] = x
    return x, ep


def _tail_layers(x, params, w, scope, ep, first, second):
    """Segment-level layers shared by every network: <first>_dense/bn/relu then <second>_dense[/bn][/relu]
    (model/tdnn.py:137-179, :521-589; model/resnet.py:307-349)."""
    x = dense_layer(x, w["%s/%s_dense/kernel" % (scope, first)], w["%s/%s_dense/bias" % (scope, first)])
    ep["%s_dense" % first] = x
    x = batch_norm_infer(x, w, "%s/%s_bn" % (scope, first))
    ep["%s_bn" % first] = x
    x = activation(x, params, w, "%s/%s_relu" % (scope, first))
    ep["%s_relu" % first] = x
    if "num_nodes_last_layer" not in params.dict:
        params.dict["num_nodes_last_layer"] = 512
    x = dense_layer(x, w["%s/%s_dense/kernel" % (scope, second)], w["%s/%s_dense/bias" % (scope, second)])
    ep["%s_dense" % second] = x
    if "last_layer_no_bn" not in params.dict:
        params.last_layer_no_bn = False
    if not params.last_layer_no_bn:
        x = batch_norm_infer(x, w, "%s/%s_bn" % (scope, second))
        ep["%s_bn" % second] = x
    if "last_layer_linear" not in params.dict:
        params.last_layer_linear = False
    if not params.last_layer_linear:
        x = activation(x, params, w, "%s/%s_relu" % (scope, second))
        ep["%s_relu" % second] = x
    return x


def conv1d_valid(x, kernel, bias):
    """tf.layers.conv1d(filters, k) on [B,T,C], VALID, stride 1 (model/tdnn.py:378-383); kernel [k,C,O]."""
    return conv_1xk(x, np.asarray(kernel)[None], bias)


# frame-level layer table of etdnn (model/tdnn.py:376-518): (name, conv kernel width or 0 for dense)
ETDNN_FRAME_LAYERS = (("tdnn1", 5), ("tdnn2", 0), ("tdnn3", 5), ("tdnn4", 0), ("tdnn5", 7), ("tdnn6", 0),
                      ("tdnn7", 9), ("tdnn8", 0), ("tdnn9", 0), ("tdnn10", 0))


def etdnn(features, params, w):
    """model/tdnn.py:343-591, inference: conv1d k=5,5,7,9 interleaved with 512-dense, tdnn9 dense, tdnn10 dense
    (num_nodes_pooling_layer), pooling, tdnn12, tdnn13.  Variable scope "etdnn"."""
    ep = OrderedDict()
    x = np.asarray(features, np.float64)
    if "num_nodes_pooling_layer" not in params.dict:
        params.dict["num_nodes_pooling_layer"] = 1500
    for name, k in ETDNN_FRAME_LAYERS:
        if k:
            x = conv1d_valid(x, w["etdnn/%s_conv/kernel" % name], w["etdnn/%s_conv/bias" % name])
            ep["%s_conv" % name] = x
        else:
            x = dense_layer(x, w["etdnn/%s_dense/kernel" % name], w["etdnn/%s_dense/bias" % name])
            ep["%s_dense" % name] = x
        x = batch_norm_infer(x, w, "etdnn/%s_bn" % name)
        ep["%s_bn" % name] = x
        x = activation(x, params, w, "etdnn/%s_relu" % name)
        ep["%s_relu" % name] = x
    x = general_pooling(x, None, ep, params, w, "etdnn")
    x = _tail_layers(x, params, w, "etdnn", ep, "tdnn12", "tdnn13")
    return x, ep


# --------------------------------------------------------------------------- ResNet-18
def same_pads(n_in, k, s):
    """TensorFlow 'SAME' padding split: out = ceil(in/s); total = max((out-1)*s + k - in, 0); before = total//2."""
    n_out = -(-n_in // s)
    total = max((n_out - 1) * s + k - n_in, 0)
    return n_out, total // 2, total - total // 2


def conv2d_same(x, kernel, strides=(1, 1), bias=None):
    """tf.layers.conv2d(padding='same', use_bias=False) on NHWC [B,T,F,C] (model/resnet.py:25-32);
    kernel [kh,kw,Cin,Cout], cross-correlation, TF asymmetric SAME padding."""
    x = np.asarray(x, np.float64)
    k = np.asarray(kernel, np.float64)
    kh, kw = k.shape[:2]
    st, sf = strides
    b, t, f, c = x.shape
    to, pt0, pt1 = same_pads(t, kh, st)
    fo, pf0, pf1 = same_pads(f, kw, sf)
    xp = np.zeros((b, t + pt0 + pt1, f + pf0 + pf1, c), np.float64)
    xp[:, pt0:pt0 + t, pf0:pf0 + f] = x
    y = np.zeros((b, to, fo, k.shape[3]), np.float64)
    for dh in range(kh):
        for dw in range(kw):
            y += xp[:, dh:dh + (to - 1) * st + 1:st, dw:dw + (fo - 1) * sf + 1:sf] @ k[dh, dw]
    if bias is not None:
        y = y + np.asarray(bias, np.float64)
    return y


def max_pool_3x3_same(x):
    """tf.layers.max_pooling2d(x, (3,3), (1,1), padding='same') (model/resnet.py:230-231)."""
    b, t, f, c = x.shape
    xp = np.full((b, t + 2, f + 2, c), -np.inf)
    xp[:, 1:t + 1, 1:f + 1] = x
    y = np.full_like(x, -np.inf)
    for dh in range(3):
        for dw in range(3):
            y = np.maximum(y, xp[:, dh:dh + t, dw:dw + f])
    return y


def _res_block(x, filters, strides, params, w, name, project):
    """conv_block (project=True, model/resnet.py:9-87) / identity_block (:90-149) with two 3x3 convs."""
    s = "resnet_18/" + name
    y = conv2d_same(x, w[s + "_conv0/kernel"], strides)
    y = batch_norm_infer(y, w, s + "_bn0")
    y = activation(y, params, w, s + "_relu0")
    y = conv2d_same(y, w[s + "_conv1/kernel"])
    y = batch_norm_infer(y, w, s + "_bn1")
    if project:
        sc = conv2d_same(x, w[s + "_conv_short/kernel"], strides)
        sc = batch_norm_infer(sc, w, s + "_bn_short")
    else:
        sc = x
    return activation(y + sc, params, w, s + "_relu_final")


def resnet_18(features, params, w, require_dim=None):
    """model/resnet.py:152-351, inference.  The reference asserts a 40-dim input (:190); pass require_dim=40 to
    keep that check (BASELINE.json config 4 also names 64-dim fbank, so the default accepts any F % 8 == 0)."""
    x = np.asarray(features, np.float64)
    if require_dim is not None:
        assert x.shape[-1] == require_dim
    assert x.shape[-1] % 8 == 0
    time_stride = 2 if params.resnet_time_stride else 1
    if "resnet_blocks" not in params.dict:
        params.dict["resnet_blocks"] = [2, 2, 2, 2]
    ep = OrderedDict()
    x = x[:, :, :, None]
    x = conv2d_same(x, w["resnet_18/conv0_1/kernel"])
    x = batch_norm_infer(x, w, "resnet_18/conv0_bn")
    x = activation(x, params, w, "resnet_18/conv0_relu")
    if params.resnet_maxpooling:
        x = max_pool_3x3_same(x)
    for stage, filt in enumerate((64, 128, 256, 512)):
        strides = (1, 1) if stage == 0 else (time_stride, 2)
        x = _res_block(x, filt, strides, params, w, "conv%da" % (stage + 1), True)
        for i in range(params.resnet_blocks[stage] - 1):
            x = _res_block(x, filt, (1, 1), params, w, "conv%db_%d" % (stage + 1, i), False)
    # conv5: (1, F') VALID with bias (model/resnet.py:256-267)
    k5 = np.asarray(w["resnet_18/conv5/kernel"], np.float64)          # [1, F', 512, 512]
    assert k5.shape[1] == x.shape[2]
    x = np.einsum("btfc,fco->bto", x, k5[0]) + np.asarray(w["resnet_18/conv5/bias"], np.float64)
    x = batch_norm_infer(x, w, "resnet_18/conv5_bn")
    x = activation(x, params, w, "resnet_18/conv5_relu")
    for name in ("dense1", "dense2"):
        x = dense_layer(x, w["resnet_18/%s/kernel" % name], w["resnet_18/%s/bias" % name])
        x = batch_norm_infer(x, w, "resnet_18/%s_bn" % name)
        x = activation(x, params, w, "resnet_18/%s_relu" % name)
    x = general_pooling(x, None, ep, params, w, "resnet_18")
    x = _tail_layers(x, params, w, "resnet_18", ep, "tdnn6", "tdnn7")
    return x, ep


NETWORKS = {"tdnn": tdnn, "extended_tdnn": etdnn, "resnet_18": resnet_18}      # model/trainer.py:101-108


def entire_network(features, params, w):
    """model/trainer.py:385-405."""
    if params.network_type not in NETWORKS:
        raise NotImplementedError("Not implement %s network" % params.network_type)
    x, ep = NETWORKS[params.network_type](features, params, w)
    ep["output"] = x
    if "feature_norm" in params.dict and params.feature_norm:
        assert "feature_scaling_factor" in params.dict
        x = l2_scaling(x, params.feature_scaling_factor)
        ep["output"] = x
    return x, ep


def predict(features, params, w, dim, node=None):
    """model/trainer.py:886-913 (rank-2/3 handling, slicing to dim, node selection :379-380)."""
    features = np.asarray(features)
    rank = features.ndim
    assert rank == 2 or rank == 3
    if rank == 2:
        features = features[None]
    if features.shape[-1] != dim:
        features = features[:, :, :dim]
    _, ep = entire_network(features, params, w)
    out = ep[node if node else params.embedding_node]
    if rank == 2:
        out = out[0]
    return out


# --------------------------------------------------------------------------- extractor
def chunk_plan(length, chunk_size):
    """egs/voxceleb/v1/nnet/lib/extract.py:68-78 with Python-2 integer division.

    Returns [(start, size), ...] for an utterance longer than chunk_size.
    """
    half = chunk_size // 2                       # `args.chunk_size / 2` under Python 2
    num_chunks = int(np.ceil(float(length - chunk_size) / half)) + 1
    plan = []
    for i in range(num_chunks):
        start = i * half
        size = chunk_size if length - start > chunk_size else length - start
        plan.append((start, size))
    return plan


def extract_embedding(feature, predict_fn, chunk_size=10000, min_chunk_size=25, normalize=False):
    """extract.py:64-93 for one utterance; returns None when the utterance is skipped."""
    feature = np.asarray(feature)
    n = feature.shape[0]
    if n < min_chunk_size:
        return None
    if n > chunk_size:
        plan = chunk_plan(n, chunk_size)
        chunks = [feature[s:s + z] for s, z in plan]
        lengths = np.array([z for _, z in plan], np.float64)[:, None]
        emb = predict_fn(np.array(chunks[:-1], dtype=np.float32))
        emb_last = predict_fn(chunks[-1])
        emb = np.concatenate([emb, emb_last[None]], axis=0)
        if normalize:
            emb = emb / np.sqrt(np.sum(np.square(emb), axis=1, keepdims=True))
        embedding = np.sum(emb * lengths, axis=0) / np.sum(lengths)
    else:
        embedding = predict_fn(feature)
    if normalize:
        embedding = embedding / np.sqrt(np.sum(np.square(embedding)))
    return embedding


# --------------------------------------------------------------------------- work model
def tdnn_flops(t, d, pool_nodes=1500):
    """algorithmic FLOPs of one utterance up to tdnn6_dense."""
    return (2 * 512 * (5 * d * (t - 4) + 2560 * (t - 8) + 3584 * (t - 14) + 512 * (t - 14))
            + 2 * 512 * pool_nodes * (t - 14) + 2 * (2 * pool_nodes) * 512)
