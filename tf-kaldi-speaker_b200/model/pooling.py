"""Pooling layers with the reference's signatures (model/pooling.py:8-240) on numpy arrays; arithmetic on the GPU.

When `endpoints` is the lazy Endpoints object produced by this package's tdnn() (the way the reference calls
general_pooling from inside tdnn(), model/tdnn.py:133) the fused engine path is used; for plain arrays the
standalone C-ABI entry points xv_stat_pooling / xv_attention_pooling run the same reductions."""
import numpy as np

from . import variables as V
from .common import dense, dense_bn_relu, dense_relu, dense_tanh, prelu, shape_list  # noqa: F401

VAR2STD_EPSILON = 1e-12


def general_pooling(features, aux_features, endpoints, params, is_training):
    """model/pooling.py:8-25."""
    if params.pooling_type == "statistics_pooling":
        features = statistics_pooling(features, aux_features, endpoints, params, is_training)
    elif params.pooling_type == "self_attention":
        features = self_attention(features, aux_features, endpoints, params, is_training)
    else:
        raise NotImplementedError("Not implement %s pooling" % params.pooling_type)
    endpoints['pooling'] = features
    return features


def statistics_pooling(features, aux_features, endpoints, params, is_training):
    """model/pooling.py:27-52: [batch, length, dim] -> [batch, 2*dim] = [mean | stddev] (population variance,
    floored at 1e-12, no length mask -- every row of `features` counts)."""
    x = np.asarray(features, np.float32)
    assert x.ndim == 3, "statistics_pooling expects [batch, length, dim]"
    b, l, c = x.shape
    cu = np.arange(b + 1, dtype=np.int32) * l
    return V.utility_engine().stat_pooling(x.reshape(b * l, c), cu)


def statistics_pooling_ragged(rows, cu_lens, dtype="fp32"):
    """Extension: masked mean/std over variable-length utterances packed back to back ([rows, dim], cu_lens[n+1]).
    Equals the reference's statistics_pooling applied to each utterance on its own."""
    return V.utility_engine().stat_pooling(rows, cu_lens, dtype=dtype)


def self_attention(features, aux_features, endpoints, params, is_training=None):
    """model/pooling.py:55-240 (inference; the penalty term is training-only).  Key and value are taken from
    `endpoints` by name, not from `features` (model/pooling.py:97-98)."""
    value_features = np.asarray(endpoints[params.att_value_input], np.float32)
    key_features = np.asarray(endpoints[params.att_key_input], np.float32)
    with V.variable_scope("attention"):
        def side(x, nodes, last_type, prefix):
            nodes = list(nodes)
            if len(nodes) > 1:
                for index, num_nodes in enumerate(nodes[:-1]):
                    x = dense_bn_relu(x, num_nodes, endpoints, params, is_training, name=("%s%d" % (prefix, index)))
            last = "%s%d" % (prefix, len(nodes) - 1)
            if last_type == 0:
                x = dense(x, nodes[-1], endpoints, params, is_training, name=last)
            elif last_type == 1:
                x = dense_relu(x, nodes[-1], endpoints, params, is_training, name=last)
            elif last_type == 2:
                x = dense_bn_relu(x, nodes[-1], endpoints, params, is_training, name=last)
            elif last_type == 3:
                x = dense_tanh(x, nodes[-1], endpoints, params, is_training, name=last)
            return x

        key_features = side(key_features, params.att_key_num_nodes, params.att_key_network_type, "att_key")
        if len(params.att_value_num_nodes) > 0:
            value_features = side(value_features, params.att_value_num_nodes, params.att_value_network_type, "att_value")
        n_heads = params.att_num_heads
        if params.att_split_value:
            assert shape_list(value_features)[2] % n_heads == 0, "The dim of the value must be divided by the num of heads."
        if params.att_split_key:
            assert shape_list(key_features)[2] % n_heads == 0
        query = np.asarray(V.get_variable("query"), np.float32)
        b, l, cv = value_features.shape
        cu = np.arange(b + 1, dtype=np.int32) * l
        att, weights = V.utility_engine().attention_pooling(
            value_features.reshape(b * l, cv), key_features.reshape(b * l, -1), query, cu,
            params.att_split_value, params.att_split_key, params.att_use_scale, want_weights=True)
        endpoints["attention_weights"] = np.stack(weights, axis=0)        # [b, h, l]
        endpoints["att_output_before_nonlinear"] = att
        if params.att_apply_nonlinear:
            from .common import _fold_bn, _slope
            s, h = _fold_bn("att_post_bn")
            eye = np.eye(att.shape[1], dtype=np.float32)                 # BN as a diagonal affine through the dense kernel
            zero = np.zeros(att.shape[1], np.float32)
            eng = V.utility_engine()
            bn = eng.dense_rows(att, eye, zero, bn_scale=s, bn_shift=h)
            endpoints["att_post_bn"] = bn
            att = eng.dense_rows(att, eye, zero, bn_scale=s, bn_shift=h, slope=_slope(params, "att_post_relu", att.shape[1]))
            endpoints["att_post_relu"] = att
    return att
