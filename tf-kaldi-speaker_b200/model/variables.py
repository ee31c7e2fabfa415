"""Stand-in for TensorFlow's variable store.  The reference's layer functions create/read variables through
tf.get_variable under tf.variable_scope (model/tdnn.py:36, model/pooling.py:96, model/common.py:132-142); here the
"graph state" is a dict {TF variable name: ndarray} installed with `variable_store(weights)` plus a scope stack,
and a cache of engines (one xv_ctx per (config, dim, precision, weights))."""
import collections
import contextlib
import json
import zlib

import numpy as np

_state = {"weights": None, "scopes": [], "precision": "bf16", "device": 0}
_engines = collections.OrderedDict()      # LRU: least recently used first
MAX_CACHED_ENGINES = 8                    # every engine holds a device copy of its weights; evicted contexts are closed


@contextlib.contextmanager
def variable_store(weights, precision=None, device=None):
    """Install `weights` ({tf variable name: array}) as the variables the layer functions read."""
    old = dict(_state)
    _state["weights"] = weights
    if precision is not None:
        _state["precision"] = precision
    if device is not None:
        _state["device"] = device
    try:
        yield
    finally:
        _state.update(old)


def set_variables(weights, precision=None, device=None):
    _state["weights"] = weights
    if precision is not None:
        _state["precision"] = precision
    if device is not None:
        _state["device"] = device


def current_weights():
    if _state["weights"] is None:
        raise RuntimeError("no variables installed: use model.variables.variable_store(weights) "
                           "(the TF checkpoint restore of the reference is replaced by an explicit weight dict)")
    return _state["weights"]


def precision():
    return _state["precision"]


def device():
    return _state["device"]


@contextlib.contextmanager
def variable_scope(name):
    _state["scopes"].append(name)
    try:
        yield
    finally:
        _state["scopes"].pop()


def scope_prefix():
    return "".join(s + "/" for s in _state["scopes"])


def get_variable(name):
    full = scope_prefix() + name
    w = current_weights()
    if full not in w:
        raise KeyError("variable '%s' not found in the variable store" % full)
    return w[full]


def _config_key(params_dict):
    clean = {}
    for k, v in params_dict.items():
        try:
            json.dumps(v)
        except (TypeError, ValueError):
            continue
        clean[k] = v
    return json.dumps(clean, sort_keys=True)


def weights_fingerprint(weights):
    """Content fingerprint of a variable dict (names, shapes, CRC-32 of the values): a dict that is mutated in place, or
    a fresh dict with the same content, maps to the right engine (~20 ms for the 4.2 M parameters of the TDNN)."""
    crc = 0
    for name in sorted(weights):
        a = np.ascontiguousarray(weights[name])
        crc = zlib.crc32(("%s%s%s" % (name, a.shape, a.dtype)).encode("utf-8"), crc)
        crc = zlib.crc32(a.view(np.uint8).reshape(-1), crc)
    return crc


def _remember(key, eng):
    _engines[key] = eng
    _engines.move_to_end(key)
    while len(_engines) > MAX_CACHED_ENGINES:
        _, old = _engines.popitem(last=False)
        old.close()


def get_engine(params, dim, weights=None, prec=None, dev=None, config=None):
    """Engine for (config, dim, precision, device, weight CONTENT); built and weight-loaded on first use, kept in an LRU
    of MAX_CACHED_ENGINES contexts (the evicted context is closed and its device memory freed).
    `config` overrides params.dict (the network functions force their own network_type)."""
    from ..engine import Engine
    weights = current_weights() if weights is None else weights
    prec = prec or precision()
    dev = device() if dev is None else dev
    cfg = params.dict if config is None else config
    key = (_config_key(cfg), int(dim), prec, dev, weights_fingerprint(weights))
    eng = _engines.get(key)
    if eng is None:
        eng = Engine(cfg, dim, precision=prec, device=dev)
        eng.set_weights(weights)
    _remember(key, eng)
    return eng


def utility_engine():
    """A context with no model of its own, for the standalone pooling / dense entry points."""
    from ..engine import Engine
    key = ("__utility__", device())
    eng = _engines.get(key)
    if eng is None:
        eng = Engine({"network_type": "tdnn", "pooling_type": "statistics_pooling"}, 1, precision="fp32", device=device())
    _remember(key, eng)
    return eng


def clear_engines():
    for e in _engines.values():
        e.close()
    _engines.clear()


# ------------------------------------------------------------------------------------------------ initial values
def initial_variables(params, dim, num_speakers, seed=0, hidden=512):
    """What sess.run(tf.global_variables_initializer()) leaves in the graph of Trainer.build("train")
    (model/trainer.py:601) for network_type "tdnn" with statistics pooling, as {TF variable name: array}:
    tf.layers.conv2d / dense kernels glorot-uniform (the tf.layers default), biases 0; batch norm gamma 1, beta 0,
    moving_mean 0, moving_variance 1; PReLU alpha 0.01 (model/common.py:27-42); softmax/output/kernel
    tf.contrib.layers.xavier_initializer (uniform), softmax/output/bias 0 (model/loss.py:29-34, 236-239).
    The random stream is numpy's (TensorFlow's is not reproducible outside TensorFlow): same distributions, other draws."""
    rng = np.random.default_rng(seed)
    d = params.dict if hasattr(params, "dict") else dict(params)
    pool_nodes = d.get("num_nodes_pooling_layer", 1500)
    last_nodes = d.get("num_nodes_last_layer", 512)
    prelu = d.get("network_relu_type", None) == "prelu"
    w = {}

    def glorot(shape, fan_in, fan_out):
        lim = np.sqrt(6.0 / (fan_in + fan_out))
        return rng.uniform(-lim, lim, size=shape).astype(np.float32)

    def bn(scope, n):
        w[scope + "/gamma"] = np.ones(n, np.float32)
        w[scope + "/beta"] = np.zeros(n, np.float32)
        w[scope + "/moving_mean"] = np.zeros(n, np.float32)
        w[scope + "/moving_variance"] = np.ones(n, np.float32)

    def alpha(name, n):
        if prelu:
            w["tdnn/%s_relu/alpha" % name] = np.full(n, 0.01, np.float32)

    cin = dim
    for i, k in ((1, 5), (2, 5), (3, 7)):
        w["tdnn/tdnn%d_conv/kernel" % i] = glorot((1, k, cin, hidden), k * cin, k * hidden)
        w["tdnn/tdnn%d_conv/bias" % i] = np.zeros(hidden, np.float32)
        bn("tdnn/tdnn%d_bn" % i, hidden)
        alpha("tdnn%d" % i, hidden)
        cin = hidden
    for i, n_in, n_out in ((4, hidden, hidden), (5, hidden, pool_nodes), (6, 2 * pool_nodes, 512), (7, 512, last_nodes)):
        w["tdnn/tdnn%d_dense/kernel" % i] = glorot((n_in, n_out), n_in, n_out)
        w["tdnn/tdnn%d_dense/bias" % i] = np.zeros(n_out, np.float32)
        if i < 7 or not d.get("last_layer_no_bn", False):
            bn("tdnn/tdnn%d_bn" % i, n_out)
        if i < 7 or not d.get("last_layer_linear", False):
            alpha("tdnn%d" % i, n_out)
    w["softmax/output/kernel"] = glorot((last_nodes, num_speakers), last_nodes, num_speakers)
    if d.get("loss_func", "softmax") == "softmax":
        w["softmax/output/bias"] = np.zeros(num_speakers, np.float32)
    return w
