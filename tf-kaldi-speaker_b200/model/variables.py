"""Stand-in for TensorFlow's variable store.  The reference's layer functions create/read variables through
tf.get_variable under tf.variable_scope (model/tdnn.py:36, model/pooling.py:96, model/common.py:132-142); here the
"graph state" is a dict {TF variable name: ndarray} installed with `variable_store(weights)` plus a scope stack,
and a cache of engines (one xv_ctx per (config, dim, precision, weights))."""
import contextlib
import json

_state = {"weights": None, "scopes": [], "precision": "bf16", "device": 0}
_engines = {}


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


def get_engine(params, dim, weights=None, prec=None, dev=None, config=None):
    """Engine for (config, dim, precision, weight dict identity); built and weight-loaded on first use.
    `config` overrides params.dict (the network functions force their own network_type)."""
    from ..engine import Engine
    weights = current_weights() if weights is None else weights
    prec = prec or precision()
    dev = device() if dev is None else dev
    cfg = params.dict if config is None else config
    key = (_config_key(cfg), int(dim), prec, dev, id(weights))
    eng = _engines.get(key)
    if eng is None:
        eng = Engine(cfg, dim, precision=prec, device=dev)
        eng.set_weights(weights)
        eng._weights_ref = weights          # keep the dict alive so id() stays unique
        _engines[key] = eng
    return eng


def utility_engine():
    """A context with no model of its own, for the standalone pooling / dense entry points."""
    from ..engine import Engine
    key = ("__utility__", device())
    eng = _engines.get(key)
    if eng is None:
        eng = Engine({"network_type": "tdnn", "pooling_type": "statistics_pooling"}, 1, precision="fp32", device=device())
        _engines[key] = eng
    return eng


def clear_engines():
    for e in _engines.values():
        e.close()
    _engines.clear()
