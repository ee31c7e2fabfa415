"""tdnn() with the reference's signature and endpoint names (model/tdnn.py:10-181), executed by libxvector.

There is no TF graph: `features` is a numpy array [batch, length, dim]; the variables tdnn() would create/read
are taken from the installed variable store (model.variables.variable_store).  The returned `endpoints` is an
OrderedDict in the reference's insertion order whose values are computed on first access (each access is one
forward pass through the fused engine up to that node) and then cached."""
from collections import OrderedDict

import numpy as np

from . import variables as V


class Endpoints(OrderedDict):
    """OrderedDict of lazily evaluated network outputs."""

    def __init__(self, names, compute):
        super(Endpoints, self).__init__()
        self._compute = compute
        for n in names:
            OrderedDict.__setitem__(self, n, None)

    def __getitem__(self, key):
        v = OrderedDict.__getitem__(self, key)
        if v is None and self._compute is not None:
            v = self._compute(key)
            OrderedDict.__setitem__(self, key, v)
        return v

    def items(self):
        return [(k, self[k]) for k in self.keys()]

    def values(self):
        return [self[k] for k in self.keys()]


# tdnn() runs its three conv2d layers on [b, 1, l, d] (model/tdnn.py:38), so these endpoints are rank 4 [B, 1, T', 512];
# tdnn3_relu is overwritten with the squeezed tensor (:94-96).  etdnn() uses conv1d: rank 3 throughout.
TDNN_RANK4_NODES = frozenset(("tdnn1_conv", "tdnn1_bn", "tdnn1_relu", "tdnn2_conv", "tdnn2_bn", "tdnn2_relu",
                              "tdnn3_conv", "tdnn3_bn"))

TDNN_FRAME_LAYERS = (("tdnn1", True), ("tdnn2", True), ("tdnn3", True), ("tdnn4", False), ("tdnn5", False))
# etdnn: conv1d k=5,5,7,9 interleaved with dense layers (model/tdnn.py:376-518)
ETDNN_FRAME_LAYERS = (("tdnn1", True), ("tdnn2", False), ("tdnn3", True), ("tdnn4", False), ("tdnn5", True),
                      ("tdnn6", False), ("tdnn7", True), ("tdnn8", False), ("tdnn9", False), ("tdnn10", False))


def endpoint_names(params, frame_layers=TDNN_FRAME_LAYERS, seg=("tdnn6", "tdnn7")):
    """Names tdnn()/etdnn() add to endpoints, in order (model/tdnn.py:48-179, :384-589,
    model/pooling.py:24,198,220,227-229)."""
    names = []
    for name, conv in frame_layers:
        names += ["%s_%s" % (name, "conv" if conv else "dense"), "%s_bn" % name, "%s_relu" % name]
    if params.pooling_type == "self_attention":
        def side(prefix, nodes, last_type):
            out = []
            nodes = list(nodes)
            for j in range(len(nodes)):
                n = "%s%d" % (prefix, j)
                last = j == len(nodes) - 1
                out.append(n + "_dense")
                if (not last) or last_type == 2:
                    out.append(n + "_bn")
                if (not last) or last_type in (1, 2):
                    out.append(n + "_relu")
                elif last_type == 3:
                    out.append(n + "_tanh")
            return out
        names += side("att_key", params.att_key_num_nodes, params.att_key_network_type)
        names += side("att_value", params.att_value_num_nodes, params.att_value_network_type)
        names += ["attention_weights", "att_output_before_nonlinear"]
        if params.att_apply_nonlinear:
            names += ["att_post_bn", "att_post_relu"]
    names += ["pooling", "%s_dense" % seg[0], "%s_bn" % seg[0], "%s_relu" % seg[0], "%s_dense" % seg[1]]
    if not params.last_layer_no_bn:
        names.append("%s_bn" % seg[1])
    if not params.last_layer_linear:
        names.append("%s_relu" % seg[1])
    return names


def _inject_defaults(params, pooling_layer=True):
    """Defaults the reference injects lazily by mutating params.dict (model/tdnn.py:114-116,152-154,163-164,173-174)."""
    if pooling_layer and "num_nodes_pooling_layer" not in params.dict:
        params.dict["num_nodes_pooling_layer"] = 1500
    if "num_nodes_last_layer" not in params.dict:
        params.dict["num_nodes_last_layer"] = 512
    if "last_layer_no_bn" not in params.dict:
        params.last_layer_no_bn = False
    if "last_layer_linear" not in params.dict:
        params.last_layer_linear = False
    if params.pooling_type not in ("statistics_pooling", "self_attention"):
        raise NotImplementedError("Not implement %s pooling" % params.pooling_type)


def _run_network(features, params, network_type, names, last_prefix):
    x = np.ascontiguousarray(features, dtype=np.float32)
    assert x.ndim == 3, "features must be [batch, length, dim]"
    b, t, d = x.shape
    cfg = dict(params.dict)
    cfg["network_type"] = network_type
    eng = V.get_engine(params, d, config=cfg)
    packed = x.reshape(b * t, d)
    lengths = np.full(b, t, np.int32)

    def compute(node):
        _, frame_level, _ = eng.node_info(node)
        if not frame_level:
            return eng.forward(packed, lengths, node)
        out = np.stack(eng.forward_frames(packed, lengths, node), axis=0)
        if network_type == "tdnn" and node in TDNN_RANK4_NODES:
            out = np.expand_dims(out, 1)
        return out

    endpoints = Endpoints(names, compute)
    last = last_prefix + "_relu" if not params.last_layer_linear else (
        last_prefix + "_bn" if not params.last_layer_no_bn else last_prefix + "_dense")
    return endpoints[last], endpoints


def tdnn(features, params, is_training=None, reuse_variables=None, aux_features=None):
    """Build (= run) a TDNN network.  Same arguments and return value as the reference (model/tdnn.py:10):
    returns (output of the last layer [batch, num_nodes_last_layer], endpoints).  `is_training` must be falsy:
    this build covers the inference path; `reuse_variables` / `aux_features` are accepted and ignored."""
    if is_training:
        raise NotImplementedError("tdnn(is_training=True): the training graph is outside this build's scope")
    _inject_defaults(params)
    return _run_network(features, params, "tdnn", endpoint_names(params), "tdnn7")


def etdnn(features, params, is_training=None, reuse_variables=None, aux_features=None):
    """Extended TDNN (model/tdnn.py:343-591): 10 frame layers, pooling, tdnn12, tdnn13; variable scope "etdnn".
    Same arguments and return value as the reference."""
    if is_training:
        raise NotImplementedError("etdnn(is_training=True): the training graph is outside this build's scope")
    _inject_defaults(params)
    return _run_network(features, params, "extended_tdnn",
                        endpoint_names(params, ETDNN_FRAME_LAYERS, ("tdnn12", "tdnn13")), "tdnn13")
