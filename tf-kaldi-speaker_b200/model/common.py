"""Layer helpers with the reference's names and argument meaning (model/common.py:7-265), operating on numpy
arrays.  Arithmetic runs on the GPU through libxvector (fp32 SIMT kernels: these direct-call helpers are the
validation/utility surface; the fused bf16 path is reached through tdnn()/Trainer.predict)."""
import numpy as np

from . import variables as V


def shape_list(x):
    """model/common.py:7-24 -- static shape as a list."""
    return list(np.shape(x))


def _relu_kind(params):
    if "network_relu_type" in params.dict:
        if params.network_relu_type == "prelu":
            return "prelu"
        if params.network_relu_type == "lrelu":
            return "lrelu"
    return "relu"


def prelu(x, name="prelu", shared=False):
    """model/common.py:27-42: relu(x) + alpha*(x-|x|)/2, alpha = variable '<scope>/<name>/alpha'."""
    with V.variable_scope(name):
        alpha = np.asarray(V.get_variable("alpha"), np.float32).reshape(-1)
    return V.utility_engine().prelu(x, alpha, shared=shared)


def l2_scaling(x, scaling_factor, epsilon=1e-12, name="l2_norm"):
    """model/common.py:45-58."""
    return V.utility_engine().l2_scaling(x, scaling_factor, epsilon)


def _fold_bn(scope):
    g = np.asarray(V.get_variable(scope + "/gamma"), np.float64)
    b = np.asarray(V.get_variable(scope + "/beta"), np.float64)
    m = np.asarray(V.get_variable(scope + "/moving_mean"), np.float64)
    v = np.asarray(V.get_variable(scope + "/moving_variance"), np.float64)
    s = g / np.sqrt(v + 1e-3)            # tf.layers.batch_normalization epsilon default
    return s.astype(np.float32), (b - s * m).astype(np.float32)


def _slope(params, scope, n):
    kind = _relu_kind(params)
    if kind == "prelu":
        return np.asarray(V.get_variable(scope + "/alpha"), np.float32).reshape(-1)
    return np.full(n, 0.2 if kind == "lrelu" else 0.0, np.float32)


def _dense_generic(features, num_nodes, endpoints, params, name, bn, act):
    x = np.asarray(features, np.float32)
    lead = x.shape[:-1]
    x2 = x.reshape(-1, x.shape[-1])
    eng = V.utility_engine()
    with V.variable_scope(name):
        kernel = np.asarray(V.get_variable("%s_dense/kernel" % name), np.float32)
        bias = np.asarray(V.get_variable("%s_dense/bias" % name), np.float32)
        assert kernel.shape[-1] == num_nodes, "variable %s has %d nodes, asked for %d" % (name, kernel.shape[-1], num_nodes)
        y = eng.dense_rows(x2, kernel, bias)
        endpoints["%s_dense" % name] = y.reshape(lead + (num_nodes,))
        out = y
        s = h = None
        if bn:
            s, h = _fold_bn("%s_bn" % name)
            out = eng.dense_rows(x2, kernel, bias, bn_scale=s, bn_shift=h)
            endpoints["%s_bn" % name] = out.reshape(lead + (num_nodes,))
        if act == "relu":
            sl = _slope(params, "%s_relu" % name, num_nodes)
            out = eng.dense_rows(x2, kernel, bias, bn_scale=s, bn_shift=h, slope=sl)
            endpoints["%s_relu" % name] = out.reshape(lead + (num_nodes,))
        elif act == "tanh":
            out = eng.dense_rows(x2, kernel, bias, bn_scale=s, bn_shift=h, act_tanh=True)
            endpoints["%s_tanh" % name] = out.reshape(lead + (num_nodes,))
    return out.reshape(lead + (num_nodes,))


def dense_bn_relu(features, num_nodes, endpoints, params, is_training=None, name="dense"):
    """model/common.py:113-146."""
    return _dense_generic(features, num_nodes, endpoints, params, name, True, "relu")


def dense(features, num_nodes, endpoints, params, is_training=None, name="dense"):
    """model/common.py:149-169."""
    return _dense_generic(features, num_nodes, endpoints, params, name, False, None)


def dense_relu(features, num_nodes, endpoints, params, is_training=None, name="dense"):
    """model/common.py:172-199."""
    return _dense_generic(features, num_nodes, endpoints, params, name, False, "relu")


def dense_tanh(features, num_nodes, endpoints, params, is_training=None, name="dense"):
    """model/common.py:202-223."""
    return _dense_generic(features, num_nodes, endpoints, params, name, False, "tanh")


def split_last_dimension(x, n):
    """model/common.py:226-241 (pure reshape)."""
    x = np.asarray(x)
    m = x.shape[-1]
    assert m % n == 0
    return x.reshape(x.shape[:-1] + (n, m // n))


def split_heads(x, num_heads):
    """model/common.py:244-253: [batch, length, channels] -> [batch, heads, length, channels/heads]."""
    return np.transpose(split_last_dimension(x, num_heads), [0, 2, 1, 3])


def combine_last_two_dimensions(x):
    """model/common.py:256-265."""
    x = np.asarray(x)
    return x.reshape(x.shape[:-2] + (x.shape[-2] * x.shape[-1],))
