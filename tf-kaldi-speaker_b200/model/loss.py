"""The two softmax-family losses of the training step on host arrays, for validation (Trainer.valid): the loss network of
the reference run on the eval-mode network output, a few dozen utterances against the speaker table at a time
(model/loss.py:9-48 softmax, :201-286 additive_margin_softmax; the training step computes the same two inside
xv_train_step).  Arithmetic in float64; the inputs come from the device as float32."""
import numpy as np


def _cross_entropy(logits, labels):
    """tf.losses.sparse_softmax_cross_entropy: mean over the batch of -log softmax(logits)[label]."""
    z = logits - logits.max(axis=1, keepdims=True)
    logp = z - np.log(np.sum(np.exp(z), axis=1, keepdims=True))
    return float(-np.mean(logp[np.arange(len(labels)), labels]))


def softmax(features, labels, kernel, bias):
    """model/loss.py:9-48: logits = features . kernel + bias (tf.layers.dense "output")."""
    f = np.asarray(features, np.float64)
    logits = f @ np.asarray(kernel, np.float64) + np.asarray(bias, np.float64)
    return _cross_entropy(logits, np.asarray(labels, np.int64))


def amsoftmax_lambda(params, global_step):
    """model/loss.py:265-267: max(lambda_min, lambda_base * (1 + gamma * step) ** -power)."""
    p = params.dict if hasattr(params, "dict") else params
    return max(float(p["amsoftmax_lambda_min"]),
               float(p["amsoftmax_lambda_base"]) * (1.0 + float(p["amsoftmax_lambda_gamma"]) * float(global_step))
               ** (-float(p["amsoftmax_lambda_power"])))


def additive_margin_softmax(features, labels, kernel, params, global_step=0):
    """model/loss.py:201-286: logits against the column-normalised kernel; at the label the logit becomes
    |x| (cos(theta) - m) with cos clipped to [-1 + 1e-12, 1 - 1e-12]; the modified logits are blended with the plain ones
    by fa = 1 / (1 + lambda)."""
    p = params.dict if hasattr(params, "dict") else params
    f = np.asarray(features, np.float64)
    w = np.asarray(kernel, np.float64)
    y = np.asarray(labels, np.int64)
    eps = 1e-12
    w_norm = w / np.sqrt(np.maximum(np.sum(w * w, axis=0, keepdims=True), eps))          # tf.nn.l2_normalize(w, dim=0)
    logits = f @ w_norm
    rows = np.arange(len(y))
    sel = logits[rows, y]
    f_norm = np.maximum(np.sqrt(np.sum(f * f, axis=1)), eps)
    cos = np.clip(sel / f_norm, -1 + eps, 1 - eps)
    scaled = (cos - float(p["amsoftmax_m"])) * f_norm
    margin = logits.copy()
    margin[rows, y] = scaled                                                             # logits + scatter(scaled - sel)
    fa = 1.0 / (1.0 + amsoftmax_lambda(p, global_step))
    return _cross_entropy((1.0 - fa) * logits + fa * margin, y)


def valid_loss(loss_func, features, labels, variables, params, global_step=0):
    """The loss of one validation batch from the network output and the loss head's variables (TF names)."""
    if loss_func == "softmax":
        return softmax(features, labels, variables["softmax/output/kernel"], variables["softmax/output/bias"])
    if loss_func == "additive_margin_softmax":
        return additive_margin_softmax(features, labels, variables["softmax/output/kernel"], params, global_step)
    raise NotImplementedError("Not implement %s loss" % loss_func)
