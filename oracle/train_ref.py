"""fp64 numpy restatement of ONE training step of the reference (BASELINE.json config 5): TDNN in training mode
(batch-statistics BN), additive-margin softmax or plain softmax, L2 regularisation, hand-derived backward pass,
SGD / momentum / Adam update, per-tensor gradient clipping, BN moving-average update.  TEST INFRASTRUCTURE -- see oracle/__init__.py.

Follows, line by line:
  model/tdnn.py:36-181 with is_training=True            (network, BN -> activation order)
  model/pooling.py:40-50                                 (statistics pooling)
  model/loss.py:9-48, 201-286                            (softmax, additive_margin_softmax)
  model/trainer.py:225-246, 472-527, 563-570             (optimizer, total loss = loss + regularisation, gradients)
  model/trainer.py:1133-1150, 1183-1190, 1225-1248       (data-parallel towers: batch slices, per-tower BN
                                                          statistics, moving averages from tower 0, add_n/N)
TensorFlow-1.x defaults restated here (the library is un-vendored): BN epsilon 1e-3; the moving
variance update of FUSED batch norm (rank-4 inputs = the conv layers tdnn1-3) uses the unbiased batch variance
(n/(n-1)), the non-fused implementation (rank-2/3 inputs = tdnn4-7) the biased one; tf.contrib.layers.
l2_regularizer(s)(w) = s * sum(w^2) / 2 on kernels only; tf.nn.l2_normalize(x) = x * rsqrt(max(sum x^2, 1e-12));
GradientDescentOptimizer: v -= lr * g; MomentumOptimizer: a = m*a + g, v -= lr*a (Nesterov: v -= lr*(g + m*a)).

Parity status: the AM-softmax LOSS VALUE is pinned by the reference's own numpy statement
model/test_utils.py:229-270 (tests/golden/amsoftmax_golden.npz); gradients and the update are PARITY UNPINNED by
the reference (it ships no gradient test) and are cross-checked against torch autograd in fp64
(oracle/torch_train_ref.py).
"""
from __future__ import annotations

from collections import OrderedDict

import numpy as np

from .numpy_ref import BN_EPSILON, LRELU_ALPHA, VAR2STD_EPSILON

TDNN_LAYERS = (("tdnn1", "conv", 5), ("tdnn2", "conv", 5), ("tdnn3", "conv", 7), ("tdnn4", "dense", 1),
               ("tdnn5", "dense", 1))


def amsoftmax_lambda(params, global_step):
    """model/loss.py:265-268."""
    lam = max(float(params.amsoftmax_lambda_min),
              float(params.amsoftmax_lambda_base) * (1.0 + float(params.amsoftmax_lambda_gamma) * float(global_step))
              ** (-float(params.amsoftmax_lambda_power)))
    return lam


def _act_fwd(x, kind, alpha):
    if kind == "prelu":
        return np.maximum(x, 0.0) + alpha * (x - np.abs(x)) * 0.5
    if kind == "lrelu":
        return np.where(x >= 0.0, x, LRELU_ALPHA * x)
    return np.maximum(x, 0.0)


def _act_bwd(dy, x, kind, alpha):
    """Returns (dx, dalpha or None).  ReLU's gradient at exactly 0 is 0 (tf.nn.relu)."""
    if kind == "prelu":
        neg = x < 0.0
        dx = np.where(neg, alpha * dy, dy)
        axes = tuple(range(x.ndim - 1))
        return dx, np.sum(np.where(neg, dy * x, 0.0), axis=axes)
    if kind == "lrelu":
        return np.where(x >= 0.0, dy, LRELU_ALPHA * dy), None
    return np.where(x > 0.0, dy, 0.0), None


def bf16_round(x):
    """fp64 -> fp32 -> bf16 (round to nearest even), returned as fp64: the operand rounding of the mixed-precision
    training step (tcgen05 GEMMs with bf16 operands and fp32 accumulation)."""
    x32 = np.ascontiguousarray(x, np.float32)
    u = x32.view(np.uint32).astype(np.uint64)
    u = (u + 0x7FFF + ((u >> 16) & 1)) & 0xFFFF0000
    return u.astype(np.uint32).view(np.float32).reshape(x32.shape).astype(np.float64)


class _Affine(object):
    """conv(1xk)/dense + BN(batch statistics) [+ activation] with a cached forward for the backward pass."""

    def __init__(self, w, scope, name, kind, taps, has_bn=True, act=True, relu_kind=None, fused_bn=False, bf16_gemm=False):
        self.scope, self.name, self.kind, self.taps = scope, name, kind, taps
        # bf16_gemm: the three GEMMs of this layer (forward, dgrad, wgrad) read bf16-rounded operands (activations,
        # dZ and a per-step bf16 copy of the fp32 master kernel); everything else stays exact
        self.rnd = bf16_round if bf16_gemm else (lambda v: v)
        self.has_bn, self.act, self.relu_kind, self.fused_bn = has_bn, act, relu_kind, fused_bn
        aff = "%s/%s_%s" % (scope, name, kind)
        self.kname, self.bname = aff + "/kernel", aff + "/bias"
        self.bn = "%s/%s_bn" % (scope, name)
        self.aname = "%s/%s_relu/alpha" % (scope, name)
        k = np.asarray(w[self.kname], np.float64)
        self.k = k.reshape(taps, -1, k.shape[-1])                 # [taps, Cin, Cout]
        self.b = np.asarray(w[self.bname], np.float64)
        if has_bn:
            self.gamma = np.asarray(w[self.bn + "/gamma"], np.float64)
            self.beta = np.asarray(w[self.bn + "/beta"], np.float64)
        self.alpha = np.asarray(w[self.aname], np.float64) if (act and relu_kind == "prelu") else None

    def forward(self, a):
        """a [..., L, Cin] (rank 3) or [B, Cin] (rank 2)."""
        a = self.rnd(a)
        kq = self.rnd(self.k)
        self.a = a
        if a.ndim == 3:
            lo = a.shape[1] - self.taps + 1
            z = np.zeros((a.shape[0], lo, self.k.shape[2]))
            for j in range(self.taps):
                z += a[:, j:j + lo, :] @ kq[j]
        else:
            z = a @ kq[0]
        z = z + self.b
        self.z = z
        y = z
        if self.has_bn:
            axes = tuple(range(z.ndim - 1))
            self.n = int(np.prod([z.shape[i] for i in axes]))
            self.mu = z.mean(axis=axes)
            self.var = z.var(axis=axes)                            # biased
            self.rstd = 1.0 / np.sqrt(self.var + BN_EPSILON)
            self.xhat = (z - self.mu) * self.rstd
            y = self.gamma * self.xhat + self.beta
        self.pre_act = y
        self.near_zero = 0
        if self.act:
            # pre-activations within fp32 rounding of 0: an fp32 implementation may take the other branch of the
            # activation's derivative there (tests use this count to choose between the strict and the loose bar)
            mag = (np.abs(self.gamma * self.xhat) + np.abs(self.beta)) if self.has_bn else np.abs(z)
            self.near_zero = int(np.sum(np.abs(y) < 2e-6 * np.maximum(mag, 1e-30)))
            y = _act_fwd(y, self.relu_kind, self.alpha)
        return y

    def backward(self, dy, grads):
        if self.act:
            dy, dalpha = _act_bwd(dy, self.pre_act, self.relu_kind, self.alpha)
            if dalpha is not None:
                grads[self.aname] = dalpha
        if self.has_bn:
            axes = tuple(range(dy.ndim - 1))
            grads[self.bn + "/gamma"] = np.sum(dy * self.xhat, axis=axes)
            grads[self.bn + "/beta"] = np.sum(dy, axis=axes)
            dxhat = dy * self.gamma
            dz = self.rstd / self.n * (self.n * dxhat - dxhat.sum(axis=axes) - self.xhat * np.sum(dxhat * self.xhat, axis=axes))
        else:
            dz = dy
        a = self.a
        grads[self.bname] = dz.reshape(-1, dz.shape[-1]).sum(axis=0)
        dk = np.zeros_like(self.k)
        da = np.zeros_like(a)
        dzq, kq = self.rnd(dz), self.rnd(self.k)
        if a.ndim == 3:
            lo = dz.shape[1]
            for j in range(self.taps):
                dk[j] = np.einsum("blc,blo->co", a[:, j:j + lo, :], dzq)
                da[:, j:j + lo, :] += dzq @ kq[j].T
        else:
            dk[0] = a.T @ dzq
            da = dzq @ kq[0].T
        grads[self.kname] = dk
        return da

    def moving_updates(self, w, momentum, out):
        """tf.layers.batch_normalization update ops: moving = moving*momentum + batch*(1-momentum)."""
        if not self.has_bn:
            return
        var = self.var * (self.n / max(self.n - 1.0, 1.0)) if self.fused_bn else self.var
        mm = np.asarray(w[self.bn + "/moving_mean"], np.float64)
        mv = np.asarray(w[self.bn + "/moving_variance"], np.float64)
        out[self.bn + "/moving_mean"] = mm * momentum + self.mu * (1.0 - momentum)
        out[self.bn + "/moving_variance"] = mv * momentum + var * (1.0 - momentum)


def loss_forward_backward(features, labels, w, params, global_step, want_grad=True):
    """softmax (model/loss.py:9-48) or additive_margin_softmax (:201-286).  Returns (loss, dfeatures, {var: grad})."""
    f = np.asarray(features, np.float64)
    b = f.shape[0]
    labels = np.asarray(labels, np.int64)
    idx = np.arange(b)
    grads = {}
    wk = np.asarray(w["softmax/output/kernel"], np.float64)
    if params.loss_func == "softmax":
        bias = np.asarray(w["softmax/output/bias"], np.float64)
        logits = f @ wk + bias
        clipped = None
    elif params.loss_func == "additive_margin_softmax":
        ss = np.sum(wk * wk, axis=0)
        r = 1.0 / np.sqrt(np.maximum(ss, 1e-12))                  # tf.nn.l2_normalize(w, dim=0)
        wn = wk * r
        base = f @ wn
        sel = base[idx, labels]
        eps = 1e-12
        fnorm_raw = np.sqrt(np.sum(f * f, axis=1))
        fnorm = np.maximum(fnorm_raw, eps)
        cos = sel / fnorm
        clipped = (cos < -1 + eps) | (cos > 1 - eps)
        cosc = np.clip(cos, -1 + eps, 1 - eps)
        m = float(params.amsoftmax_m)
        scaled = (cosc - m) * fnorm
        lam = amsoftmax_lambda(params, global_step)
        fa = 1.0 / (1.0 + lam)
        logits = base.copy()
        logits[idx, labels] = sel + fa * (scaled - sel)             # fs*logits + fa*logits_amsoftmax
    else:
        raise NotImplementedError("Not implement %s loss" % params.loss_func)
    mx = logits.max(axis=1, keepdims=True)
    lse = mx[:, 0] + np.log(np.exp(logits - mx).sum(axis=1))
    loss = float(np.mean(lse - logits[idx, labels]))                # tf.losses.sparse_softmax_cross_entropy (mean)
    if not want_grad:
        return loss, None, grads
    p = np.exp(logits - lse[:, None])
    dl = p
    dl[idx, labels] -= 1.0
    dl /= b
    if params.loss_func == "softmax":
        grads["softmax/output/kernel"] = f.T @ dl
        grads["softmax/output/bias"] = dl.sum(axis=0)
        return loss, dl @ wk.T, grads
    g = dl[idx, labels].copy()
    dbase = dl
    dbase[idx, labels] = g * (1.0 - fa * clipped)
    dfn = np.where(clipped, g * fa * (cosc - m), -g * fa * m)
    df = dbase @ wn.T
    nz = fnorm_raw > eps
    df += np.where(nz, dfn / np.where(nz, fnorm_raw, 1.0), 0.0)[:, None] * f
    dwn = f.T @ dbase
    dw = r * (dwn - wn * np.sum(dwn * wn, axis=0))
    dw = np.where(ss > 1e-12, dw, r * dwn)
    grads["softmax/output/kernel"] = dw
    return loss, df, grads


def build_layers(w, params, bf16_gemm=False):
    kind = params.dict.get("network_relu_type", None) if "network_relu_type" in params.dict else None
    if kind not in ("prelu", "lrelu"):
        kind = None
    frame = [_Affine(w, "tdnn", n, k, t, relu_kind=kind, fused_bn=(k == "conv"), bf16_gemm=bf16_gemm) for n, k, t in TDNN_LAYERS]
    no_bn = bool(params.dict.get("last_layer_no_bn", False))
    linear = bool(params.dict.get("last_layer_linear", False))
    seg = [_Affine(w, "tdnn", "tdnn6", "dense", 1, relu_kind=kind),
           _Affine(w, "tdnn", "tdnn7", "dense", 1, has_bn=not no_bn, act=not linear, relu_kind=kind)]
    return frame, seg


def tower_forward_backward(features, labels, w, params, global_step=0, want_grad=True, bf16_gemm=False):
    """One tower: forward in training mode, loss, full backward.  Returns dict(loss, reg, total, grads, moving,
    endpoints).  grads include the L2-regularisation gradient wd * W of every kernel."""
    if params.network_type != "tdnn" or params.pooling_type != "statistics_pooling":
        raise NotImplementedError("training oracle: tdnn + statistics_pooling only")
    frame, seg = build_layers(w, params, bf16_gemm)
    ep = OrderedDict()
    x = np.asarray(features, np.float64)
    for l in frame:
        x = l.forward(x)
        ep[l.name + "_relu"] = x
    # statistics pooling (model/pooling.py:40-50)
    length = x.shape[1]
    mean = x.mean(axis=1)
    var = ((x - mean[:, None, :]) ** 2).mean(axis=1)
    floor = var <= VAR2STD_EPSILON
    std = np.sqrt(np.where(floor, VAR2STD_EPSILON, var))
    pooled = np.concatenate([mean, std], axis=1)
    ep["pooling"] = pooled
    y = pooled
    for l in seg:
        y = l.forward(y)
    ep["output"] = y
    if params.dict.get("feature_norm", False):
        raise NotImplementedError("training oracle: feature_norm is not covered")
    loss, dy, grads = loss_forward_backward(y, labels, w, params, global_step, want_grad)
    wd = float(params.weight_l2_regularizer)
    wd_out = float(params.dict.get("output_weight_l2_regularizer", wd))
    reg = 0.0
    for l in frame + seg:
        reg += wd * 0.5 * float(np.sum(l.k * l.k))
    reg += wd_out * 0.5 * float(np.sum(np.asarray(w["softmax/output/kernel"], np.float64) ** 2))
    out = {"loss": loss, "reg": reg, "total": loss + reg, "endpoints": ep, "grads": None, "moving": {},
           "near_zero": sum(l.near_zero for l in frame + seg)}
    momentum = float(params.batchnorm_momentum)
    for l in frame + seg:
        l.moving_updates(w, momentum, out["moving"])
    if not want_grad:
        return out
    for l in reversed(seg):
        dy = l.backward(dy, grads)
    c = mean.shape[1]
    dmean, dstd = dy[:, :c], dy[:, c:]
    dvar = np.where(floor, 0.0, dstd / (2.0 * std))
    dx = dmean[:, None, :] / length + dvar[:, None, :] * 2.0 * (x - mean[:, None, :]) / length
    for l in reversed(frame):
        dx = l.backward(dx, grads)
    for l in frame + seg:
        grads[l.kname] = grads[l.kname].reshape(np.asarray(w[l.kname]).shape) + wd * np.asarray(w[l.kname], np.float64)
    grads["softmax/output/kernel"] = grads["softmax/output/kernel"] + wd_out * np.asarray(w["softmax/output/kernel"], np.float64)
    out["grads"] = grads
    return out


def train_step(features, labels, w, params, learning_rate, global_step=0, num_towers=1, opt_state=None, bf16_gemm=False):
    """TrainerMGPU semantics (model/trainer.py:1133-1248) for `num_towers` towers (1 = Trainer): the batch is cut
    into equal slices, every tower uses its own BN batch statistics, gradients are averaged (add_n / N), moving
    averages come from tower 0, one optimizer step.  Returns (new_weights, info)."""
    n = features.shape[0]
    assert n % num_towers == 0
    per = n // num_towers
    outs = [tower_forward_backward(features[i * per:(i + 1) * per], labels[i * per:(i + 1) * per], w, params, global_step,
                                   bf16_gemm=bf16_gemm)
            for i in range(num_towers)]
    grads = {k: sum(o["grads"][k] for o in outs) / num_towers for k in outs[0]["grads"]}
    if params.dict.get("clip_gradient", False):
        c = float(params.clip_gradient_norm)
        for k in grads:                                              # tf.clip_by_norm per tensor (trainer.py:529-543)
            nrm = np.sqrt(np.sum(grads[k] ** 2))
            grads[k] = grads[k] * c / max(nrm, c)
    opt = params.dict.get("optimizer", "sgd")
    new_w = {k: np.asarray(v, np.float64).copy() for k, v in w.items()}
    state = {} if opt_state is None else opt_state
    for k, g in grads.items():
        if opt == "sgd":
            new_w[k] = new_w[k] - learning_rate * g
        elif opt == "momentum":
            mom = float(params.momentum)
            acc = mom * state.get(k, 0.0) + g
            state[k] = acc
            new_w[k] = new_w[k] - learning_rate * ((g + mom * acc) if params.dict.get("use_nesterov", False) else acc)
        elif opt == "adam":
            # tf.train.AdamOptimizer(learning_rate) defaults (model/trainer.py:241-243): beta1 0.9, beta2 0.999, eps 1e-8;
            # lr_t = lr * sqrt(1 - b2^t) / (1 - b1^t), m = b1 m + (1-b1) g, v = b2 v + (1-b2) g^2, w -= lr_t m / (sqrt(v) + eps)
            t = state.get("__adam_t", 0) + (1 if k == next(iter(grads)) else 0)
            state["__adam_t"] = t
            m, v = state.get(k, (0.0, 0.0))
            m = 0.9 * m + 0.1 * g
            v = 0.999 * v + 0.001 * g * g
            state[k] = (m, v)
            lr_t = learning_rate * np.sqrt(1.0 - 0.999 ** t) / (1.0 - 0.9 ** t)
            new_w[k] = new_w[k] - lr_t * m / (np.sqrt(v) + 1e-8)
        else:
            raise NotImplementedError("training oracle: optimizer %s" % opt)
    for k, v in outs[0]["moving"].items():
        new_w[k] = v
    info = {"loss": float(np.mean([o["loss"] for o in outs])), "reg": outs[0]["reg"], "grads": grads,
            "tower_losses": [o["loss"] for o in outs], "opt_state": state, "near_zero": sum(o["near_zero"] for o in outs)}
    info["total"] = info["loss"] + info["reg"]
    return new_w, info
