"""Independent fp64 torch restatement of the training-mode forward + loss of config 5; gradients come from
autograd.  TEST INFRASTRUCTURE (oracle/__init__.py): used only to cross-check the hand-derived backward pass of
oracle/train_ref.py.  Built from torch's conv2d / linear / batch_norm / cross_entropy, not from train_ref.py."""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F

BN_EPSILON = 1e-3


def total_loss_and_grads(features, labels, w, cfg, global_step=0):
    """Returns (loss, reg, {variable: d(loss+reg)/d variable}) for one tower, TF variable names/layouts."""
    relu_kind = cfg.get("network_relu_type", None)
    t = {k: torch.tensor(np.asarray(v, np.float64), requires_grad=True) for k, v in w.items()
         if not k.endswith("moving_mean") and not k.endswith("moving_variance")}

    def act(x, scope, ch_axis):
        if relu_kind == "prelu":
            a = t[scope + "/alpha"]
            shape = [1] * x.dim()
            shape[ch_axis] = -1
            return torch.relu(x) + a.view(shape) * (x - x.abs()) * 0.5
        if relu_kind == "lrelu":
            return F.leaky_relu(x, 0.2)
        return torch.relu(x)

    def bn(x, scope):                                     # channels on dim 1, batch statistics
        return F.batch_norm(x, None, None, t[scope + "/gamma"], t[scope + "/beta"], True, 0.0, BN_EPSILON)

    x = torch.tensor(np.asarray(features, np.float64)).permute(0, 2, 1).unsqueeze(2)       # [B,D,1,T]
    for i in (1, 2, 3):
        k = t["tdnn/tdnn%d_conv/kernel" % i].permute(3, 2, 0, 1)
        x = act(bn(F.conv2d(x, k, t["tdnn/tdnn%d_conv/bias" % i]), "tdnn/tdnn%d_bn" % i), "tdnn/tdnn%d_relu" % i, 1)
    x = x.squeeze(2)
    for i in (4, 5):
        x = F.linear(x.transpose(1, 2), t["tdnn/tdnn%d_dense/kernel" % i].t(), t["tdnn/tdnn%d_dense/bias" % i]).transpose(1, 2)
        x = act(bn(x, "tdnn/tdnn%d_bn" % i), "tdnn/tdnn%d_relu" % i, 1)
    mean = x.mean(dim=2)
    var = ((x - mean.unsqueeze(2)) ** 2).mean(dim=2)
    mask = (var <= 1e-12).double()
    var = (1.0 - mask) * var + mask * 1e-12
    p = torch.cat([mean, var.sqrt()], dim=1)
    y = act(bn(F.linear(p, t["tdnn/tdnn6_dense/kernel"].t(), t["tdnn/tdnn6_dense/bias"]), "tdnn/tdnn6_bn"), "tdnn/tdnn6_relu", 1)
    y = F.linear(y, t["tdnn/tdnn7_dense/kernel"].t(), t["tdnn/tdnn7_dense/bias"])
    if not cfg.get("last_layer_no_bn", False):
        y = bn(y, "tdnn/tdnn7_bn")
    if not cfg.get("last_layer_linear", False):
        y = act(y, "tdnn/tdnn7_relu", 1)
    lab = torch.tensor(np.asarray(labels, np.int64))
    wk = t["softmax/output/kernel"]
    if cfg["loss_func"] == "softmax":
        logits = y @ wk + t["softmax/output/bias"]
    else:
        wn = wk * torch.rsqrt(torch.clamp((wk * wk).sum(dim=0, keepdim=True), min=1e-12))
        logits = y @ wn
        idx = torch.arange(y.shape[0])
        sel = logits[idx, lab]
        fn = torch.clamp(y.norm(dim=1), min=1e-12)
        cos = torch.clamp(sel / fn, -1 + 1e-12, 1 - 1e-12)
        scaled = (cos - float(cfg["amsoftmax_m"])) * fn
        onehot = F.one_hot(lab, logits.shape[1]).double()
        logits_am = logits + onehot * (scaled - sel).unsqueeze(1)
        lam = max(float(cfg["amsoftmax_lambda_min"]),
                  float(cfg["amsoftmax_lambda_base"]) * (1.0 + float(cfg["amsoftmax_lambda_gamma"]) * global_step)
                  ** (-float(cfg["amsoftmax_lambda_power"])))
        fa = 1.0 / (1.0 + lam)
        logits = (1.0 - fa) * logits + fa * logits_am
    loss = F.cross_entropy(logits, lab)
    wd = float(cfg["weight_l2_regularizer"])
    wd_out = float(cfg.get("output_weight_l2_regularizer", wd))
    reg = sum(wd * 0.5 * (v * v).sum() for k, v in t.items() if k.startswith("tdnn/") and k.endswith("/kernel"))
    reg = reg + wd_out * 0.5 * (wk * wk).sum()
    (loss + reg).backward()
    grads = {k: (v.grad.numpy() if v.grad is not None else None) for k, v in t.items()}
    return float(loss), float(reg), grads
