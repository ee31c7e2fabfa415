"""CPU: the training-step oracle (config 5) -- AM-softmax loss pinned on the reference's own numpy statement,
hand-derived backward pass cross-checked against torch autograd, tower semantics of TrainerMGPU."""
import os

import numpy as np
import pytest

from oracle import numpy_ref, synth, torch_train_ref, train_ref

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def test_amsoftmax_loss_matches_reference_numpy_statement():
    """Pinned: model/test_utils.py:229-270 compute_amsoftmax executed from the reference tree (make_golden.py)."""
    g = np.load(os.path.join(GOLDEN, "amsoftmax_golden.npz"))
    names = sorted({k.split("/")[0] for k in g.files})
    assert len(names) == 4
    for n in names:
        m, lmin, lb, lg, lp, step = g[n + "/hyper"]
        p = numpy_ref.OracleParams(loss_func="additive_margin_softmax", amsoftmax_m=m, amsoftmax_lambda_min=lmin,
                                   amsoftmax_lambda_base=lb, amsoftmax_lambda_gamma=lg, amsoftmax_lambda_power=lp)
        loss, _, _ = train_ref.loss_forward_backward(g[n + "/embeddings"], g[n + "/labels"],
                                                     {"softmax/output/kernel": g[n + "/w"]}, p, step, want_grad=False)
        assert abs(loss - float(g[n + "/loss"])) < 1e-10


@pytest.mark.parametrize("loss_func,relu,linear,no_bn,step", [
    ("additive_margin_softmax", None, True, False, 0),           # tdnn_amsoftmax_m0.20_linear_bn_1e-2.json
    ("additive_margin_softmax", "prelu", False, False, 30000),
    ("softmax", "lrelu", True, True, 0),                          # last layer without BN: its bias gradient is non-zero
], ids=["amsoftmax_cfg5", "amsoftmax_prelu_step30000", "softmax_lrelu_nobn"])
def test_manual_backward_matches_autograd(loss_func, relu, linear, no_bn, step):
    cfg = dict(synth.TDNN_AMSOFTMAX_CONFIG, loss_func=loss_func, last_layer_linear=linear, last_layer_no_bn=no_bn)
    if relu:
        cfg["network_relu_type"] = relu
    s = 37
    w = synth.add_softmax_weights(synth.make_tdnn_weights(20, cfg, seed=3), s, bias=(loss_func == "softmax"))
    x, y = synth.config5_batch(6, 40, 20, s, seed=1)
    out = train_ref.tower_forward_backward(x, y, w, numpy_ref.OracleParams(**cfg), step)
    loss, reg, grads = torch_train_ref.total_loss_and_grads(x, y, w, cfg, step)
    assert abs(out["loss"] - loss) < 1e-10 and abs(out["reg"] - reg) < 1e-9
    scale = max(np.abs(v).max() for v in grads.values() if v is not None)
    for k, v in grads.items():
        if no_bn and k.startswith("tdnn/tdnn7_bn"):
            continue
        assert v is not None, k
        mine = out["grads"][k].reshape(v.shape)
        # biases in front of a batch norm have an exactly-zero gradient: compare on the global scale
        assert np.abs(mine - v).max() <= 1e-9 * max(np.abs(v).max(), 1e-3 * scale), k


def test_tower_semantics_and_sgd_update():
    cfg = dict(synth.TDNN_AMSOFTMAX_CONFIG)
    s = 23
    w = synth.add_softmax_weights(synth.make_tdnn_weights(16, cfg, seed=5), s)
    x, y = synth.config5_batch(8, 36, 16, s, seed=2)
    p = numpy_ref.OracleParams(**cfg)
    new1, info1 = train_ref.train_step(x, y, w, p, 0.01, 0, num_towers=1)
    new2, info2 = train_ref.train_step(x, y, w, p, 0.01, 0, num_towers=2)
    # per-tower BN statistics make the 2-tower gradient differ from the full-batch one, but not by much
    k = "tdnn/tdnn3_conv/kernel"
    rel = np.abs(info1["grads"][k] - info2["grads"][k]).max() / np.abs(info1["grads"][k]).max()
    assert rel > 1e-6 and np.isfinite(rel)
    np.testing.assert_allclose(new1[k], np.asarray(w[k], np.float64) - 0.01 * info1["grads"][k].reshape(w[k].shape))
    # moving averages: tower 0 only, momentum 0.99, unbiased variance for the fused (conv) layers
    o0 = train_ref.tower_forward_backward(x[:4], y[:4], w, p, 0, want_grad=False)
    np.testing.assert_allclose(new2["tdnn/tdnn1_bn/moving_mean"], o0["moving"]["tdnn/tdnn1_bn/moving_mean"])
    mv = np.asarray(w["tdnn/tdnn4_bn/moving_variance"], np.float64)
    assert np.all(np.abs(new1["tdnn/tdnn4_bn/moving_variance"] - mv) <= 0.0100001 * np.maximum(mv, 5.0))
    assert info2["loss"] == pytest.approx(np.mean(info2["tower_losses"]))


def test_bf16_gemm_option_rounds_only_frame_layer_operands():
    """train_ref's bf16_gemm restates the mixed-precision step: close to the exact step, not identical, same variables."""
    from oracle import numpy_ref, synth, train_ref
    cfg = dict(synth.TDNN_AMSOFTMAX_CONFIG)
    w = synth.add_softmax_weights(synth.make_tdnn_weights(24, cfg, seed=5), 40, bias=False)
    x, y = synth.config5_batch(8, 40, 24, 40, seed=6)
    p = numpy_ref.OracleParams(**cfg)
    _, exact = train_ref.train_step(x, y, w, p, 0.1, 0)
    _, mixed = train_ref.train_step(x, y, w, p, 0.1, 0, bf16_gemm=True)
    assert set(exact["grads"]) == set(mixed["grads"])
    assert 0 < abs(exact["loss"] - mixed["loss"]) <= 1e-2 * abs(exact["loss"])
    r = train_ref.bf16_round(np.array([1.0, 1.0 + 2.0 ** -8, 1.0 + 3 * 2.0 ** -9, -0.1]))
    assert r[0] == 1.0 and r[1] == 1.0 and r[2] == 1.0 + 2.0 ** -7 and abs(r[3] + 0.1) < 2.0 ** -11


def test_adam_and_clip_updates_against_torch():
    """The oracle's Adam / clip_by_norm branches against torch.optim.Adam and torch.nn.utils.clip_grad_norm_ driven by the
    oracle's own gradients.  tf.train.AdamOptimizer's epsilon is the paper's "epsilon hat" (added to the UNcorrected
    sqrt(v)); torch adds eps to the bias-corrected one, so torch is run with eps_t = 1e-8 / sqrt(1 - beta2^t)."""
    import torch
    cfg = dict(synth.TDNN_AMSOFTMAX_CONFIG, optimizer="adam", clip_gradient=True, clip_gradient_norm=0.5)
    s = 23
    w = synth.add_softmax_weights(synth.make_tdnn_weights(20, cfg, seed=5), s, bias=False)
    x, y = synth.config5_batch(4, 40, 20, s, seed=2)
    params = numpy_ref.OracleParams(**cfg)
    lr = 1e-3
    w1, i1 = train_ref.train_step(x, y, w, params, lr, 0)
    raw1 = train_ref.train_step(x, y, w, numpy_ref.OracleParams(**dict(cfg, clip_gradient=False, optimizer="sgd")), lr, 0)[1]["grads"]
    w2, i2 = train_ref.train_step(x, y, w1, params, lr, 1, opt_state=i1["opt_state"])
    raw2 = train_ref.train_step(x, y, w1, numpy_ref.OracleParams(**dict(cfg, clip_gradient=False, optimizer="sgd")), lr, 1)[1]["grads"]
    for k in ("tdnn/tdnn6_dense/kernel", "tdnn/tdnn2_conv/kernel", "softmax/output/kernel", "tdnn/tdnn5_bn/gamma"):
        t = torch.tensor(np.asarray(w[k], np.float64), requires_grad=True)
        opt = torch.optim.Adam([t], lr=lr, betas=(0.9, 0.999), eps=1e-8)
        for step, (raw, ref) in enumerate(((raw1, w1), (raw2, w2)), start=1):
            opt.param_groups[0]["eps"] = 1e-8 / np.sqrt(1.0 - 0.999 ** step)
            t.grad = torch.tensor(raw[k].reshape(t.shape))
            torch.nn.utils.clip_grad_norm_([t], 0.5)          # one tensor per call == tf.clip_by_norm per variable
            opt.step()
            d_ref = ref[k].reshape(t.shape) - np.asarray(w[k], np.float64)
            d = t.detach().numpy() - np.asarray(w[k], np.float64)
            assert np.linalg.norm(d - d_ref) <= 1e-5 * np.linalg.norm(d_ref), k   # clip_grad_norm_ adds 1e-6 to the norm
