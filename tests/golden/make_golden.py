#!/usr/bin/env python
"""Generate the committed golden fixtures under tests/golden/.

Runs ONLY in the build container (needs /root/reference).  Two kinds of fixture:

1. attention_golden.npz -- outputs of the REFERENCE's own numpy statement
   `compute_self_attention` (model/test_utils.py:321-392), executed from the file
   where it lies.  The file is Python-2 (`value_dim/n_heads` integer division at
   :335,338,367-368), so its source text is patched IN MEMORY ('/'->'//' on those
   reshape arguments) before exec; nothing from the reference is copied into the repo.
   This is the only reference-authored numeric pin on the hot path.

1b. amsoftmax_golden.npz -- outputs of the REFERENCE's own numpy statement `compute_amsoftmax`
   (model/test_utils.py:229-270), executed from the file where it lies: pins the loss VALUE of
   additive_margin_softmax (model/loss.py:201-286) for the training-step oracle (oracle/train_ref.py).

2. tdnn_small_golden.npz -- fp64 oracle (oracle/numpy_ref.py) outputs on small seeded
   inputs.  These are regression fixtures for the oracle itself ("parity unpinned":
   the reference has no runnable implementation or fixture for tdnn/stat pooling).
"""
import os
import re
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference/model/test_utils.py"


def load_reference_attention():
    src = open(REF).read()
    m = re.search(r"^def softmax\(.*?(?=^def )", src, re.S | re.M)
    softmax_src = m.group(0) if m else None
    m = re.search(r"^def compute_self_attention\(.*?(?=^def )", src, re.S | re.M)
    fn_src = m.group(0)
    fn_src = fn_src.replace("value_dim/n_heads", "value_dim//n_heads").replace("key_dim/n_heads", "key_dim//n_heads")
    mod = types.ModuleType("ref_test_utils")
    mod.np = np
    if softmax_src is not None:
        exec(compile(softmax_src, REF, "exec"), mod.__dict__)
    else:  # the reference imports softmax from scipy in some versions
        from scipy.special import softmax as _sm
        mod.softmax = _sm
    exec(compile(fn_src, REF, "exec"), mod.__dict__)
    return mod.compute_self_attention


def load_reference_amsoftmax():
    src = open(REF).read()
    mod = types.ModuleType("ref_test_utils_am")
    mod.np = np
    m = re.search(r"^def softmax\(.*?(?=^def )", src, re.S | re.M)
    if m:
        exec(compile(m.group(0), REF, "exec"), mod.__dict__)
    else:
        from scipy.special import softmax as _sm
        mod.softmax = lambda x: _sm(x, axis=1)
    m = re.search(r"^def compute_amsoftmax\(.*?(?=^def )", src, re.S | re.M)
    exec(compile(m.group(0), REF, "exec"), mod.__dict__)
    return mod.compute_amsoftmax


class P(object):
    pass


def main():
    from oracle import numpy_ref, synth
    fn = load_reference_attention()
    rng = np.random.default_rng(7)
    out = {}
    cases = [
        # (name, B, L, Cv, Ck, H, split_value, split_key, use_scale)
        ("h1_split", 3, 37, 24, 16, 1, True, True, True),
        ("h4_split", 2, 50, 32, 24, 4, True, True, True),
        ("h4_nosplit_value", 2, 41, 12, 24, 4, False, True, False),
        ("h3_nosplit_key", 2, 29, 18, 10, 3, True, False, True),
        ("h2_nosplit_both", 2, 33, 10, 14, 2, False, False, False),
    ]
    for name, b, l, cv, ck, h, sv, sk, sc in cases:
        p = P()
        p.att_split_value, p.att_split_key, p.att_use_scale, p.att_penalty_term = sv, sk, sc, 0.0
        value = np.abs(rng.standard_normal((b, l, cv)))
        key = rng.standard_normal((b, l, ck))
        query = 0.5 * rng.standard_normal((h, ck // h if sk else ck))
        att, _ = fn(value, key, query, p)
        out[name + "/value"] = value
        out[name + "/key"] = key
        out[name + "/query"] = query
        out[name + "/flags"] = np.array([sv, sk, sc], np.int32)
        out[name + "/att"] = att
    np.savez_compressed(os.path.join(HERE, "attention_golden.npz"), **out)
    print("wrote attention_golden.npz with", len(cases), "cases")

    # --- additive margin softmax loss values from the reference's numpy statement ---
    am = load_reference_amsoftmax()
    a = {}
    am_cases = [("m020_step0", 8, 16, 11, 0.20, 0), ("m020_step30000", 8, 16, 11, 0.20, 30000),
                ("m035_step200000", 5, 24, 7, 0.35, 200000), ("m015_lmin", 6, 12, 9, 0.15, 10 ** 7)]
    for name, b, e, s_, m_, step in am_cases:
        p = P()
        p.feature_norm = False
        p.amsoftmax_m = m_
        p.amsoftmax_lambda_min = 0.0 if "lmin" not in name else 0.5
        p.amsoftmax_lambda_base, p.amsoftmax_lambda_gamma, p.amsoftmax_lambda_power = 1000.0, 0.0001, 5.0
        p.global_step = step
        emb = rng.standard_normal((b, e))
        wmat = rng.standard_normal((e, s_)) * 0.3
        lab = rng.integers(0, s_, b)
        a[name + "/embeddings"] = emb
        a[name + "/w"] = wmat
        a[name + "/labels"] = lab.astype(np.int32)
        a[name + "/hyper"] = np.array([m_, p.amsoftmax_lambda_min, 1000.0, 0.0001, 5.0, step], np.float64)
        a[name + "/loss"] = np.float64(am(emb.copy(), lab, p, wmat.copy()))
    np.savez_compressed(os.path.join(HERE, "amsoftmax_golden.npz"), **a)
    print("wrote amsoftmax_golden.npz with", len(am_cases), "cases")

    # --- oracle regression fixtures (small) ---
    g = {}
    for tag, cfg in (("stat", synth.TDNN_STAT_CONFIG), ("att", dict(synth.TDNN_ATT_CONFIG, att_num_heads=3))):
        cfg = dict(cfg)
        cfg["num_nodes_pooling_layer"] = 1500
        w = synth.make_tdnn_weights(23, cfg, seed=99)
        rng = np.random.default_rng(5)
        x = synth.mfcc_like(rng, 2 * 40, 23).reshape(2, 40, 23)
        params = numpy_ref.OracleParams(**cfg)
        _, ep = numpy_ref.entire_network(x, params, w)
        g[tag + "/x"] = x
        for node in ("tdnn1_relu", "tdnn3_relu", "tdnn5_relu", "pooling", "tdnn6_dense", "output"):
            v = ep[node]
            v = v[:, 0] if v.ndim == 4 else v            # conv2d endpoints are [B, 1, T', 512]
            g[tag + "/" + node] = v if v.ndim < 3 else v[:, ::7, ::37]     # subsample frame-level nodes
    np.savez_compressed(os.path.join(HERE, "tdnn_small_golden.npz"), **g)
    print("wrote tdnn_small_golden.npz")


if __name__ == "__main__":
    main()
