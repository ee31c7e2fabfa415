#!/usr/bin/env python
"""Generate the committed golden fixtures under tests/golden/.

Runs ONLY in the build container (needs /root/reference).  Two kinds of fixture:

1. attention_golden.npz -- outputs of the REFERENCE's own numpy statement
   `compute_self_attention` (model/test_utils.py:321-392), executed from the file
   where it lies.  The file is Python-2 (`value_dim/n_heads` integer division at
   :335,338,367-368), so its source text is patched IN MEMORY ('/'->'//' on those
   reshape arguments) before exec; nothing from the reference is copied into the repo.
   This is the only reference-authored numeric pin on the hot path.

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
            g[tag + "/" + node] = v if v.ndim < 3 else v[:, ::7, ::37]     # subsample frame-level nodes
    np.savez_compressed(os.path.join(HERE, "tdnn_small_golden.npz"), **g)
    print("wrote tdnn_small_golden.npz")


if __name__ == "__main__":
    main()
