#!/usr/bin/env python
"""train_dp_golden.npz -- the N-tower data-parallel training step (model/trainer.py:1133-1248: batch slices, per-tower BN
statistics, add_n/N, moving averages from tower 0) from the fp64 oracle (oracle/train_ref.py num_towers = 2, 4, 8) on the
small seeded problem bench.py's train_dp leg re-creates (dim 24, 101 speakers, 16 utterances x 50 frames, lr 0.05,
step 0).  bench.py may not import the oracle for checking, so the expected loss / gradients / tower-0 moving mean are
committed here.  parity unpinned by the reference (it ships no gradient test): these are fixtures of the restatement.

  python tests/golden/make_train_dp_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SUBSAMPLE = 97
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))


def main():
    from oracle import numpy_ref, synth, train_ref
    cfg = dict(synth.TDNN_AMSOFTMAX_CONFIG)
    dim, speakers, batch, frames = 24, 101, 16, 50
    w = synth.add_softmax_weights(synth.make_tdnn_weights(dim, cfg, seed=31), speakers)
    x, y = synth.config5_batch(batch, frames, dim, speakers, seed=32)
    out = {}
    for n in (2, 4, 8):
        new_ref, info = train_ref.train_step(x, y, w, numpy_ref.OracleParams(**cfg), 0.05, 0, num_towers=n)
        out["n%d/loss" % n] = np.float64(info["loss"])
        for name in ("tdnn/tdnn3_conv/kernel", "tdnn/tdnn6_dense/kernel", "tdnn/tdnn5_bn/gamma"):
            g = np.asarray(info["grads"][name], np.float32).reshape(np.asarray(w[name]).shape)
            out["n%d/g:%s" % (n, name)] = g.reshape(-1)[::SUBSAMPLE].copy()     # every 97th element: small fixture, every row/column hit
        out["n%d/mm" % n] = np.asarray(new_ref["tdnn/tdnn1_bn/moving_mean"], np.float32)
    np.savez_compressed(os.path.join(HERE, "train_dp_golden.npz"), **out)
    print("wrote train_dp_golden.npz", {k: out[k].shape for k in out if k.startswith("n2/")})


if __name__ == "__main__":
    main()
