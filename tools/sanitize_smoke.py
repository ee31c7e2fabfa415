#!/usr/bin/env python
"""Small-shape pass over every kernel family of libxvector.so, meant to be run under compute-sanitizer
("race detection / sanitizers"):

  compute-sanitizer --tool memcheck  python tools/sanitize_smoke.py
  compute-sanitizer --tool racecheck python tools/sanitize_smoke.py
  compute-sanitizer --tool synccheck python tools/sanitize_smoke.py

Shapes are the smallest that still take every code path: ragged utterances that straddle row tiles, an odd number
of row tiles (phantom tile of a 2-CTA pair), the column-stationary pooled schedule, attention with fused scores,
a mixed-precision training step (tcgen05 forward / dgrad / MN-major wgrad + every SIMT kernel of the step).
Results are checked for finiteness only; parity is the job of tests/ (this script must stay fast under a 50x
slowdown).  `--only a,b` restricts the sections."""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    only = set(x for x in args.only.split(",") if x)
    want = lambda name: not only or name in only
    import xvector_b200
    from oracle import synth                      # generators only

    dim = 24
    lengths = np.array([150, 333, 61, 40, 517], np.int32)
    feats = synth.packed_features(lengths, dim, seed=7)
    if want("tdnn"):
        cfg = dict(synth.TDNN_STAT_CONFIG)
        w = synth.make_tdnn_weights(dim, cfg, seed=1)
        for prec in ("bf16", "fp32"):
            eng = xvector_b200.Engine(cfg, dim, precision=prec)
            eng.set_weights(w)
            out = eng.forward(feats, lengths)
            assert np.isfinite(out).all()
            fr = eng.forward_frames(feats, lengths, "tdnn3_relu")
            assert all(np.isfinite(f).all() for f in fr)
            print("tdnn", prec, "ok", eng.launch_summary())
            eng.close()
    if want("attention"):
        cfg = dict(synth.TDNN_ATT_CONFIG, att_num_heads=4, att_key_num_nodes=[512, 512])
        w = synth.make_tdnn_weights(dim, cfg, seed=2)
        eng = xvector_b200.Engine(cfg, dim, precision="bf16")
        eng.set_weights(w)
        out = eng.forward(feats, lengths)
        assert np.isfinite(out).all()
        print("attention ok", eng.launch_summary())
        eng.close()
    if want("pool"):
        eng = xvector_b200.Engine(dict(synth.TDNN_STAT_CONFIG), dim, precision="bf16")
        x = np.random.default_rng(3).standard_normal((700, 1500)).astype(np.float32)
        cu = np.array([0, 100, 333, 700], np.int32)
        for dt in ("fp32", "bf16"):
            out = eng.stat_pooling(x, cu, dtype=dt)
            assert np.isfinite(out).all()
        print("pool ok")
        eng.close()
    if want("train"):
        cfg = dict(synth.TDNN_AMSOFTMAX_CONFIG)
        speakers = 101
        w = synth.add_softmax_weights(synth.make_tdnn_weights(dim, cfg, seed=4), speakers)
        x, y = synth.config5_batch(8, 50, dim, speakers, seed=5)
        for prec in ("bf16", "fp32"):
            eng = xvector_b200.Engine(cfg, dim, precision=prec)
            eng.set_weights(w)
            eng.train_init(speakers)
            for i in range(2):
                loss, total = eng.train_step(x, y, 0.05, i)
                assert np.isfinite(loss) and np.isfinite(total)
            print("train", prec, "ok", loss)
            eng.close()
    if want("resnet"):
        cfg = dict(synth.RESNET_CONFIG)
        rdim = 40
        w = synth.make_resnet_weights(rdim, cfg, seed=6)
        eng = xvector_b200.Engine(cfg, rdim, precision="bf16")
        eng.set_weights(w)
        ln = np.array([48, 48, 64], np.int32)
        f = synth.fbank_like(int(ln.sum()), rdim, seed=8)
        out = eng.forward(f, ln)
        assert np.isfinite(out).all()
        print("resnet ok", eng.launch_summary())
        eng.close()
    print("sanitize_smoke OK")


if __name__ == "__main__":
    main()
