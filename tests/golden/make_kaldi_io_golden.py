#!/usr/bin/env python
"""Golden fixtures for the Kaldi ark I/O, produced by IMPORTING the reference module
dataset/kaldi_io.py from /root/reference (build container only; the module is Python-3 clean):

  kaldi_io_mats.ark   written by the reference's write_mat      (FM and DM matrices, one with zero rows)
  kaldi_io_vecs.ark   written by the reference's write_vec_flt  (FV and DV vectors)
  kaldi_io_cm.ark     format-1 compressed matrices from oracle/kaldi_io_ref.compress_cm (the reference has no
                      compressor), DECODED by the reference's read_mat_ark
  kaldi_io_text.ark   text matrices, decoded by the reference's read_mat_ark
  kaldi_io_golden.npz what the reference's readers return for every key of the four arks, plus sub-matrix reads
                      through the reference's _read_submat_binary / _read_compressed_submat
"""
import io
import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")


def main():
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        from dataset import kaldi_io as R
    from oracle import kaldi_io_ref as O
    rng = np.random.default_rng(20260)
    gold = {}

    mats = [("fm_%d" % i, (rng.standard_normal((n, 30)) * np.linspace(3, 0.5, 30)).astype(np.float32))
            for i, n in enumerate((1, 17, 200))]
    mats.append(("dm_0", rng.standard_normal((9, 23))))
    mats.append(("fm_empty", np.zeros((0, 30), np.float32)))
    mats.append(("spk1-utt_07.wav", rng.standard_normal((33, 40)).astype(np.float32)))
    with open(os.path.join(HERE, "kaldi_io_mats.ark"), "wb") as f:
        for k, m in mats:
            R.write_mat(f, m, key=k)
    for k, m in R.read_mat_ark(os.path.join(HERE, "kaldi_io_mats.ark")):
        gold["mats/" + k] = m

    vecs = [("fv_0", rng.standard_normal(512).astype(np.float32)), ("dv_0", rng.standard_normal(200)),
            ("fv_1", rng.standard_normal(1).astype(np.float32))]
    with open(os.path.join(HERE, "kaldi_io_vecs.ark"), "wb") as f:
        for k, v in vecs:
            R.write_vec_flt(f, v, key=k)
    for k, v in R.read_vec_flt_ark(os.path.join(HERE, "kaldi_io_vecs.ark")):
        gold["vecs/" + k] = v

    cms = [("cm_%d" % i, (rng.standard_normal((n, d)) * np.linspace(20, 0.1, d) + np.linspace(-5, 5, d)).astype(np.float32))
           for i, (n, d) in enumerate(((300, 30), (5, 24), (64, 40)))]
    const = np.tile(np.linspace(-1, 1, 8, dtype=np.float32), (40, 1))
    const[:, 3] = 0.25                                         # a constant column
    cms.append(("cm_const", const))
    with open(os.path.join(HERE, "kaldi_io_cm.ark"), "wb") as f:
        for k, m in cms:
            f.write(O.write_cm_bytes(m, key=k))
    for k, m in R.read_mat_ark(os.path.join(HERE, "kaldi_io_cm.ark")):
        gold["cm/" + k] = np.ascontiguousarray(m)

    text = b"t_0  [\n  1.5 -2 3e-3\n  4 5.25 6 ]\nt_1  [\n  0.1 0.2 ]\n"
    with open(os.path.join(HERE, "kaldi_io_text.ark"), "wb") as f:
        f.write(text)
    for k, m in R.read_mat_ark(os.path.join(HERE, "kaldi_io_text.ark")):
        gold["text/" + k] = m

    # sub-matrix reads (the FeatureReader path): rows [start, start+length) of fm_2 and cm_0
    for ark, key, start, length in (("kaldi_io_mats.ark", "fm_2", 37, 100), ("kaldi_io_cm.ark", "cm_0", 11, 250),
                                    ("kaldi_io_cm.ark", "cm_2", 0, 64)):
        with open(os.path.join(HERE, ark), "rb") as f:
            k = R.read_key(f)
            while k != key:
                R.read_mat(f)
                k = R.read_key(f)
            assert f.read(2) == b"\0B"
            gold["sub/%s/%d/%d" % (key, start, length)] = np.ascontiguousarray(R._read_submat_binary(f, start, length))

    np.savez_compressed(os.path.join(HERE, "kaldi_io_golden.npz"), **gold)
    print("wrote %d arrays" % len(gold), {k: v.shape for k, v in gold.items()})


if __name__ == "__main__":
    main()
