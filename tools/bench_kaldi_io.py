#!/usr/bin/env python
"""Throughput of the native ark reader (lib/libxvkaldi.so) on FM and CM arks, next to the reference's Python reader when
/root/reference is present (build container).  Scratch files go to gpurun_out/ (git-ignored).
usage: python tools/bench_kaldi_io.py [n_utts] [frames] [dim]"""
import os
import sys
import time
import warnings

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    from oracle import kaldi_io_ref as O, synth
    from xvector_b200 import kaldi_io as K
    n, frames, dim = (int(a) for a in (sys.argv[1:4] + ["2000", "600", "30"][len(sys.argv) - 1:]))
    scratch = os.path.join(ROOT, "gpurun_out")
    os.makedirs(scratch, exist_ok=True)
    rng = np.random.default_rng(1)
    base = synth.mfcc_like(rng, frames, dim)
    fm, cm = os.path.join(scratch, "bench_fm.ark"), os.path.join(scratch, "bench_cm.ark")
    cm_body = O.write_cm_bytes(base, "K")[2:]
    with K.open_or_fd(fm, "wb") as f, open(cm, "wb") as g:
        for i in range(n):
            K.write_mat(f, base, key="utt%06d" % i)
            g.write(("utt%06d " % i).encode() + cm_body)
    total = n * frames
    res = {}
    for name, path in (("FM", fm), ("CM", cm)):
        for _ in range(2):                                  # second pass: page cache warm
            t0 = time.perf_counter()
            rows = 0
            for keys, feats, lens in K.BatchReader(path, dim, max_rows=1 << 18, max_utts=4096,
                                                   threads=int(os.environ.get("XVK_BENCH_THREADS", "8"))):
                rows += feats.shape[0]
            dt = time.perf_counter() - t0
        assert rows == total
        res["native_batch_" + name] = total / dt
        t0 = time.perf_counter()
        rows = sum(m.shape[0] for _, m in K.read_mat_ark(path))
        res["native_generator_" + name] = total / (time.perf_counter() - t0)
    # threaded scp random access
    ark, scp = os.path.join(scratch, "bench_x.ark"), os.path.join(scratch, "bench_x.scp")
    with K.open_or_fd("ark,scp:%s,%s" % (ark, scp), "wb") as f:
        for i in range(n):
            K.write_mat(f, base, key="utt%06d" % i)
    table = K.ScpTable(scp)
    idx = np.random.default_rng(2).permutation(len(table))
    rows = np.full(len(idx), frames, np.int32)
    out = np.empty((total, dim), np.float32)
    for threads in (1, 8):
        t0 = time.perf_counter()
        table.read_many(idx, out=out, rows=rows, threads=threads)
        res["native_scp_random_%dthr" % threads] = total / (time.perf_counter() - t0)
    if os.path.isdir("/root/reference/dataset"):
        sys.path.insert(0, "/root/reference")
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            from dataset import kaldi_io as R
        for name, path, cap in (("FM", fm, n), ("CM", cm, max(n // 10, 1))):
            t0 = time.perf_counter()
            rows = 0
            for i, (k, m) in enumerate(R.read_mat_ark(path)):
                rows += m.shape[0]
                if i + 1 == cap:
                    break
            res["reference_python_" + name] = rows / (time.perf_counter() - t0)
    for k, v in res.items():
        print("%-28s %10.1f M frames/s  (%6.2f GB/s of fp32 features)" % (k, v / 1e6, v * dim * 4 / 1e9))
    for p in (fm, cm, ark, scp):
        os.remove(p)


if __name__ == "__main__":
    main()
