#!/usr/bin/env python
"""statpool_golden.npz -- statistics pooling pinned on the reference's own numpy statement.

`compute_phone_pooling` (model/test_utils.py:1110-1131) opens with
    p1 = np.concatenate([np.mean(features, axis=1), np.std(features, axis=1)], axis=1)
which is the reference author's numeric statement of statistics_pooling (model/pooling.py:40-50): mean and
POPULATION standard deviation over the time axis, [mean | std].  The function is executed from the file where it lies
(/root/reference, build container only; nothing is copied) on seeded inputs with a single all-ones posterior class, so
that its phone-weighted block (m, s) is the same statistic computed by explicit loops (+1e-16 under the root).
Cases include a constant channel (variance 0 -> the 1e-12 floor of model/pooling.py:46-47 decides the std) and a
large-offset channel (cancellation in one-pass E[x^2]-E[x]^2 formulations)."""
import os
import re
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/model/test_utils.py"


def load_reference():
    src = open(REF).read()
    m = re.search(r"^def compute_phone_pooling\(.*?(?=^def )", src, re.S | re.M)
    mod = types.ModuleType("ref_test_utils_pool")
    mod.np = np
    exec(compile(m.group(0), REF, "exec"), mod.__dict__)
    return mod.compute_phone_pooling


def main():
    fn = load_reference()
    rng = np.random.default_rng(11)
    out = {}
    for name, b, l, c in (("small", 3, 37, 24), ("relu_like", 2, 286, 60), ("offset", 2, 120, 16)):
        x = rng.standard_normal((b, l, c))
        if name == "relu_like":
            x = np.maximum(x, 0.0)                      # post-ReLU: non-negative, mean ~ std
            x[:, :, 5] = 0.0                            # dead channel: variance exactly 0
        if name == "offset":
            x[:, :, 3] += 1.0e3                         # mean >> std
            x[:, :, 7] = 2.5                            # constant non-zero channel
        res = fn(x, np.ones((b, l, 1)))
        out[name + "/x"] = x
        out[name + "/p1"] = res[:, :2 * c]              # [mean | np.std]
        out[name + "/loops"] = res[:, 2 * c:]           # the same statistic from the explicit loops (+1e-16 under the root)
    np.savez_compressed(os.path.join(HERE, "statpool_golden.npz"), **out)
    print("wrote statpool_golden.npz")


if __name__ == "__main__":
    main()
