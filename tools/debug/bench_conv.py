"""Times xv_conv2d_bf16_async on one shape (profile scope 'conv2d'): python tools/debug/bench_conv.py nb t f cin cout [res] [reps]."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import xvector_b200  # noqa: E402


def main():
    nb, t, f, cin, cout = (int(v) for v in sys.argv[1:6])
    res = int(sys.argv[6]) if len(sys.argv) > 6 else 0
    reps = int(sys.argv[7]) if len(sys.argv) > 7 else 20
    eng = xvector_b200.Engine({"network_type": "tdnn", "pooling_type": "statistics_pooling"}, 8, precision="bf16")
    rng = np.random.default_rng(0)
    x = rng.standard_normal((nb, t, f, cin)).astype(np.float32)
    w = (rng.standard_normal((cout, 9 * cin)) / np.sqrt(9 * cin)).astype(np.float32)
    d_x, d_w, d_o = eng.device_alloc(x.size * 2), eng.device_alloc(w.size * 2), eng.device_alloc(nb * t * f * cout * 2)
    d_s, d_h, d_l = (eng.device_alloc(cout * 4) for _ in range(3))
    eng.upload_bf16(d_x, x)
    eng.upload_bf16(d_w, w)
    eng.h2d(d_s, np.ones(cout, np.float32))
    eng.h2d(d_h, np.zeros(cout, np.float32))
    eng.h2d(d_l, np.zeros(cout, np.float32))
    d_r = None
    if res:
        d_r = eng.device_alloc(nb * t * f * cout * 2)
        eng.upload_bf16(d_r, rng.standard_normal((nb, t, f, cout)).astype(np.float32))

    def run(n):
        for _ in range(n):
            eng._check(eng.lib.xv_conv2d_bf16_async(eng._ctx, d_x, nb, t, f, cin, d_w, 3, 3, 1, 1, cout, d_s, d_h, d_l, d_r, d_o))
        eng.sync()
    run(5)
    eng.profile_enable(True)
    run(reps)
    ms, cnt = eng.profile_query("conv2d")
    flops = 2.0 * nb * t * f * 9 * cin * cout
    print("shape nb=%d t=%d f=%d cin=%d cout=%d res=%d halo=%s: %.1f us/launch, %.0f TFLOP/s" % (
        nb, t, f, cin, cout, res, os.environ.get("XV_CONV_HALO", "1"),
        ms / cnt * 1e3, flops / (ms / cnt * 1e-3) / 1e12))
    eng.close()


if __name__ == "__main__":
    main()
