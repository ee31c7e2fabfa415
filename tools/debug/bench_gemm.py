"""Times xv_gemm_bf16_async (profile scope 'gemm') on one shape: python tools/debug/bench_gemm.py M cin taps N [reps].
A small M keeps the A operand L2-resident between launches; a large one streams it from HBM."""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import xvector_b200  # noqa: E402


def main():
    m, cin, taps, n = (int(v) for v in sys.argv[1:5])
    reps = int(sys.argv[5]) if len(sys.argv) > 5 else 20
    eng = xvector_b200.Engine({"network_type": "tdnn", "pooling_type": "statistics_pooling"}, 8, precision="bf16")
    rng = np.random.default_rng(0)
    rows = m + taps
    d_a, d_w, d_o = eng.device_alloc(rows * cin * 2), eng.device_alloc(n * taps * cin * 2), eng.device_alloc(m * n * 2)
    d_s, d_h, d_l = (eng.device_alloc(n * 4) for _ in range(3))
    chunk = 1 << 16
    blk = rng.standard_normal((chunk, cin)).astype(np.float32)
    for r0 in range(0, rows, chunk):
        k = min(chunk, rows - r0)
        eng.upload_bf16(C.c_void_p(d_a.value + r0 * cin * 2), blk[:k])
    eng.upload_bf16(d_w, (rng.standard_normal((n, taps * cin)) / np.sqrt(taps * cin)).astype(np.float32))
    eng.h2d(d_s, np.ones(n, np.float32))
    eng.h2d(d_h, np.zeros(n, np.float32))
    eng.h2d(d_l, np.zeros(n, np.float32))

    def run(k):
        for _ in range(k):
            eng._check(eng.lib.xv_gemm_bf16_async(eng._ctx, d_a, rows, cin, taps, 1, d_w, n, d_s, d_h, d_l, d_o, m))
        eng.sync()
    run(5)
    eng.profile_enable(True)
    run(reps)
    ms, cnt = eng.profile_query("gemm")
    flops = 2.0 * m * taps * cin * n
    print("gemm M=%d cin=%d taps=%d N=%d cluster=%s: %.1f us/launch, %.0f TFLOP/s" % (
        m, cin, taps, n, os.environ.get("XV_GEMM_CLUSTER", "1"), ms / cnt * 1e3, flops / (ms / cnt * 1e-3) / 1e12))
    eng.close()


if __name__ == "__main__":
    main()
