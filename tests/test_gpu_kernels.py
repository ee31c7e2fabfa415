"""GPU: kernel-level parity through the C-ABI -- the tcgen05 GEMM (splice-in-load), pooling reductions."""
import ctypes as C
import os

import numpy as np
import pytest

from util import bf16_round, tensor_rel

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module")
def eng():
    import xvector_b200
    e = xvector_b200.Engine({"network_type": "tdnn", "pooling_type": "statistics_pooling"}, 8, precision="bf16")
    yield e
    e.close()


def _run_gemm(eng, a, wt, taps, scale, shift, neg_slope, m_rows, dilation=1):
    """a [rows, cin] fp32 (bf16-representable), wt [n, taps*cin]; returns bf16 output as fp32 [m_rows, n]."""
    rows, cin = a.shape
    n = wt.shape[0]
    d_a = eng.device_alloc(a.size * 2)
    d_w = eng.device_alloc(wt.size * 2)
    d_s = eng.device_alloc(n * 4)
    d_h = eng.device_alloc(n * 4)
    d_o = eng.device_alloc(rows * n * 2)
    try:
        eng.upload_bf16(d_a, a)
        eng.upload_bf16(d_w, wt)
        eng.h2d(d_s, scale.astype(np.float32))
        eng.h2d(d_h, shift.astype(np.float32))
        eng._check(eng.lib.xv_gemm_bf16(eng._ctx, d_a, rows, cin, taps, dilation, d_w, n, d_s, d_h, C.c_float(neg_slope),
                                        d_o, n, m_rows))
        raw = np.empty((rows, n), np.uint16)
        eng.d2h(raw, d_o)
    finally:
        for p in (d_a, d_w, d_s, d_h, d_o):
            eng.device_free(p)
    return (raw.astype(np.uint32) << 16).view(np.float32)[:m_rows]


def _ref_gemm(a, wt, taps, scale, shift, neg_slope, m_rows, dilation=1):
    """out[m] = sum_j a[m + j*dilation] . wt[:, tap j]; rows outside [0, rows) read as zero (a negative dilation is the
    data-gradient mode of the training step: taps reach backwards)."""
    rows, cin = a.shape
    pad = (taps - 1) * abs(dilation)
    a64 = np.concatenate([np.zeros((pad, cin)), a.astype(np.float64), np.zeros((pad, cin))], axis=0)
    acc = np.zeros((m_rows, wt.shape[0]))
    for j in range(taps):
        lo = pad + j * dilation
        acc += a64[lo:lo + m_rows] @ wt[:, j * cin:(j + 1) * cin].astype(np.float64).T
    y = acc * scale + shift
    return np.where(y >= 0, y, neg_slope * y)


@pytest.mark.parametrize("rows,cin,taps,n,dil", [
    (128, 64, 1, 256, 1),          # one tile, one k-block: the smallest possible problem
    (256, 512, 1, 512, 1),         # tdnn4 shape
    (1000, 512, 5, 512, 1),        # tdnn2: spliced context 5, ragged last tile
    (777, 512, 7, 512, 1),         # tdnn3
    (300, 512, 1, 1536, 1),        # tdnn5 (N padded 1500 -> 1536)
    (640, 192, 1, 512, 1),         # tdnn1 on pre-spliced rows (K = 150 -> 192)
    (20000, 512, 5, 512, 1),       # more tiles than SMs: persistent loop, both TMEM stages, ring wrap
    (900, 512, 3, 256, 2),         # dilation 2 (Kaldi-style context; the reference uses 1)
    (1000, 512, 5, 512, -1),       # tap step -1: the data-gradient GEMM of the training step (negative TMA rows = 0)
    (700, 1536, 1, 512, -1),       # tdnn5's data gradient: K = 1536
])
def test_tcgen05_gemm_matches_fp64(eng, rows, cin, taps, n, dil):
    rng = np.random.default_rng(rows + cin + taps + n)
    a = bf16_round(rng.standard_normal((rows, cin)).astype(np.float32))
    wt = bf16_round((rng.standard_normal((n, taps * cin)) / np.sqrt(taps * cin)).astype(np.float32))
    scale = rng.uniform(0.5, 1.5, n)
    shift = rng.standard_normal(n) * 0.1
    m_rows = rows - 3            # leave a partial row tile
    out = _run_gemm(eng, a, wt, taps, scale, shift, 0.2, m_rows, dil)
    ref = _ref_gemm(a, wt, taps, scale, shift, 0.2, m_rows, dil)
    # inputs are exact in bf16, accumulation is fp32, the only rounding is the bf16 store (2^-9 relative)
    err = np.abs(out - ref)
    assert (err <= 2.0 ** -8 * np.abs(ref) + 1e-3).all(), "max abs err %.4g (rel %.3g)" % (err.max(), tensor_rel(out, ref))


def test_stat_pooling_ragged_fp32_and_bf16(eng):
    rng = np.random.default_rng(5)
    for c in (1500, 512, 96, 7):
        lens = np.array([1, 2, 37, 300, 1111, 64, 5], np.int32)
        cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
        x = np.abs(rng.standard_normal((cu[-1], c)).astype(np.float32)) + 3.0       # mean >> std stresses cancellation
        x[:, 0] = 2.5                                                                  # constant channel -> floor
        for dtype, tol in (("fp32", 2e-6), ("bf16", 2e-6)):
            xin = x if dtype == "fp32" else bf16_round(x)
            out = eng.stat_pooling(xin, cu, dtype=dtype)
            for u in range(len(lens)):
                seg = xin[cu[u]:cu[u + 1]].astype(np.float64)
                mean = seg.mean(axis=0)
                var = ((seg - mean) ** 2).mean(axis=0)
                var = np.where(var <= 1e-12, 1e-12, var)
                ref = np.concatenate([mean, np.sqrt(var)])
                got = out[u].astype(np.float64)
                np.testing.assert_allclose(got[:c], ref[:c], rtol=tol, atol=1e-6)
                np.testing.assert_allclose(got[c:], ref[c:], rtol=2e-4, atol=2e-6)


def test_stat_pooling_large_matches_oracle(eng):
    from oracle import numpy_ref
    rng = np.random.default_rng(9)
    lens = np.array([6000, 200, 3333], np.int32)
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    x = np.maximum(rng.standard_normal((cu[-1], 1500)).astype(np.float32), 0)
    out = eng.stat_pooling(x, cu)
    for u in range(3):
        ref = numpy_ref.statistics_pooling(x[None, cu[u]:cu[u + 1]])[0]
        assert tensor_rel(out[u], ref) < 3e-5      # fp32 one-pass sums over 6000 rows; end-to-end budget is 1e-4


def test_stat_pooling_matches_reference_golden(eng):
    """Golden vectors = [np.mean | np.std] of the reference's compute_phone_pooling (model/test_utils.py:1110-1112);
    channels whose variance is exactly 0 take the 1e-12 floor of model/pooling.py:46-47.  fp32 kernel: errors are judged
    on the channel's scale max(|mean|, std) (one channel carries a 1e3 offset, i.e. fp32 inputs quantised at 6e-5)."""
    g = np.load(os.path.join(GOLDEN, "statpool_golden.npz"))
    for name in sorted({k.split("/")[0] for k in g.files}):
        x, p1 = g[name + "/x"], g[name + "/p1"]
        b, l, c = x.shape
        cu = (np.arange(b + 1) * l).astype(np.int32)
        out = eng.stat_pooling(x.reshape(b * l, c).astype(np.float32), cu, dtype="fp32").astype(np.float64)
        scale = np.maximum(np.abs(p1[:, :c]), p1[:, c:])
        scale = np.maximum(scale, 1e-3)
        assert (np.abs(out[:, :c] - p1[:, :c]) / scale).max() < 2e-6, name
        live = p1[:, c:] > 1e-5
        assert (np.abs(out[:, c:] - p1[:, c:])[live] / scale[live]).max() < 2e-5, name
        if (~live).any():
            np.testing.assert_allclose(out[:, c:][~live], 1e-6, rtol=1e-3)


def test_attention_pooling_matches_reference_golden(eng):
    """Golden vectors = outputs of the reference's compute_self_attention (model/test_utils.py:321-392)."""
    g = np.load(os.path.join(GOLDEN, "attention_golden.npz"))
    for name in sorted({k.split("/")[0] for k in g.files}):
        sv, sk, sc = [bool(v) for v in g[name + "/flags"]]
        value, key, query = g[name + "/value"], g[name + "/key"], g[name + "/query"]
        b, l, cv = value.shape
        cu = (np.arange(b + 1) * l).astype(np.int32)
        att, w = eng.attention_pooling(value.reshape(b * l, cv), key.reshape(b * l, -1), query, cu, sv, sk, sc, True)
        assert tensor_rel(att, g[name + "/att"]) < 2e-5, name
        for u in range(b):
            np.testing.assert_allclose(w[u].sum(axis=1), 1.0, rtol=1e-5)


def test_attention_pooling_ragged(eng):
    from oracle import numpy_ref
    rng = np.random.default_rng(21)
    lens = np.array([50, 7, 400], np.int32)
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    value = np.abs(rng.standard_normal((cu[-1], 60))).astype(np.float32)
    key = rng.standard_normal((cu[-1], 48)).astype(np.float32)
    for h, sv, sk in ((1, True, True), (4, True, True), (3, False, True), (2, True, False)):
        q = (0.3 * rng.standard_normal((h, 48 // h if sk else 48))).astype(np.float32)
        att = eng.attention_pooling(value, key, q, cu, sv, sk, True)
        for u in range(3):
            ref, _ = numpy_ref.attention_core(value[None, cu[u]:cu[u + 1]], key[None, cu[u]:cu[u + 1]], q, sv, sk, True)
            assert tensor_rel(att[u], ref[0]) < 2e-5
