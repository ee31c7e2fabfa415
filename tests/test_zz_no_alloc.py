"""GPU: xv_stat_pooling_noalloc (include/xvector.h) against xv_stat_pooling_dev through the C-ABI.

The entry point has no counterpart in the reference: statistics
pooling whose [n, 2C] workspace is a file-scope __device__ array instead of a cudaMalloc'ed ctx buffer.  Named
test_zz_* so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

from util import bf16_round

pytestmark = pytest.mark.gpu

MAX_FLOATS = 512 * 1024          # XV_NOALLOC_POOL_MAX_FLOATS


@pytest.fixture(scope="module")
def eng():
    import xvector_b200
    e = xvector_b200.Engine({"network_type": "tdnn", "pooling_type": "statistics_pooling"}, 8, precision="bf16")
    yield e
    e.close()


def _pool_both(eng, x, lens, dtype):
    """-> (xv_stat_pooling_dev result, xv_stat_pooling_noalloc result), each [n, 2C] fp32, same device inputs."""
    from xvector_b200 import _lib
    n, c = len(lens), x.shape[1]
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    esz = 2 if dtype == "bf16" else 4
    d_x = eng.device_alloc(x.size * esz)
    d_cu = eng.device_alloc(cu.nbytes)
    d_ref = eng.device_alloc(n * 2 * c * 4)
    d_new = eng.device_alloc(n * 2 * c * 4)
    try:
        if dtype == "bf16":
            eng.upload_bf16(d_x, x)
        else:
            eng.h2d(d_x, x)
        eng.h2d(d_cu, cu)
        code = _lib.DTYPE_BF16 if dtype == "bf16" else _lib.DTYPE_F32
        # poison the outputs so that a call that wrote nothing cannot pass
        junk = np.full((n, 2 * c), np.nan, np.float32)
        eng.h2d(d_ref, junk)
        eng.h2d(d_new, junk)
        eng._check(eng.lib.xv_stat_pooling_dev(eng._ctx, d_x, d_cu, n, c, code, 0, d_ref))
        eng._check(eng.lib.xv_stat_pooling_noalloc(eng._ctx, d_x, d_cu, n, c, code, C.c_int64(int(cu[-1])), d_new))
        ref = np.empty((n, 2 * c), np.float32)
        new = np.empty((n, 2 * c), np.float32)
        eng.d2h(ref, d_ref)          # synchronous copies on the ctx stream: both launches have finished
        eng.d2h(new, d_new)
    finally:
        for p in (d_x, d_cu, d_ref, d_new):
            eng.device_free(p)
    return ref, new


def _assert_same(ref, new, c):
    assert np.isfinite(ref).all() and np.isfinite(new).all()
    # the tolerances of tests/test_gpu_kernels.py::test_stat_pooling_ragged_fp32_and_bf16 (fp32 atomics: the order of
    # the partial sums differs from launch to launch, so two runs of the SAME entry point already differ in the last bits)
    np.testing.assert_allclose(new[:, :c], ref[:, :c], rtol=2e-6, atol=1e-6)
    np.testing.assert_allclose(new[:, c:], ref[:, c:], rtol=2e-4, atol=2e-6)


def test_noalloc_pooling_matches_dev_entry_point_on_two_calls_in_a_row(eng):
    rng = np.random.default_rng(11)
    # call 1: 9 ragged utterances x 1500 channels (the x-vector's tdnn5 width), bf16 frames
    lens1 = np.array([1, 2, 37, 300, 1111, 64, 5, 800, 129], np.int32)
    x1 = bf16_round(np.abs(rng.standard_normal((int(lens1.sum()), 1500)).astype(np.float32)) + 3.0)
    ref1, new1 = _pool_both(eng, x1, lens1, "bf16")
    _assert_same(ref1, new1, 1500)
    # call 2, straight after: fewer utterances and another width, fp32 -- the shared array still holds call 1's sums
    lens2 = np.array([411, 7, 90], np.int32)
    x2 = rng.standard_normal((int(lens2.sum()), 512)).astype(np.float32) * 4.0 - 1.0
    x2[:, 3] = 2.5                                   # constant channel -> variance floor
    ref2, new2 = _pool_both(eng, x2, lens2, "fp32")
    _assert_same(ref2, new2, 512)
    # and the first problem once more: same answer as before (nothing of call 2 leaks into it)
    ref3, new3 = _pool_both(eng, x1, lens1, "bf16")
    _assert_same(ref3, new3, 1500)
    _assert_same(new1, new3, 1500)
    # independent check of one utterance against float64
    seg = x1[int(lens1[:3].sum()):int(lens1[:4].sum())].astype(np.float64)
    np.testing.assert_allclose(new1[3, :1500], seg.mean(axis=0), rtol=2e-6, atol=1e-6)


def test_noalloc_pooling_at_and_beyond_the_limit(eng):
    from xvector_b200 import _lib
    rng = np.random.default_rng(12)
    c = 1024
    n_max = MAX_FLOATS // (2 * c)                    # 256 utterances: exactly the scratch size
    lens = rng.integers(1, 9, n_max).astype(np.int32)
    x = rng.standard_normal((int(lens.sum()), c)).astype(np.float32)
    ref, new = _pool_both(eng, x, lens, "fp32")
    _assert_same(ref, new, c)
    # one utterance more than fits: rejected before anything is launched, with a message naming the limit
    d = eng.device_alloc(1024)
    try:
        rc = eng.lib.xv_stat_pooling_noalloc(eng._ctx, d, d, n_max + 1, c, _lib.DTYPE_F32, C.c_int64(n_max + 1), d)
    finally:
        eng.device_free(d)
    assert rc == _lib.XV_EINVAL
    msg = eng.lib.xv_last_error(eng._ctx).decode()
    assert "XV_NOALLOC_POOL_MAX_FLOATS" in msg and str(MAX_FLOATS) in msg
    # empty job: nothing to do, no error
    d = eng.device_alloc(1024)
    try:
        assert eng.lib.xv_stat_pooling_noalloc(eng._ctx, d, d, 0, c, _lib.DTYPE_F32, C.c_int64(0), d) == 0
    finally:
        eng.device_free(d)
