"""GPU: the tcgen05 weight-gradient kernel (MN-major UMMA operands, split-K) against fp64 numpy, through the C-ABI."""
import numpy as np
import pytest

from util import bf16_round, tensor_rel

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    import xvector_b200
    e = xvector_b200.Engine({"network_type": "tdnn", "pooling_type": "statistics_pooling"}, 8, precision="bf16")
    yield e
    e.close()


def _wgrad(eng, a, dz, rows, taps, cin, n_valid, ldo):
    bufs = []

    def dev(nbytes):
        p = eng.device_alloc(nbytes)
        bufs.append(p)
        return p
    try:
        d_a, d_z, d_o = dev(a.size * 2), dev(dz.size * 2), dev(taps * cin * ldo * 4)
        eng.upload_bf16(d_a, a)
        eng.upload_bf16(d_z, dz)
        eng._check(eng.lib.xv_wgrad_bf16_async(eng._ctx, d_a, a.shape[0], a.shape[1], d_z, dz.shape[1], rows, taps, cin,
                                               n_valid, d_o, ldo))
        eng.sync()
        out = np.empty((taps * cin, ldo), np.float32)
        eng.d2h(out, d_o)
    finally:
        for p in bufs:
            eng.device_free(p)
    return out


@pytest.mark.parametrize("rows,taps,cin,lda,n,ldz,ldo", [
    (1000, 1, 512, 512, 512, 512, 512),        # dense layer, rows not a multiple of 64
    (777, 5, 30, 64, 512, 512, 512),           # tdnn1: 30-dim features, 5 taps, channel mask inside one M tile
    (3000, 7, 512, 512, 512, 512, 512),        # tdnn3: 7 taps, 28 M tiles
    (500, 1, 512, 512, 1500, 1536, 1500),      # tdnn5: 1500 outputs (N mask, ldo not a multiple of 256)
    (64, 1, 128, 128, 256, 256, 256),          # a single K block
    (20000, 5, 192, 192, 320, 320, 320),       # many splits, cin and n not multiples of the tile
])
def test_wgrad_matches_fp64(eng, rows, taps, cin, lda, n, ldz, ldo):
    rng = np.random.default_rng(rows + taps * 7 + cin)
    a = np.zeros((rows, lda), np.float32)
    a[:, :cin] = bf16_round(rng.standard_normal((rows, cin)).astype(np.float32))
    dz = np.zeros((rows, ldz), np.float32)
    dz[:, :n] = bf16_round(rng.standard_normal((rows, n)).astype(np.float32))
    out = _wgrad(eng, a, dz, rows, taps, cin, n, ldo)
    a64, dz64 = a[:, :cin].astype(np.float64), dz[:, :n].astype(np.float64)
    ref = np.zeros((taps * cin, n))
    for j in range(taps):
        ref[j * cin:(j + 1) * cin] = a64[j:].T @ dz64[:rows - j]
    assert tensor_rel(out[:, :n], ref) <= 1e-5, "rel %.3e" % tensor_rel(out[:, :n], ref)
    err = np.abs(out[:, :n] - ref)
    assert err.max() <= 1e-3 * np.sqrt(rows), "max abs err %.4g" % err.max()
    assert not out[:, n:].any()
