"""GPU (>= 2 devices): the data-parallel training step -- two ranks, one ncclAllReduce -- against the oracle's
two-tower semantics (model/trainer.py:1133-1248).  Skipped on a single-GPU box."""
import multiprocessing as mp
import os
import tempfile

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _n_gpus():
    try:
        out = os.popen("nvidia-smi -L 2>/dev/null").read()
        return sum(1 for l in out.splitlines() if l.startswith("GPU "))
    except Exception:
        return 0


def _rank_main(rank, world, path, q):
    import xvector_b200
    from oracle import synth
    from xvector_b200 import train_dp
    cfg = dict(synth.TDNN_AMSOFTMAX_CONFIG)
    dim, speakers = 24, 101
    w = synth.add_softmax_weights(synth.make_tdnn_weights(dim, cfg, seed=31), speakers)
    x, y = synth.config5_batch(8, 50, dim, speakers, seed=32)
    eng = xvector_b200.Engine(cfg, dim, precision="fp32", device=rank)
    eng.set_weights(w)
    eng.train_init(speakers)
    uid = train_dp.exchange_unique_id_file(eng, rank, world, path)
    eng.comm_init(uid, rank, world)
    xs, ys = train_dp.slice_batch(x, y, rank, world)
    loss, total = eng.train_step(xs, ys, 0.05, 0)
    names = ["tdnn/tdnn3_conv/kernel", "tdnn/tdnn6_dense/kernel", "softmax/output/kernel", "tdnn/tdnn5_bn/gamma"]
    out = {"loss": loss, "total": total}
    for n in names:
        out["g:" + n] = eng.train_grad(n, w[n].shape)
        out["w:" + n] = eng.train_weight(n, w[n].shape)
    out["mm"] = eng.train_weight("tdnn/tdnn1_bn/moving_mean", (512,))
    eng.close()
    q.put((rank, out))


@pytest.mark.skipif(_n_gpus() < 2, reason="needs 2 GPUs (run with gpurun --gpus 2)")
def test_two_rank_step_matches_two_tower_oracle():
    from oracle import numpy_ref, synth, train_ref
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    path = os.path.join(tempfile.mkdtemp(), "nccl_uid")
    procs = [ctx.Process(target=_rank_main, args=(r, 2, path, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=300) for _ in procs)
    for p in procs:
        p.join(60)
    cfg = dict(synth.TDNN_AMSOFTMAX_CONFIG)
    w = synth.add_softmax_weights(synth.make_tdnn_weights(24, cfg, seed=31), 101)
    x, y = synth.config5_batch(8, 50, 24, 101, seed=32)
    new_ref, info = train_ref.train_step(x, y, w, numpy_ref.OracleParams(**cfg), 0.05, 0, num_towers=2)
    for r in range(2):
        assert abs(res[r]["loss"] - info["loss"]) <= 1e-5 * abs(info["loss"])      # mean of the tower losses on every rank
        for k, v in res[r].items():
            if k.startswith("g:"):
                ref = info["grads"][k[2:]].reshape(v.shape)
                assert np.linalg.norm(v - ref) <= 5e-3 * np.linalg.norm(ref), k
            elif k.startswith("w:"):
                ref = new_ref[k[2:]].reshape(v.shape)
                assert np.linalg.norm(v - ref) <= 5e-3 * np.linalg.norm(ref), k
        # moving averages come from tower 0 on EVERY rank (model/trainer.py:1183-1190)
        np.testing.assert_allclose(res[r]["mm"], new_ref["tdnn/tdnn1_bn/moving_mean"], rtol=1e-5, atol=1e-7)
    # replicas stay bit-identical
    for k in res[0]:
        if k.startswith("w:"):
            np.testing.assert_array_equal(res[0][k], res[1][k])
