"""CPU, world_size 2 over gloo: host-side logic of the data-parallel training step -- tower slices of the global
batch (model/trainer.py:1148-1150), hand-off of the 128-byte NCCL unique id, and the exchange semantics
(sum over ranks * 1/N == add_n/N, moving statistics and loss riding in the same flat buffer) against the oracle's
two-tower step."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class _FakeEngine(object):
    def comm_unique_id(self):
        return bytes(range(128))


def _worker(rank, world, port, ret):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import xvector_b200  # noqa: F401
    from xvector_b200 import train_dp
    from oracle import numpy_ref, synth, train_ref
    uid = train_dp.exchange_unique_id_torch(_FakeEngine(), rank, world)
    cfg = dict(synth.TDNN_AMSOFTMAX_CONFIG)
    w = synth.add_softmax_weights(synth.make_tdnn_weights(16, cfg, seed=5), 23)
    x, y = synth.config5_batch(8, 36, 16, 23, seed=2)
    xs, ys = train_dp.slice_batch(x, y, rank, world)
    tower = train_ref.tower_forward_backward(xs, ys, w, numpy_ref.OracleParams(**cfg), 0)
    # the flat buffer of xv_train_step: [gradients | moving statistics (rank 0 only) | loss]
    names = sorted(tower["grads"])
    mov = sorted(tower["moving"])
    flat = np.concatenate([tower["grads"][k].ravel() for k in names] +
                          [tower["moving"][k].ravel() * (1.0 if rank == 0 else 0.0) for k in mov] + [np.array([tower["loss"]])])
    t = torch.from_numpy(flat)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)                  # the ONE collective of the step
    flat = t.numpy()
    n_grad = sum(tower["grads"][k].size for k in names)
    if rank == 0:
        _, info = train_ref.train_step(x, y, w, numpy_ref.OracleParams(**cfg), 0.01, 0, num_towers=world)
        ref = np.concatenate([info["grads"][k].ravel() for k in names])
        ret["grad_err"] = float(np.abs(flat[:n_grad] / world - ref).max() / np.abs(ref).max())
        ret["loss_err"] = abs(flat[-1] / world - info["loss"])
        ret["mov_err"] = float(np.abs(flat[n_grad:-1] - np.concatenate([tower["moving"][k].ravel() for k in mov])).max())
    ret["uid%d" % rank] = uid
    ret["n%d" % rank] = (xs.shape[0], int(ys[0]))
    dist.barrier()
    dist.destroy_process_group()


def test_world_size_2_training_exchange():
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    port = 29600 + (os.getpid() % 300)
    mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)
    assert ret["uid0"] == ret["uid1"] == bytes(range(128))
    assert ret["n0"][0] == ret["n1"][0] == 4
    assert ret["grad_err"] < 1e-12 and ret["loss_err"] < 1e-12 and ret["mov_err"] == 0.0


def test_slice_batch_rejects_uneven_batches():
    import pytest
    import xvector_b200  # noqa: F401
    from xvector_b200 import train_dp
    with pytest.raises(ValueError):
        train_dp.slice_batch(np.zeros((7, 3, 2)), np.zeros(7), 0, 2)
    a, b = train_dp.slice_batch(np.arange(24).reshape(4, 3, 2), np.arange(4), 1, 2)
    assert a.shape == (2, 3, 2) and list(b) == [2, 3]
