"""Data-parallel plumbing for the training step (BASELINE.json config 5): what TrainerMGPU.build does with towers
(model/trainer.py:1133-1223), restated for one process per GPU.

  * the GLOBAL batch is cut into `world` equal slices (model/trainer.py:1133-1134, 1148-1150);
  * every rank runs xv_train_step on its slice with its own BN batch statistics (:1183-1190);
  * gradients are averaged by ONE ncclAllReduce over the flat gradient buffer inside xv_train_step (== add_n/N,
    :1240-1244); BN moving averages and the reported loss ride in the tail of the same buffer;
  * weights are replicated (the reference shards them over GPUs as "parameter servers": mgpu_tools.py:122-153).

No torch types cross into the library: the 128-byte ncclUniqueId is created by rank 0 through the C-ABI and handed
to the other ranks by whatever the launcher offers (a torch.distributed broadcast, a file, an env var)."""
import os
import time

import numpy as np


def slice_batch(features, labels, rank, world):
    """features [B, T, D], labels [B] -> this rank's tower slice (model/trainer.py:1148-1150)."""
    b = features.shape[0]
    if b % world != 0:
        raise ValueError("the global batch (%d) must be divisible by the number of GPUs (%d)" % (b, world))
    per = b // world
    return features[rank * per:(rank + 1) * per], labels[rank * per:(rank + 1) * per]


def exchange_unique_id_file(engine, rank, world, path, timeout_s=120.0):
    """Rank 0 writes the ncclUniqueId to `path`; the other ranks poll for it."""
    if world == 1:
        return b"\0" * 128
    if rank == 0:
        uid = engine.comm_unique_id()
        tmp = path + ".tmp"
        with open(tmp, "wb") as f:
            f.write(uid)
        os.replace(tmp, path)
        return uid
    t0 = time.time()
    while not os.path.exists(path):
        if time.time() - t0 > timeout_s:
            raise RuntimeError("timed out waiting for the NCCL unique id at %s" % path)
        time.sleep(0.01)
    with open(path, "rb") as f:
        return f.read()


def exchange_unique_id_torch(engine, rank, world):
    """Same through an already initialised torch.distributed process group (any backend)."""
    import torch
    import torch.distributed as dist
    if world == 1:
        return b"\0" * 128
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    buf = torch.zeros(128, dtype=torch.uint8, device=dev)
    if rank == 0:
        buf = torch.tensor(list(engine.comm_unique_id()), dtype=torch.uint8, device=dev)
    dist.broadcast(buf, src=0)
    return bytes(buf.cpu().numpy().tolist())


def make_replica(cfg, dim, weights, num_speakers, rank, world, uid, device=None):
    """Engine with training state + communicator for this rank."""
    from .engine import Engine
    eng = Engine(cfg, dim, precision="fp32", device=rank if device is None else device)
    eng.set_weights(weights)
    eng.train_init(num_speakers)
    eng.comm_init(uid, rank, world)
    return eng


def train_flops(batch, frames, dim, speakers, pool_nodes=1500, last_nodes=512):
    """Algorithmic FLOPs of one step (forward + dgrad + wgrad ~ 3x forward GEMMs)."""
    t = float(frames)
    fwd = (2 * 512 * (5 * dim * (t - 4) + 2560 * (t - 8) + 3584 * (t - 14) + 512 * (t - 14)) + 2 * 512 * pool_nodes * (t - 14)
           + 2 * (2 * pool_nodes) * 512 + 2 * 512 * last_nodes + 2 * last_nodes * speakers)
    first = 2 * 512 * 5 * dim * (t - 4)                     # the first layer has no dgrad
    return batch * (3 * fwd - first)
