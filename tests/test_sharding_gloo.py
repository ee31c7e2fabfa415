"""CPU, world_size 2 over gloo: the multi-GPU extraction path has NO data-path collective -- every rank takes
its LPT shard of the global utterance list and the only communication is the timing/barrier plumbing of bench.py.
This checks that the shards partition the list and that a gathered result is placed at the right indices."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, ret):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import xvector_b200  # noqa: F401
    from xvector_b200 import scheduler
    from oracle import synth
    lengths = synth.config2_lengths(256 * world, seed=11)
    shards = scheduler.shard_lpt(lengths, world)
    mine = shards[rank]
    # stand-in for the per-rank device work: a deterministic function of (index, length)
    local = torch.tensor([float(i) * 3 + float(lengths[i]) for i in mine], dtype=torch.float64)
    frames = torch.tensor([float(lengths[mine].sum())], dtype=torch.float64)
    dist.all_reduce(frames, op=dist.ReduceOp.SUM)              # bench.py: total frames processed by the job
    t = torch.tensor([0.1 * (rank + 1)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)                   # bench.py: max over ranks of the timed region
    sizes = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(sizes, torch.tensor([len(mine)], dtype=torch.int64))
    if rank == 0:
        ret["frames"] = float(frames.item())
        ret["tmax"] = float(t.item())
        ret["sizes"] = [int(s.item()) for s in sizes]
        ret["expect_frames"] = float(lengths.sum())
        ret["n"] = len(lengths)
    ret["local%d" % rank] = (mine.tolist(), local.tolist())
    dist.barrier()
    dist.destroy_process_group()


def test_world_size_2_sharding_no_collective_on_data_path():
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    port = 29500 + (os.getpid() % 500)
    mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)
    assert ret["frames"] == ret["expect_frames"]
    assert abs(ret["tmax"] - 0.2) < 1e-12
    assert sum(ret["sizes"]) == ret["n"]
    seen = np.zeros(ret["n"], bool)
    for r in range(world):
        idx, vals = ret["local%d" % r]
        assert not seen[idx].any()
        seen[idx] = True
    assert seen.all()
