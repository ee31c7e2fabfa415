"""Host-side scheduling for extraction: chunk planning (extract.py:68-78) and sharding of utterances across GPUs.

Utterances are packed back to back (no padding), so length bucketing is not needed for efficiency inside one GPU;
across GPUs the work is balanced by longest-processing-time-first on the algorithmic FLOPs of each utterance.
There is no collective on this path: every rank extracts its own shard."""
import math

import numpy as np


def chunk_plan(length, chunk_size):
    """[(start, size), ...] for an utterance longer than chunk_size (extract.py:68-78, Python-2 integer
    division for chunk_size / 2)."""
    half = chunk_size // 2
    num_chunks = int(math.ceil(float(length - chunk_size) / half)) + 1
    plan = []
    for i in range(num_chunks):
        start = i * half
        size = chunk_size if length - start > chunk_size else length - start
        plan.append((start, size))
    return plan


def plan_arrays(lengths, chunk_size, min_chunk_size):
    """chunk_plan over a whole batch without a Python loop: -> (owners, starts, sizes), int64 arrays with one entry per piece
    in utterance order; utterances shorter than min_chunk_size own no piece (extract.py:65-78)."""
    n = np.asarray(lengths, np.int64)
    keep = n >= min_chunk_size
    half = chunk_size // 2
    chunked = keep & (n > chunk_size)
    pieces = keep.astype(np.int64)
    if chunked.any():
        pieces = np.where(chunked, -(-(n - chunk_size) // half) + 1, pieces)
    owners = np.repeat(np.arange(len(n), dtype=np.int64), pieces)
    first = np.cumsum(pieces) - pieces
    starts = (np.arange(owners.size, dtype=np.int64) - first[owners]) * half
    left = n[owners] - starts
    return owners, starts, np.where(left > chunk_size, chunk_size, left)


def utterance_flops(t, d, pool_nodes=1500):
    """Algorithmic FLOPs up to tdnn6_dense."""
    t = np.asarray(t, np.float64)
    return (2 * 512 * (5 * d * (t - 4) + 2560 * (t - 8) + 3584 * (t - 14) + 512 * (t - 14))
            + 2 * 512 * pool_nodes * (t - 14) + 2 * (2 * pool_nodes) * 512)


def shard_lpt(lengths, world_size, dim=30):
    """Longest-processing-time-first partition.  Returns a list (per rank) of utterance index arrays; within a
    rank utterances are ordered by decreasing length (a deterministic length bucketing)."""
    lengths = np.asarray(lengths)
    cost = utterance_flops(lengths, dim)
    order = np.argsort(-cost, kind="stable")
    loads = np.zeros(world_size)
    shards = [[] for _ in range(world_size)]
    for i in order:
        r = int(np.argmin(loads))
        shards[r].append(int(i))
        loads[r] += cost[i]
    return [np.asarray(s, dtype=np.int64) for s in shards]


def resnet_layer_flops(t, f, blocks=(2, 2, 2, 2), time_stride=False):
    """Algorithmic FLOPs of one resnet_18 utterance of t frames x f bins up to tdnn6_dense, per kernel family
    (2 FLOPs per MAC, SAME convolutions count every output position; SURVEY.md section 8(d):
    65.46 GFLOP at (400, 40), 104.25 GFLOP at (400, 64))."""
    out = {"stem": 2.0 * t * f * 9 * 64}
    cin, tt, ff = 64, t, f
    for stage, filt in enumerate((64, 128, 256, 512)):
        if stage > 0:
            ff = ff // 2
            if time_stride:
                tt = (tt + 1) // 2
        pos = float(tt) * ff
        fam = "conv.c%d" % filt
        fl = 2 * pos * 9 * cin * filt + 2 * pos * cin * filt + 2 * pos * 9 * filt * filt      # conv0, 1x1 shortcut, conv1
        fl += (blocks[stage] - 1) * 2 * (2 * pos * 9 * filt * filt)
        out[fam] = out.get(fam, 0.0) + fl
        cin = filt
    out["gemm.conv5"] = 2.0 * tt * ff * 512 * 512
    out["gemm.dense1"] = 2.0 * tt * 512 * 512
    out["gemm.dense2"] = 2.0 * tt * 512 * 1500
    out["gemm.tdnn6"] = 2.0 * 3000 * 512
    return out


def resnet_flops(t, f, blocks=(2, 2, 2, 2), time_stride=False):
    return sum(resnet_layer_flops(t, f, blocks, time_stride).values())
