#!/usr/bin/env python
"""bench.py -- x-vector extraction throughput (frames/s) of the B200-native hot path.

  python bench.py --gpus 1 --steps 10 --warmup 3              (N=1)
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
         bench.py --gpus N --steps K --warmup W               (N>1, one rank per GPU)
  python bench.py --impl reference ...                        (CPU restatement of the reference graph)

Workload (BASELINE.json configs[1]): the standard x-vector TDNN (5x512 frame layers, 1500-dim pre-pool,
statistics pooling, 512-dim tdnn6_dense embedding) on synthetic 30-dim MFCC utterances of log-uniform length
200..6000 frames, 4096 utterances per GPU (weak scaling: a global list of 4096*N utterances is LPT-sharded over
the N ranks; no collective on the data path).  A "step" = one pass of the hot path over the rank's shard.

Printed JSON line (rank 0): `value` = whole-job input frames/s with features resident in HBM, timed with CUDA
events on the library's stream, max over ranks; `e2e` = same metric through the public C-ABI call xv_forward from
pinned HOST memory (H2D of features and D2H of embeddings inside the timed region); `roofline` = the step's
algorithmic GEMM FLOPs / the step's time against the measured bf16 peak (the summed GEMM launch
durations are reported beside it); `cpu_baseline` = the fp32 torch-CPU restatement run the way the reference deploys
extraction.

The default run (workload tdnn) also takes short, bounded LEGS for the other BASELINE.json configs and attaches them
under `legs` of the same JSON line, so that one driver run puts every config under its clock:
  N = 1 : cfg0 (1000 x 300 frames: whole job + single-utterance Trainer.predict latency), cfg1_chunked
          (chunk_size = 2000 through the extractor's chunk-and-average), attention (configs[2]), resnet18 (configs[3]),
          train (configs[4], one GPU);
  N > 1 : train_dp -- the library's own NCCL communicator (xv_comm_init): weak (64 utterances per GPU) and the
          reference-faithful strong split (global 64, model/trainer.py:1133-1134) ms/step, the allreduce alone
          (bus GB/s against 900 GB/s per direction), and a parity check of the N-tower step against a committed golden
          (tests/golden/train_dp_golden.npz).
A leg that fails reports {"error": ...} and never takes the headline down.  `--no-legs` skips them.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

UTTS_PER_GPU = 4096
DIM = 30
METRIC = "xvector_extraction_frames_per_sec"
UNIT = "frames/s"


def read_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(path):
        with open(path) as f:
            p = json.load(f)
        return {"bf16_burst": p["bf16_tflops"], "bf16_sustained": p.get("bf16_tflops_sustained", p["bf16_tflops"]),
                "hbm": p["hbm_gbs"], "source": "measured"}
    return {"bf16_burst": 1590.0, "bf16_sustained": 1400.0, "hbm": 6650.0, "source": "fallback"}


class ClockSampler(threading.Thread):
    """Samples nvidia-smi SM clocks / throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.rows = []
        self.stop_flag = False
        self.proc = None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                if self.stop_flag:
                    break
                self.rows.append([c.strip() for c in line.split(",")])
        except Exception:
            pass

    def stop(self):
        self.stop_flag = True
        if self.proc is not None:
            try:
                self.proc.terminate()
            except Exception:
                pass

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except (ValueError, IndexError):
                continue
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        busy = sorted(sm)[len(sm) // 2:]                     # the upper half = samples taken under load
        return {"sm_mhz": float(np.median(busy)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm)}


def make_shard(rank, world, utts_per_gpu=None):
    from oracle import synth                                   # input generators only (no path computation)
    from xvector_b200 import scheduler
    lengths = synth.config2_lengths((utts_per_gpu or UTTS_PER_GPU) * world, seed=202)
    mine = scheduler.shard_lpt(lengths, world, DIM)[rank]
    my_len = lengths[mine].astype(np.int32)
    return my_len


def gemm_flops(lengths):
    """Algorithmic FLOPs of the GEMM layers per step (unpadded shapes, 2 FLOPs/MAC)."""
    from xvector_b200.scheduler import utterance_flops
    return float(utterance_flops(lengths, DIM).sum())


class Workload(object):
    """What one bench step runs.  tdnn = BASELINE.json configs[1] (the headline); resnet18 = configs[3]
    (secondary line, `--workload resnet18`): ResNet-18 + statistics pooling on fbank x 400 frames."""

    def __init__(self, args, rank, world):
        from oracle import synth                               # input/weight generators only (no path computation)
        from xvector_b200 import scheduler
        self.name = args.workload
        if self.name == "tdnn":
            self.dim = DIM
            self.cfg = dict(synth.TDNN_STAT_CONFIG)
            self.weights = synth.make_tdnn_weights(DIM, self.cfg, seed=1234)
            self.utts = args.utts or UTTS_PER_GPU
            self.lengths = make_shard(rank, world, self.utts)
            self.batch_frames = args.batch_frames or 262144
            self.scale = np.linspace(3.0, 0.5, DIM).astype(np.float32)
            t = self.lengths.astype(np.float64)
            self.layer_flops = {"gemm.tdnn1": 2 * 512 * 5 * DIM * (t - 4), "gemm.tdnn2": 2 * 512 * 2560 * (t - 8),
                                "gemm.tdnn3": 2 * 512 * 3584 * (t - 14), "gemm.tdnn4": 2 * 512 * 512 * (t - 14),
                                "gemm_pool.tdnn5": 2 * 512 * 1500 * (t - 14), "gemm.tdnn6": 2 * 3000 * 512 + 0 * t}
            self.layer_flops = {k: float(v.sum()) for k, v in self.layer_flops.items()}
            self.flops = float(scheduler.utterance_flops(self.lengths, DIM).sum())
            self.tensor_prefixes = ("gemm",)
            self.kernel = "xv_tdnn_gemm_kernel (tcgen05 spliced GEMM + fused epilogue), all 6 layer launches of a step"
            self.workload = Workload.describe(args)
        elif self.name == "attention":
            # configs[2]: TDNN + multi-head self-attention pooling (egs/voxceleb/v1/nnet_conf/tdnn_softmax_1e-2_tdnn4_att_pretrain.json
            # with att_num_heads = 4: keys from tdnn4_relu through a 1500-1500 network, values = tdnn5_relu, both split by head)
            self.dim = args.dim or 64
            self.cfg = dict(synth.TDNN_ATT_CONFIG, att_num_heads=4)
            self.weights = synth.make_tdnn_weights(self.dim, self.cfg, seed=1234)
            self.utts = args.utts or 2500
            self.lengths = make_shard(rank, world, self.utts)
            self.batch_frames = args.batch_frames or 262144
            self.scale = np.ones(self.dim, np.float32)
            t = self.lengths.astype(np.float64)
            key = 2.0 * (512 * 1500 + 1500 * 1500) * (t - 14)
            lf = {"gemm.tdnn1": 2 * 512 * 5 * self.dim * (t - 4), "gemm.tdnn2": 2 * 512 * 2560 * (t - 8),
                  "gemm.tdnn3": 2 * 512 * 3584 * (t - 14), "gemm.tdnn4": 2 * 512 * 512 * (t - 14),
                  "gemm_pool.tdnn5": 2 * 512 * 1500 * (t - 14), "gemm.att_key0": 2.0 * 512 * 1500 * (t - 14),
                  "gemm.att_key1": 2.0 * 1500 * 1500 * (t - 14), "gemm.tdnn6": 2 * 3000 * 512 + 0 * t}
            self.layer_flops = {k: float(v.sum()) for k, v in lf.items()}
            self.flops = float((scheduler.utterance_flops(self.lengths, self.dim) + key).sum())
            self.tensor_prefixes = ("gemm",)
            self.kernel = ("xv_tdnn_gemm_kernel: 5 frame layers + the 2 attention-key layers + tdnn6 (att_scores / att_softmax / "
                           "weighted pooling are separate HBM-bound kernels, listed in kernel_families_ms_per_step)")
            self.workload = Workload.describe(args)
        else:
            self.dim = args.dim or 64
            self.cfg = dict(synth.RESNET_CONFIG)
            self.weights = synth.make_resnet_weights(self.dim, self.cfg, seed=1234)
            self.utts = args.utts or 1024
            self.lengths = np.full(self.utts, 400, np.int32)      # fixed 400-frame segments: every rank gets the same shard
            self.batch_frames = args.batch_frames or 16384
            self.scale = np.ones(self.dim, np.float32)
            per = scheduler.resnet_layer_flops(400, self.dim)
            per["gemm_pool.dense2"] = per.pop("gemm.dense2")          # statistics pooling is fused into dense2's epilogue
            self.layer_flops = {k: v * self.utts for k, v in per.items()}
            self.flops = float(sum(v for k, v in self.layer_flops.items() if k != "stem"))
            self.tensor_prefixes = ("gemm", "conv")
            self.kernel = ("xv_conv_gemm_kernel (tcgen05 implicit-GEMM 3x3/1x1 convolutions, fused BN+residual+ReLU) + "
                           "xv_tdnn_gemm_kernel (conv5, dense1, dense2+pooling, tdnn6)")
            self.workload = Workload.describe(args)
        self.n_frames = int(self.lengths.sum())

    @staticmethod
    def describe(args):
        if getattr(args, "workload", "tdnn") == "resnet18":
            return ("configs[3]: ResNet-18 + statistics pooling (egs/voxceleb/v3) embedding extraction, "
                    "%d-dim fbank x 400 frames" % (args.dim or 64))
        if getattr(args, "workload", "tdnn") == "attention":
            return ("configs[2]: x-vector TDNN with 4-head self-attention pooling (keys: tdnn4_relu -> 1500 -> 1500, values: "
                    "tdnn5_relu) on %d-dim fbank utterances of log-uniform length 200-6000 frames" % (args.dim or 64))
        return ("configs[1]: standard x-vector TDNN (5x512 frame layers, 1500 pre-pool, stats pooling, "
                "tdnn6_dense 512-d) on 30-dim MFCC utterances of log-uniform length 200-6000 frames")


class Dist(object):
    """Rank bookkeeping + the torch.distributed plumbing the bench contract asks for (barrier, max / sum over ranks)."""

    def __init__(self):
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        self.dist = None
        if self.world > 1:
            import torch
            import torch.distributed as dist
            torch.cuda.set_device(self.local)
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            self.dist = dist
            self.host_group = dist.new_group(backend="gloo")

    def barrier_host(self):
        """Ranks meet on the HOST (gloo): a rank that waits here keeps its GPU idle.  An NCCL barrier is a kernel that spins on
        the GPU until the last rank arrives -- while rank 0 drove all eight GPUs from one process (the ark_e2e leg) the seven
        waiting ranks' barrier kernels sat on the SMs its persistent one-CTA-per-SM GEMMs count on, and halved their rate."""
        if self.dist is not None:
            import torch
            torch.cuda.synchronize()
            self.dist.barrier(group=self.host_group)

    def barrier(self, eng=None):
        if eng is not None:
            eng.sync()
        if self.dist is not None:
            import torch
            torch.cuda.synchronize()
            self.dist.barrier()

    def _reduce(self, v, op):
        if self.dist is None:
            return v
        import torch
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        self.dist.all_reduce(t, op=op)
        return float(t.item())

    def max(self, v):
        return self._reduce(v, self.dist.ReduceOp.MAX) if self.dist is not None else v

    def sum(self, v):
        return self._reduce(v, self.dist.ReduceOp.SUM) if self.dist is not None else v

    def close(self):
        if self.dist is not None:
            self.dist.destroy_process_group()
            self.dist = None


class Solo(object):
    """Dist stand-in for the single-GPU legs (no collective, rank 0 of 1)."""
    rank, world, dist = 0, 1, None

    def __init__(self, local=0):
        self.local = local

    def barrier(self, eng=None):
        if eng is not None:
            eng.sync()

    def max(self, v):
        return v

    def sum(self, v):
        return v


def measure_extraction(args, dc, with_pool=True, with_cpu=True):
    """One extraction workload at dc.world GPUs -> the JSON line (dict) of the bench contract."""
    import xvector_b200
    rank, world, local = dc.rank, dc.world, dc.local
    wl = Workload(args, rank, world)
    lengths, n_frames, cfg, weights, DIM = wl.lengths, wl.n_frames, wl.cfg, wl.weights, wl.dim
    eng = xvector_b200.Engine(cfg, DIM, precision="bf16", device=local, batch_frames=wl.batch_frames)
    eng.set_weights(weights)
    pin = xvector_b200.PinnedArray((n_frames, DIM))
    rng = np.random.default_rng(1000 + rank)
    scale = wl.scale
    blk = 1 << 20
    for s in range(0, n_frames, blk):                          # fill in blocks: avoids a second 0.8 GB temporary
        e = min(n_frames, s + blk)
        pin.array[s:e] = rng.standard_normal((e - s, DIM), dtype=np.float32) * scale
    feats = pin.array

    # ---------------- device-resident timing (headline `value`) ----------------
    eng.stage(feats, lengths)
    for _ in range(args.warmup):
        eng.run()
    dc.barrier(eng)
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = eng.launch_count()
    eng.event_record(0)
    for _ in range(args.steps):
        eng.run()
    eng.event_record(1)
    eng.sync()
    ms_dev = eng.elapsed_ms(0, 1)
    launches = eng.launch_count() - launches0
    dc.barrier(eng)
    sampler.stop()
    emb = eng.fetch()
    assert np.isfinite(emb).all()
    ms_dev = dc.max(ms_dev)
    total_frames = dc.sum(float(n_frames))
    total_launches = int(dc.sum(float(launches)))
    value = total_frames * args.steps / (ms_dev * 1e-3)

    # ---------------- per-kernel-family device time (roofline) ----------------
    eng.profile_enable(True)
    prof_steps = max(1, min(3, args.steps))
    for _ in range(prof_steps):
        eng.run()
    eng.sync()
    fam = {}
    for name in sorted(eng.launch_summary()):
        ms, cnt = eng.profile_query(name)
        if cnt:
            fam[name] = {"ms_per_step": ms / prof_steps, "launches_per_step": cnt / prof_steps}
    eng.profile_enable(False)
    is_tensor = lambda k: k.startswith(wl.tensor_prefixes)
    gemm_ms = sum(v["ms_per_step"] for k, v in fam.items() if is_tensor(k))
    gemm_launches = sum(v["launches_per_step"] for k, v in fam.items() if is_tensor(k))
    # per-family tensor-pipe throughput (algorithmic FLOPs of that family / its summed launch time)
    for k, v in fam.items():
        if k in wl.layer_flops and is_tensor(k) and v["ms_per_step"] > 0:
            v["tflops"] = wl.layer_flops[k] / (v["ms_per_step"] * 1e-3) / 1e12
    flops = wl.flops
    peaks = read_peaks()
    step_ms = ms_dev / args.steps
    # achieved = algorithmic FLOPs of the step / the step's time (every kernel of the step inside it)
    achieved = flops / (step_ms * 1e-3) / 1e12
    in_gemms = flops / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else 0.0
    roofline = {
        "kernel": wl.kernel,
        "bound": "tensor", "achieved": achieved, "peak": peaks["bf16_sustained"], "unit": "TFLOP/s",
        "frac": achieved / peaks["bf16_sustained"], "peak_kind": "sustained bf16, %s" % peaks["source"],
        "frac_of_burst": achieved / peaks["bf16_burst"], "basis": "algorithmic FLOPs of the step / ms_per_step (rank 0 share at N > 1)",
        "algorithmic_flops_per_step": flops, "launches_per_step": gemm_launches,
        "avg_launch_ms": gemm_ms / gemm_launches if gemm_launches else None,
        "achieved_inside_gemm_launches": in_gemms, "frac_inside_gemm_launches": in_gemms / peaks["bf16_sustained"],
        "gemm_share_of_step": gemm_ms / step_ms,
        "traffic": None,
    }
    if wl.name == "tdnn":
        # DRAM bytes per launch from the committed ncu --set full capture of the same kernels at the same batch size
        # (profiles/rNN_gemm_full_summary.csv: one launch of every layer; never measured under this timed run)
        tr = ncu_traffic("gemm_full_summary.csv", "xv_tdnn_gemm_kernel")
        if tr:
            roofline["traffic"], roofline["traffic_source"] = tr

    # ---------------- standalone statistics pooling (HBM roofline; rank 0 at N=1) ----------------
    pooling = None
    if rank == 0 and world == 1 and with_pool and wl.name == "tdnn":
        pooling = pooling_roofline(eng, peaks)

    # ---------------- end to end through the public C-ABI call (host buffers) ----------------
    e2e_steps = max(1, min(args.steps, 5))
    out = None
    for _ in range(2):
        out = eng.forward(feats, lengths)
    dc.barrier(eng)
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        out = eng.forward(feats, lengths)                        # H2D (pinned) + kernels + D2H + sync inside
    t1 = time.perf_counter()
    dc.barrier(eng)
    e2e_s = dc.max(t1 - t0)
    e2e_value = total_frames * e2e_steps / e2e_s
    assert np.isfinite(out).all()

    # ---------------- CPU baseline (rank 0, N=1 only) ----------------
    cpu = None
    if rank == 0 and world == 1 and with_cpu:
        cpu = cpu_baseline(args.cpu_seconds, args)

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": {
            "workload": wl.workload,
            "utts_per_gpu": wl.utts, "frames_rank0": n_frames, "feature_dim": DIM, "node": "tdnn6_dense",
            "batch_frames": wl.batch_frames, "weights": "random init, seed 1234",
            "parallelism": "utterance shards (LPT by FLOPs) over %d GPU(s), no collective" % world,
            "l2_policy": "inputs larger than L2 (%.2f GB fp32 features per GPU vs 126 MB L2)" % (n_frames * DIM * 4 / 1e9),
        },
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(n_frames * DIM * 4 + lengths.nbytes),
                "d2h_bytes_per_step": int(len(lengths) * 512 * 4), "steps": e2e_steps,
                "api": "xv_forward (pinned host features -> host embeddings)"},
        "gpu_launches": total_launches,
        "clocks": sampler.summary(),
        "roofline": roofline,
        "kernel_families_ms_per_step": fam,
    }
    if pooling is not None:
        line["pooling_roofline"] = pooling
    if cpu is not None:
        line["cpu_baseline"] = cpu
    pin.close()
    eng.close()
    return line


def compact(line, keep_families=True):
    """A leg's line cut down to what a reader of the combined JSON needs."""
    keys = ["metric", "value", "unit", "ms_per_step", "steps", "warmup", "scaling", "dtype", "config", "e2e", "gpu_launches",
            "clocks", "roofline", "loss_first_last"]
    out = {k: line[k] for k in keys if k in line}
    if keep_families and "kernel_families_ms_per_step" in line:
        out["kernel_families_ms_per_step"] = {k: {kk: round(vv, 4) for kk, vv in v.items()}
                                              for k, v in line["kernel_families_ms_per_step"].items()}
    return out


def leg_args(args, **kw):
    a = argparse.Namespace(**vars(args))
    for k, v in kw.items():
        setattr(a, k, v)
    return a


def run_leg(name, fn, *a, **k):
    t0 = time.time()
    try:
        out = fn(*a, **k)
    except Exception as e:                                     # a leg never takes the headline down
        import traceback
        sys.stderr.write("leg %s failed:\n%s\n" % (name, traceback.format_exc()))
        out = {"error": "%s: %s" % (type(e).__name__, str(e)[:300])}
    if isinstance(out, dict):
        out["leg_wall_s"] = round(time.time() - t0, 2)
    return out


class _P(object):
    """Params stand-in (misc/utils.py:13-41 interface: attributes + .dict)."""

    def __init__(self, **kw):
        self.__dict__.update(kw)

    @property
    def dict(self):
        return self.__dict__


def leg_cfg0(args, local):
    """BASELINE.json configs[0]: 1000 synthetic 30-dim MFCC utterances x 300 frames.  (a) the whole job in ONE packed call
    (device-resident and through xv_forward from pinned host memory); (b) the reference's actual call pattern: one
    Trainer.predict per utterance (model/trainer.py:886-913, extract.py:89), p50 / p99 wall latency."""
    import xvector_b200
    from oracle import synth
    from xvector_b200.model.trainer import Trainer
    from xvector_b200.scheduler import utterance_flops
    n, t, dim = 1000, 300, DIM
    cfg = dict(synth.TDNN_STAT_CONFIG)
    w = synth.make_tdnn_weights(dim, cfg, seed=1234)
    x = synth.config1_features(n, t, dim, seed=101)
    lengths = np.full(n, t, np.int32)
    flops = float(utterance_flops(lengths, dim).sum())
    eng = xvector_b200.Engine(cfg, dim, precision="bf16", device=local)
    eng.set_weights(w)
    pin = xvector_b200.PinnedArray((n * t, dim))
    pin.array[...] = x.reshape(n * t, dim)
    out = {"workload": "configs[0]: standard x-vector TDNN, %d utterances x %d frames x %d dims, node tdnn6_dense" % (n, t, dim),
           "frames": n * t, "algorithmic_flops": flops}
    try:
        eng.stage(pin.array, lengths)
        for _ in range(5):
            eng.run()
        eng.sync()
        steps = 50
        eng.event_record(0)
        for _ in range(steps):
            eng.run()
        eng.event_record(1)
        eng.sync()
        ms = eng.elapsed_ms(0, 1) / steps
        peaks = read_peaks()
        out["resident"] = {"ms_per_job": ms, "frames_per_s": n * t / (ms * 1e-3), "tflops": flops / (ms * 1e-3) / 1e12,
                           "frac_of_sustained_bf16": flops / (ms * 1e-3) / 1e12 / peaks["bf16_sustained"], "steps": steps}
        for _ in range(3):
            emb = eng.forward(pin.array, lengths)
        t0 = time.perf_counter()
        for _ in range(20):
            emb = eng.forward(pin.array, lengths)
        wall = (time.perf_counter() - t0) / 20
        assert np.isfinite(emb).all()
        out["e2e"] = {"ms_per_job": wall * 1e3, "frames_per_s": n * t / wall, "h2d_bytes": int(n * t * dim * 4), "d2h_bytes": int(n * 512 * 4),
                      "api": "xv_forward (pinned host features -> host embeddings), 20 calls"}
    finally:
        pin.close()
        eng.close()
    # (b) one predict per utterance, the way extract.py drives the reference
    trainer = Trainer(_P(**cfg), "/nonexistent", dim, single_cpu=True, precision="bf16", device=local)
    trainer.build("predict")
    trainer.set_weights(w)
    try:
        for i in range(20):
            trainer.predict(x[i])
        lat = []
        for i in range(300):
            t0 = time.perf_counter()
            e = trainer.predict(x[i % n])
            lat.append(time.perf_counter() - t0)
        assert e.shape == (512,)
        lat = np.sort(np.array(lat)) * 1e3
        out["predict_one_utterance"] = {"api": "Trainer.predict([300, 30]) (model/trainer.py:886-913), 300 calls", "p50_ms": float(lat[len(lat) // 2]),
                                        "p99_ms": float(lat[int(len(lat) * 0.99)]), "mean_ms": float(lat.mean()),
                                        "frames_per_s_serial": t / (float(lat.mean()) * 1e-3)}
    finally:
        trainer.close()
    if not args.no_cpu_baseline:
        # the CPU arm's per-utterance time on one core (the reference runs one single-thread process per core)
        from oracle import torch_ref
        import torch
        torch.set_num_threads(1)
        net = torch_ref.TorchTdnn(w, cfg)
        net.embed(x[:1])
        t0 = time.perf_counter()
        reps = 8
        for i in range(reps):
            net.embed(x[i:i + 1])
        out["cpu_one_utterance_ms"] = {"value": (time.perf_counter() - t0) / reps * 1e3, "kind": "port", "cores": 1,
                                       "sample": "%d x one 300-frame utterance, torch fp32 CPU restatement, 1 thread" % reps}
    return out


def leg_cfg1_chunked(args, local):
    """BASELINE.json configs[1] "chunked as in the extractor": utterances longer than chunk_size = 2000 are cut into
    overlapping chunks (hop 1000), embedded, and averaged weighted by chunk length (extract.py:68-86), through the
    drop-in extractor: extract_packed hands the chunks to the library as overlapping VIEWS of the packed (pinned) feature
    rows (xv_forward_views), so the features cross the bus once and nothing is sliced or copied on the host."""
    import xvector_b200
    from oracle import synth
    from xvector_b200.extract import extract_packed, plan_many
    from xvector_b200.model.trainer import Trainer
    dim, chunk, n_utts = DIM, 2000, 2048
    cfg = dict(synth.TDNN_STAT_CONFIG)
    w = synth.make_tdnn_weights(dim, cfg, seed=1234)
    lengths = synth.config2_lengths(UTTS_PER_GPU, seed=202)[:n_utts]
    n_in = int(lengths.sum())
    pin = xvector_b200.PinnedArray((n_in, dim))
    rng = np.random.default_rng(7)
    scale = np.linspace(3.0, 0.5, dim).astype(np.float32)
    for s0 in range(0, n_in, 1 << 20):
        e0 = min(n_in, s0 + (1 << 20))
        pin.array[s0:e0] = rng.standard_normal((e0 - s0, dim), dtype=np.float32) * scale
    work = plan_many(lengths, chunk, 25)
    processed = int(sum(z for _, _, z in work))
    trainer = Trainer(_P(**cfg), "/nonexistent", dim, single_cpu=True, precision="bf16", device=local)
    trainer.build("predict")
    trainer.set_weights(w)
    try:
        emb, kept = extract_packed(trainer, pin.array, lengths, chunk_size=chunk)
        reps = 5
        t0 = time.perf_counter()
        for _ in range(reps):
            emb, kept = extract_packed(trainer, pin.array, lengths, chunk_size=chunk)
        wall = (time.perf_counter() - t0) / reps
        assert len(kept) == len(lengths) and np.isfinite(emb).all()
        t0 = time.perf_counter()
        for _ in range(reps):
            extract_packed(trainer, pin.array, lengths, chunk_size=10000)       # the same list unchunked: the cost of the overlap
        wall_plain = (time.perf_counter() - t0) / reps
        # device-resident: the views staged once, kernels only
        offs = np.concatenate([[0], np.cumsum(lengths, dtype=np.int64)])
        row0 = np.array([offs[u] + s0 for u, s0, _ in work], np.int64)
        sizes = np.array([z for _, _, z in work], np.int32)
        eng = trainer.engine
        eng.stage_views(pin.array, row0, sizes)
        for _ in range(3):
            eng.run()
        eng.sync()
        eng.event_record(0)
        for _ in range(reps):
            eng.run()
        eng.event_record(1)
        eng.sync()
        ms_res = eng.elapsed_ms(0, 1) / reps
    finally:
        trainer.close()
        pin.close()
    return {"workload": "configs[1] lengths (first %d utterances), chunk_size %d, hop %d (extract.py:68-86)" % (n_utts, chunk, chunk // 2),
            "utterances": int(len(lengths)), "chunked_utterances": int((lengths > chunk).sum()), "pieces": len(work),
            "input_frames": n_in, "processed_frames": processed,
            "e2e": {"ms_per_job": wall * 1e3, "input_frames_per_s": n_in / wall, "processed_frames_per_s": processed / wall,
                    "unchunked_ms_per_job": wall_plain * 1e3, "h2d_bytes": int(n_in * dim * 4), "d2h_bytes": int(len(work) * 512 * 4),
                    "api": "extract_packed -> xv_forward_views (pinned packed features -> host embeddings; chunk plan + "
                           "length-weighted average on the host inside the timed region)"},
            "resident": {"ms_per_job": ms_res, "input_frames_per_s": n_in / (ms_res * 1e-3),
                         "processed_frames_per_s": processed / (ms_res * 1e-3)}}


def leg_ark_e2e(args, devices):
    """The real extraction job, ark file in -> x-vector ark out (egs/voxceleb/v1/nnet/lib/extract.py:63-93 as the
    product runs it): native reader (mmap + decode threads) into a ring of pinned buffers, one host thread + context per
    GPU (xvector_b200.extract.extract_stream_mgpu), ordered native writer.  The arks live on tmpfs so that the reader is
    measured, not a disk.  FM = float matrices, CM = Kaldi compressed matrices (what the recipes store)."""
    import shutil
    import tempfile
    from oracle import kaldi_io_ref, synth                     # generators / format writers for the INPUT files only
    from xvector_b200 import kaldi_io
    from xvector_b200.extract import CLI_BATCH_ROWS, extract_stream_mgpu      # the CLI's own batch cut
    from xvector_b200.model.trainer import Trainer
    dim = DIM
    cfg = dict(synth.TDNN_STAT_CONFIG)
    w = synth.make_tdnn_weights(dim, cfg, seed=1234)
    base = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else None
    tmp = tempfile.mkdtemp(prefix="xv_ark_e2e_", dir=base)
    out = {"devices": len(devices), "tmpfs": base is not None}
    try:
        rng = np.random.default_rng(11)
        scale = np.linspace(3.0, 0.5, dim).astype(np.float32)
        lengths = synth.config2_lengths(UTTS_PER_GPU, seed=202)
        n_fm = len(lengths)                                  # the 4096 utterances of configs[1]: 7.0 M frames, 0.84 GB
        fm = os.path.join(tmp, "feats_fm.ark")
        frames_fm = 0
        with kaldi_io.open_or_fd(fm, "wb") as fd:
            for i in range(n_fm):
                m = rng.standard_normal((int(lengths[i]), dim), dtype=np.float32) * scale
                kaldi_io.write_mat(fd, m, "utt%05d" % i)
                frames_fm += m.shape[0]
        n_cm = 1024
        cm = os.path.join(tmp, "feats_cm.ark")
        frames_cm = 0
        with open(cm, "wb") as f:
            for i in range(n_cm):
                m = rng.standard_normal((int(lengths[i]), dim), dtype=np.float32) * scale
                f.write(kaldi_io_ref.write_cm_bytes(m, "utt%05d" % i))
                frames_cm += m.shape[0]

        def make_trainer(device):
            t = Trainer(_P(**cfg), "/nonexistent", dim, single_cpu=True, precision="bf16", device=device)
            t.build("predict")
            t.set_weights(w)
            return t

        trainers = [make_trainer(d) for d in devices]          # contexts + weights once, outside the timed region
        try:
            from xvector_b200.extract import default_reader_threads
            out["reader_threads"] = default_reader_threads()
            for name, path, frames, reps in (("fm", fm, frames_fm, 4), ("cm", cm, frames_cm, 4)):
                best = None
                for _ in range(reps):                          # the first pass also warms buffers and the page cache
                    t0 = time.perf_counter()
                    written, skipped = extract_stream_mgpu(None, devices, path, os.path.join(tmp, "xvector.ark"), dim,
                                                           trainers=trainers, batch_rows=int(os.environ.get("XV_ARK_ROWS", CLI_BATCH_ROWS)))
                    dt = time.perf_counter() - t0
                    best = dt if best is None else min(best, dt)
                # the reader alone (decode into one buffer, nothing else)
                rd = None
                for _ in range(2):
                    t0 = time.perf_counter()
                    rows = 0
                    for _, feats, _ in kaldi_io.BatchReader(path, dim, max_rows=1 << 20, max_utts=4096, threads=default_reader_threads(),
                                                            batch_rows=CLI_BATCH_ROWS):
                        rows += feats.shape[0]
                    dt = time.perf_counter() - t0
                    rd = dt if rd is None else min(rd, dt)
                out[name] = {"utterances": int(written), "frames": int(frames), "bytes": os.path.getsize(path),
                             "seconds_best_of_%d" % reps: best, "frames_per_s": frames / best,
                             "reader_alone_frames_per_s": rows / rd,
                             "note": "wall clock of ark -> reader -> pinned ring -> xv_forward on %d GPU(s) -> ordered writer -> ark; "
                                     "contexts and weights are set up once outside" % len(devices)}
        finally:
            for t in trainers:
                t.close()
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    return out


def run_ours(args):
    dc = Dist()
    line = measure_extraction(args, dc, with_pool=not args.no_pool_bench, with_cpu=not args.no_cpu_baseline)
    legs = {}
    if args.workload == "tdnn" and not args.no_legs and not args.utts:
        if dc.world == 1:
            legs["cfg0"] = run_leg("cfg0", leg_cfg0, args, dc.local)
            legs["cfg1_chunked"] = run_leg("cfg1_chunked", leg_cfg1_chunked, args, dc.local)
            solo = Solo(dc.local)
            for name in ("attention", "resnet18"):
                la = leg_args(args, workload=name, steps=3, warmup=3, dim=0, batch_frames=0, utts=0)
                legs[name] = run_leg(name, lambda la=la: compact(measure_extraction(la, solo, with_pool=False, with_cpu=False)))
            la = leg_args(args, steps=200, warmup=20, train_global_batch=0)
            legs["train"] = run_leg("train", lambda: compact(measure_train(la, Solo(dc.local))))
            legs["ark_e2e"] = run_leg("ark_e2e", leg_ark_e2e, args, [dc.local])
        else:
            legs["train_dp"] = run_leg("train_dp", leg_train_dp, args, dc)
            dc.barrier_host()
            if dc.rank == 0:                                    # one process drives every GPU of the box: the product's shape
                legs["ark_e2e"] = run_leg("ark_e2e", leg_ark_e2e, args, list(range(dc.world)))
            dc.barrier_host()
    if dc.rank == 0:
        if legs:
            line["legs"] = legs
        emit(line)
    dc.close()


def ncu_traffic(suffix, kernel_substr):
    """(mean dram__bytes_read.sum + dram__bytes_write.sum per launch, source file) from the newest committed ncu summary
    under profiles/ whose name ends with `suffix`, over the rows whose kernel name contains `kernel_substr`."""
    import csv
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_" + suffix)))
    for path in reversed(files):
        try:
            rows = [r for r in csv.reader(l for l in open(path) if not l.startswith("#"))]
            hdr, units = rows[0], rows[1]
            ir, iw = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
            scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
            vals = [float(r[ir]) * scale[units[ir]] + float(r[iw]) * scale[units[iw]] for r in rows[2:] if kernel_substr in r[0]]
            if vals:
                return sum(vals) / len(vals), "%s (%d launches)" % (os.path.relpath(path, ROOT), len(vals))
        except (OSError, ValueError, KeyError, IndexError):
            continue
    return None


def pooling_roofline(eng, peaks, channels=1500, iters=10):
    """xv_stat_pooling_dev (the drop-in for direct statistics_pooling() callers, model/pooling.py:27-52) on a bf16
    [rows, 1500] matrix of 1.26 GB (10x the L2) cut into utterances with the configs[1] length distribution.
    Algorithmic bytes = rows*C*2 read + n*2C*4 written; time = events around every pooling call
    (memset of the accumulators + accumulate kernel + finalize kernel)."""
    import ctypes as C
    from oracle import synth
    from xvector_b200 import _lib
    lengths = synth.config2_lengths(4096, seed=202).astype(np.int64)
    target = 420000
    k = int(np.searchsorted(np.cumsum(lengths), target))
    lengths = lengths[:max(k, 1)]
    cu = np.concatenate([[0], np.cumsum(lengths)]).astype(np.int32)
    rows, n = int(cu[-1]), len(lengths)
    x_dev = eng.device_alloc(rows * channels * 2)
    cu_dev = eng.device_alloc(cu.nbytes)
    out_dev = eng.device_alloc(n * 2 * channels * 4)
    try:
        blk_rows = 16384
        blk = (np.random.default_rng(5).standard_normal((blk_rows, channels), dtype=np.float32)).clip(min=0)
        for r0 in range(0, rows, blk_rows):                   # values do not matter for bandwidth; reuse one block
            r1 = min(rows, r0 + blk_rows)
            eng.upload_bf16(C.c_void_p(x_dev.value + r0 * channels * 2), blk[:r1 - r0])
        eng.h2d(cu_dev, cu)
        call = lambda: eng._check(eng.lib.xv_stat_pooling_dev(eng._ctx, x_dev, C.cast(cu_dev, C.POINTER(C.c_int32)), n,
                                                               channels, _lib.DTYPE_BF16, int(lengths.max()),
                                                               C.cast(out_dev, C.POINTER(C.c_float))))
        for _ in range(3):
            call()
        eng.sync()
        eng.profile_enable(True)
        for _ in range(iters):
            call()
        eng.sync()
        ms, cnt = eng.profile_query("pool")
        eng.profile_enable(False)
        out = np.empty((n, 2 * channels), np.float32)
        eng.d2h(out, out_dev)
        assert np.isfinite(out).all()
    finally:
        eng.device_free(x_dev)
        eng.device_free(cu_dev)
        eng.device_free(out_dev)
    nbytes = rows * channels * 2 + n * 2 * channels * 4
    gbs = nbytes / (ms / cnt * 1e-3) / 1e9
    return {"kernel": "pool_accum_bulk_kernel<bf16,4> (cp.async.bulk-staged) + pool_finalize_kernel (xv_stat_pooling_dev)", "bound": "hbm",
            "achieved": gbs, "peak": peaks["hbm"], "unit": "GB/s", "frac": gbs / peaks["hbm"],
            "peak_kind": "copy bandwidth, %s" % peaks["source"], "algorithmic_bytes_per_launch": nbytes,
            "rows": rows, "utterances": n, "channels": channels, "avg_launch_ms": ms / cnt, "launches": cnt,
            "traffic": (ncu_traffic("pool_full_summary.csv", "pool_accum_bulk_kernel") or (None, None))[0]}


def measure_train(args, dc, precision=None, comm_bench=False):
    """BASELINE.json configs[4] -- one TDNN + AM-softmax training step (fwd + bwd + update) per step, data parallel with
    the library's own NCCL communicator (xv_comm_init; bucketed ncclAllReduce overlapped with the backward pass).
    Weak scaling: 64 utterances x 200 frames per GPU; --train-global-batch 64 = the reference-faithful split of a global
    batch over the towers (model/trainer.py:1133-1134).  Returns the JSON line (dict) on every rank."""
    import xvector_b200
    from oracle import synth
    from xvector_b200 import train_dp
    rank, world, local = dc.rank, dc.world, dc.local
    speakers, frames, dim = 5994, 200, DIM
    gbatch = args.train_global_batch or 64 * world
    precision = precision or args.train_precision
    cfg = dict(synth.TDNN_AMSOFTMAX_CONFIG)
    w = synth.add_softmax_weights(synth.make_tdnn_weights(dim, cfg, seed=1234), speakers)
    x, y = synth.config5_batch(gbatch, frames, dim, speakers, seed=505)
    mixed = precision == "bf16"
    eng = xvector_b200.Engine(cfg, dim, precision=precision, device=local)
    eng.set_weights(w)
    eng.train_init(speakers)
    uid = train_dp.exchange_unique_id_torch(eng, rank, world) if world > 1 else b"\0" * 128
    eng.comm_init(uid, rank, world)
    xs, ys = train_dp.slice_batch(x, y, rank, world)
    pin = xvector_b200.PinnedArray(xs.shape)
    pin.array[...] = xs
    losses = []
    for i in range(args.warmup):
        losses.append(eng.train_step(pin.array, ys, 0.01, i)[0])
    step_no = args.warmup
    # ---- `value`: the batch is resident in HBM before the timed region; steps are enqueued back to back (loss_out = NULL)
    d_x = eng.device_alloc(xs.nbytes)
    d_y = eng.device_alloc(ys.nbytes)
    eng.h2d(d_x, xs)
    eng.h2d(d_y, np.ascontiguousarray(ys, np.int32))
    for i in range(3):
        eng.train_step_resident(d_x, d_y, xs.shape[0], xs.shape[1], 0.01, step_no)
        step_no += 1
    eng.sync()
    dc.barrier(eng)
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = eng.launch_count()
    eng.event_record(0)
    for i in range(args.steps):
        eng.train_step_resident(d_x, d_y, xs.shape[0], xs.shape[1], 0.01, step_no)
        step_no += 1
    eng.event_record(1)
    eng.sync()
    ms_dev = dc.max(eng.elapsed_ms(0, 1))
    launches = eng.launch_count() - launches0
    dc.barrier(eng)
    # ---- `e2e`: the public pipelined call on a pinned host batch -- every step copies its features / labels to the device
    # (copy stream, overlapping the previous step) and brings its loss back; the host collects step k after submitting k+1
    eng.train_submit(pin.array, ys, 0.01, step_no)
    step_no += 1
    t0 = time.perf_counter()
    for i in range(args.steps):
        eng.train_submit(pin.array, ys, 0.01, step_no)
        step_no += 1
        losses.append(eng.train_collect()[0])
    t1 = time.perf_counter()
    losses.append(eng.train_collect()[0])
    wall = dc.max(t1 - t0)
    dc.barrier(eng)
    # ---- the synchronous call (xv_train_step with loss_out: H2D, step, D2H, host waits), for comparison
    n_sync = max(10, args.steps // 4)
    t0 = time.perf_counter()
    for i in range(n_sync):
        losses.append(eng.train_step(pin.array, ys, 0.01, step_no)[0])
        step_no += 1
    wall_sync = dc.max(time.perf_counter() - t0) / n_sync
    dc.barrier(eng)
    sampler.stop()
    eng.device_free(d_x)
    eng.device_free(d_y)
    assert np.isfinite(losses).all()
    info = eng.comm_info()
    comm = None
    if comm_bench and world > 1:
        comm = comm_roofline(eng, dc, info)
    total_frames = float(gbatch * frames)
    flops = train_dp.train_flops(gbatch, frames, dim, speakers)
    peaks = read_peaks()
    achieved = flops * args.steps / (ms_dev * 1e-3) / 1e12 / world           # per GPU: the roofline is a per-device quantity
    line = {
        "metric": "tdnn_amsoftmax_train_frames_per_sec", "value": total_frames * args.steps / (ms_dev * 1e-3), "unit": UNIT,
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_dev / args.steps,
        "higher_is_better": True, "scaling": "weak" if not args.train_global_batch else "strong", "vs_baseline": None,
        "dtype": "bf16 operands / f32 accumulate + f32 master weights" if mixed else "f32", "data": "synthetic",
        "config": {"workload": "configs[4]: TDNN x-vector + AM-softmax training step fwd+bwd+update, %d x %d frames x %d dims, "
                               "%d speakers, SGD" % (gbatch, frames, dim, speakers), "global_batch": gbatch,
                   "per_gpu_batch": gbatch // world, "weights": "random init, seed 1234",
                   "parallelism": "dp%d: ncclAllReduce(sum) of the flat gradient buffer (%d elements, travelling as %s) in %d bucket(s) "
                                  "on a side stream, overlapped with the backward pass; library communicator (xv_comm_init)" % (
                                      world, info["grad_elems"], info.get("dtype", "fp32"), info["buckets"]),
                   "comm_nranks": info["world"],
                   "l2_policy": "activations + gradients of a step (0.4 GB) exceed the L2"},
        "e2e": {"value": total_frames * args.steps / wall, "unit": UNIT, "h2d_bytes_per_step": int(xs.nbytes + ys.nbytes),
                "d2h_bytes_per_step": 20, "steps": args.steps,
                "api": "xv_train_submit / xv_train_collect (pinned host batch -> loss on the host, step k collected after k+1 is submitted)",
                "synchronous_xv_train_step_ms": wall_sync * 1e3},
        "gpu_launches": int(dc.sum(float(launches))), "clocks": sampler.summary(),
        "roofline": {"kernel": ("tcgen05 GEMMs of the frame layers (forward + tap-step -1 dgrad on xv_tdnn_gemm_kernel, MN-major "
                                "xv_wgrad_kernel) inside the whole step" if mixed else
                                "fp32 SIMT GEMMs of the validation-grade training path (forward, dgrad, wgrad)"),
                     "bound": "tensor", "achieved": achieved, "peak": peaks["bf16_sustained"], "unit": "TFLOP/s",
                     "frac": achieved / peaks["bf16_sustained"], "peak_kind": "sustained bf16, %s" % peaks["source"],
                     "algorithmic_flops_per_step": flops, "traffic": None,
                     "note": "whole-step FLOPs per GPU / whole-step time" + ("" if mixed else "; this path runs on FFMA, not on the tensor cores")},
        "loss_first_last": [losses[0], losses[-1]],
    }
    if comm is not None:
        line["allreduce"] = comm
    pin.close()
    eng.close()
    return line


def comm_roofline(eng, dc, info, iters=20):
    """The exchange step alone: the step's gradient buckets all-reduced back to back through the library's communicator
    (xv_comm_bench), CUDA events on the communication stream, max over ranks.  Bus bytes of a ring/tree allreduce =
    2 (N-1)/N x payload; NVLink 5 gives 900 GB/s per direction per GPU."""
    import ctypes as C
    ms = C.c_float()
    eng._check(eng.lib.xv_comm_bench(eng._ctx, 3, C.byref(ms)))       # warm-up
    dc.barrier(eng)
    eng._check(eng.lib.xv_comm_bench(eng._ctx, iters, C.byref(ms)))
    per = dc.max(float(ms.value))
    payload = info["grad_elems"] * (2 if info.get("dtype") == "bf16" else 4)       # (the fp32 tail of a bf16 exchange is < 0.1 %)
    n = info["world"]
    bus = 2.0 * (n - 1) / n * payload / (per * 1e-3) / 1e9
    return {"ms_per_allreduce": per, "payload_bytes": payload, "dtype": info.get("dtype", "fp32"), "buckets": info["buckets"], "nranks": n,
            "bus_gbs": bus, "nvlink_peak_gbs_per_direction": 900.0, "frac_of_nvlink": bus / 900.0, "iters": iters}


def train_dp_parity(dc):
    """The N-tower step (fp32 validation path, small shapes) against the committed golden written from the fp64 oracle
    (tests/golden/make_train_dp_golden.py): loss, three gradients and the tower-0 moving mean."""
    import xvector_b200
    from oracle import synth                                   # input / weight generators only
    from xvector_b200 import train_dp
    path = os.path.join(ROOT, "tests", "golden", "train_dp_golden.npz")
    g = np.load(path)
    n = dc.world
    if "n%d/loss" % n not in g.files:
        return {"skipped": "no golden for %d towers" % n}
    cfg = dict(synth.TDNN_AMSOFTMAX_CONFIG)
    dim, speakers, batch, frames = 24, 101, 16, 50
    w = synth.add_softmax_weights(synth.make_tdnn_weights(dim, cfg, seed=31), speakers)
    x, y = synth.config5_batch(batch, frames, dim, speakers, seed=32)
    eng = xvector_b200.Engine(cfg, dim, precision="fp32", device=dc.local)
    eng.set_weights(w)
    eng.train_init(speakers)
    eng.comm_init(train_dp.exchange_unique_id_torch(eng, dc.rank, n), dc.rank, n)
    xs, ys = train_dp.slice_batch(x, y, dc.rank, n)
    loss, _ = eng.train_step(xs, ys, 0.05, 0)
    out = {"towers": n, "loss_rel": abs(loss - float(g["n%d/loss" % n])) / abs(float(g["n%d/loss" % n]))}
    worst = 0.0
    for key in g.files:
        if not key.startswith("n%d/g:" % n):
            continue
        name = key.split(":", 1)[1]
        ref = g[key]                                           # every 97th element of the flattened gradient
        got = eng.train_grad(name, np.asarray(w[name]).shape).reshape(-1)[::97]
        worst = max(worst, float(np.linalg.norm(got - ref) / max(np.linalg.norm(ref), 1e-30)))
    mm = eng.train_weight("tdnn/tdnn1_bn/moving_mean", (512,))
    out["grad_l2_rel_max"] = worst
    out["moving_mean_abs_max"] = float(np.abs(mm - g["n%d/mm" % n]).max())
    eng.close()
    worst_all = dc.max(max(out["loss_rel"] / 1e-5, worst / 5e-3, out["moving_mean_abs_max"] / 1e-6))
    out["ok"] = bool(worst_all <= 1.0)
    out["tolerances"] = {"loss_rel": 1e-5, "grad_l2_rel": 5e-3, "moving_mean_abs": 1e-6}
    return out


def leg_train_dp(args, dc):
    """N > 1: the data-parallel training step through the library's communicator -- weak (64 utterances per GPU), strong
    (global 64 split over the towers, model/trainer.py:1133-1134), the allreduce alone, and the parity check."""
    steps = 200
    out = {}
    la = leg_args(args, steps=steps, warmup=20, train_global_batch=0)
    weak = measure_train(la, dc, comm_bench=True)
    out["weak"] = compact(weak)
    out["weak"]["allreduce"] = weak.get("allreduce")
    if 64 % dc.world == 0:
        ls = leg_args(args, steps=steps, warmup=20, train_global_batch=64)
        out["strong_global64"] = compact(measure_train(ls, dc))
    out["parity"] = train_dp_parity(dc)
    return out


def run_train(args):
    dc = Dist()
    line = measure_train(args, dc, comm_bench=True)
    if dc.rank == 0:
        emit(line)
    dc.close()


def cpu_baseline(seconds, args=None):
    """fp32 torch-CPU restatement (oracle/torch_ref.py), P single-thread processes, one utterance per call."""
    from oracle import synth, torch_ref
    import torch
    cores = os.cpu_count() or 1
    workload = getattr(args, "workload", "tdnn") if args is not None else "tdnn"
    if workload == "resnet18":
        dim = (args.dim or 64)
        sub = np.full(max(cores, 8), 400, np.int32)
        feats = synth.fbank_like(int(sub.sum()), dim, seed=404)
        cfg = dict(synth.RESNET_CONFIG)
        w = synth.make_resnet_weights(dim, cfg, seed=1234)
        what = "%d-dim fbank x 400-frame utterances of configs[3]" % dim
    elif workload == "attention":
        dim = (args.dim or 64)
        lengths = synth.config2_lengths(UTTS_PER_GPU, seed=202)
        take = np.random.default_rng(0).permutation(len(lengths))[:max(4 * cores, 64)]
        sub = lengths[take].astype(np.int32)
        feats = synth.fbank_like(int(sub.sum()), dim, seed=202)
        cfg = dict(synth.TDNN_ATT_CONFIG, att_num_heads=4)
        w = synth.make_tdnn_weights(dim, cfg, seed=1234)
        what = "%d-dim utterances drawn from the configs[2] length distribution (4-head attention pooling)" % dim
    else:
        lengths = synth.config2_lengths(UTTS_PER_GPU, seed=202)
        take = np.random.default_rng(0).permutation(len(lengths))[:max(4 * cores, 64)]
        sub = lengths[take].astype(np.int32)
        feats = synth.packed_features(sub, DIM, seed=202)
        cfg = dict(synth.TDNN_STAT_CONFIG)
        w = synth.make_tdnn_weights(DIM, cfg, seed=1234)
        what = "utterances drawn from the configs[1] length distribution"
    r = torch_ref.time_cpu_extraction(w, cfg, feats, sub, processes=cores, budget_s=seconds)
    return {"value": r["frames_per_s"], "unit": UNIT, "cores": r["cores"], "kind": "port",
            "sample": "%d %s (%d frames), %.0f s budget per process, "
                      "%d single-thread processes, one utterance per call; torch %s fp32 CPU restatement of the reference "
                      "graph (reference TF unavailable on this box)" % (r["utts"], what, r["frames"], seconds, r["cores"],
                                                                          torch.__version__)}


def run_reference(args):
    """--impl reference: the reference's CPU extraction path (its TF code cannot run here; the fp32 torch-CPU
    restatement stands in), all host cores, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    per_step = max(2.0, min(args.cpu_seconds, 60.0 / max(1, args.steps + args.warmup)))
    vals = []
    cpu = None
    for i in range(args.warmup + args.steps):
        cpu = cpu_baseline(per_step, args)
        if i >= args.warmup:
            vals.append(cpu["value"])
    v = float(np.mean(vals))
    cpu["value"] = v
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": per_step * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": Workload.describe(args), "node": "tdnn6_dense", "weights": "random init, seed 1234",
                       "sample": "bounded sample of that workload per step (see cpu_baseline.sample)"},
            "cpu_baseline": cpu,
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


_REAL_STDOUT = None


def claim_stdout():
    """stdout carries exactly ONE JSON line: keep the real stdout aside and point fd 1 at stderr, so that anything a
    library prints there (NCCL's version banner, a stray warning) cannot land beside it."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line):
    out = _REAL_STDOUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="tdnn", choices=["tdnn", "attention", "resnet18", "train"],
                    help="tdnn = BASELINE.json configs[1] (the headline metric); resnet18 = configs[3] (secondary)")
    ap.add_argument("--dim", type=int, default=0, help="feature dim of the resnet18 workload (default 64; 40 = reference-legal)")
    ap.add_argument("--batch-frames", type=int, default=0)
    ap.add_argument("--train-global-batch", type=int, default=0, help="train workload: fixed global batch (strong scaling); "
                    "default 64 per GPU (weak scaling)")
    ap.add_argument("--train-precision", default="bf16", choices=["bf16", "fp32"],
                    help="train workload: bf16 = tcgen05 GEMMs with bf16 operands and fp32 master weights; fp32 = validation path")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-pool-bench", action="store_true")
    ap.add_argument("--no-legs", action="store_true", help="default workload only: skip the extra legs (cfg0, cfg1_chunked, "
                    "attention, resnet18, train at N = 1; train_dp at N > 1)")
    ap.add_argument("--utts", type=int, default=0, help="utterances per GPU (default = the configs[1] set; "
                    "smaller values are for profiling runs only)")
    args = ap.parse_args()
    claim_stdout()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "train":
        run_train(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
