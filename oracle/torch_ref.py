"""fp32 torch-CPU restatement of the reference graph: second independent oracle and
the TIMED CPU BASELINE.

TEST INFRASTRUCTURE -- see oracle/__init__.py.  "CPU restatement of the reference
graph (reference TF unavailable on this box)".  Deployment shape mirrors the
reference's extractor: one utterance per call, one thread per process
(model/trainer.py:135-139, egs/voxceleb/v1/nnet/lib/extract.py:56,89), P processes.

Built from torch's own conv2d / linear / batch_norm kernels (not from
numpy_ref.py) so that the two restatements are independent implementations of
model/tdnn.py:36-149 and model/pooling.py:40-50,96-218.
"""
from __future__ import annotations

import os
import time

import numpy as np
import torch
import torch.nn.functional as F

BN_EPSILON = 1e-3
VAR2STD_EPSILON = 1e-12


class TorchTdnn(object):
    """tdnn() up to `tdnn6_dense` (the embedding node, model/tdnn.py:137-142), fp32."""

    def __init__(self, weights, config):
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32))
        self.cfg = dict(config)
        self.w = {k: t(v) for k, v in weights.items()}
        # TF conv kernel [1,k,Cin,Cout] -> torch [Cout,Cin,1,k]
        self.conv = [self.w["tdnn/tdnn%d_conv/kernel" % i].permute(3, 2, 0, 1).contiguous() for i in (1, 2, 3)]
        kind = self.cfg.get("network_relu_type", None)
        self.kind = kind

    def _act(self, x, scope, channel_axis):
        if self.kind == "prelu":
            a = self.w[scope + "/alpha"]
            shape = [1] * x.dim()
            shape[channel_axis] = -1
            return torch.relu(x) + a.view(shape) * (x - x.abs()) * 0.5
        if self.kind == "lrelu":
            return F.leaky_relu(x, 0.2)
        return torch.relu(x)

    def _bn(self, x, scope):
        # F.batch_norm normalises dim 1
        return F.batch_norm(x, self.w[scope + "/moving_mean"], self.w[scope + "/moving_variance"],
                            self.w[scope + "/gamma"], self.w[scope + "/beta"], False, 0.0, BN_EPSILON)

    def frame_layers(self, feats):
        """feats [B,T,D] -> (tdnn4_relu [B,512,T'], tdnn5_relu [B,C,T'])  (channels-first)."""
        x = feats.permute(0, 2, 1).unsqueeze(2)                  # [B,D,1,T]
        for i in (1, 2, 3):
            x = F.conv2d(x, self.conv[i - 1], self.w["tdnn/tdnn%d_conv/bias" % i])
            x = self._bn(x, "tdnn/tdnn%d_bn" % i)
            x = self._act(x, "tdnn/tdnn%d_relu" % i, 1)
        x = x.squeeze(2)                                         # [B,512,T']
        outs = []
        for i in (4, 5):
            x = F.linear(x.transpose(1, 2), self.w["tdnn/tdnn%d_dense/kernel" % i].t(),
                         self.w["tdnn/tdnn%d_dense/bias" % i]).transpose(1, 2)
            x = self._bn(x, "tdnn/tdnn%d_bn" % i)
            x = self._act(x, "tdnn/tdnn%d_relu" % i, 1)
            outs.append(x)
        return outs[0], outs[1]

    @staticmethod
    def stat_pool(x):
        """x [B,C,L] -> [B,2C]  (model/pooling.py:40-50)."""
        mean = x.mean(dim=2, keepdim=True)
        var = ((x - mean) ** 2).mean(dim=2)
        var = torch.where(var <= VAR2STD_EPSILON, torch.full_like(var, VAR2STD_EPSILON), var)
        return torch.cat([mean.squeeze(2), var.sqrt()], dim=1)

    def _side(self, x, prefix, nodes, last_type):
        """x [B,L,C] channels-last."""
        nodes = list(nodes)
        for j in range(len(nodes)):
            name = "%s%d" % (prefix, j)
            base = "tdnn/attention/%s" % name
            x = F.linear(x, self.w[base + "/%s_dense/kernel" % name].t(), self.w[base + "/%s_dense/bias" % name])
            last = j == len(nodes) - 1
            if (not last) or last_type == 2:
                x = self._bn(x.transpose(1, 2), base + "/%s_bn" % name).transpose(1, 2)
            if (not last) or last_type in (1, 2):
                x = self._act(x, base + "/%s_relu" % name, 2)
            elif last_type == 3:
                x = torch.tanh(x)
        return x

    def att_pool(self, key_in, value_in):
        """key_in / value_in channels-first [B,C,L] (model/pooling.py:96-218)."""
        c = self.cfg
        key = self._side(key_in.transpose(1, 2), "att_key", c["att_key_num_nodes"], c["att_key_network_type"])
        val = self._side(value_in.transpose(1, 2), "att_value", c["att_value_num_nodes"], c["att_value_network_type"])
        q = self.w["tdnn/attention/query"]
        h = q.shape[0]
        b, l, cv = val.shape
        if c["att_split_value"]:
            v = val.reshape(b, l, h, cv // h).permute(0, 2, 1, 3)
        else:
            v = val.unsqueeze(1).expand(b, h, l, cv)
        if c["att_split_key"]:
            k = key.reshape(b, l, h, key.shape[2] // h).permute(0, 2, 1, 3)
            s = torch.einsum("bhld,hd->bhl", k, q)
        else:
            s = torch.einsum("bld,hd->bhl", key, q)
        if c["att_use_scale"]:
            s = s * (1.0 / float(np.sqrt(q.shape[1])))
        wgt = torch.softmax(s, dim=2)
        mean = torch.einsum("bhld,bhl->bhd", v, wgt)
        var = torch.einsum("bhld,bhl->bhd", (v - mean.unsqueeze(2)) ** 2, wgt)
        mean = mean.reshape(b, -1)
        var = var.reshape(b, -1)
        var = torch.where(var <= VAR2STD_EPSILON, torch.full_like(var, VAR2STD_EPSILON), var)
        att = torch.cat([mean, var.sqrt()], dim=1)
        if c.get("att_apply_nonlinear", False):
            att = self._bn(att, "tdnn/attention/att_post_bn")
            att = self._act(att, "tdnn/attention/att_post_relu", 1)
        return att

    @torch.no_grad()
    def embed(self, feats):
        """feats np [T,D] or [B,T,D] -> tdnn6_dense np [512] / [B,512]  (trainer.py:886-913)."""
        x = torch.from_numpy(np.ascontiguousarray(feats, dtype=np.float32))
        squeeze = x.dim() == 2
        if squeeze:
            x = x.unsqueeze(0)
        t4, t5 = self.frame_layers(x)
        if self.cfg["pooling_type"] == "statistics_pooling":
            p = self.stat_pool(t5)
        elif self.cfg["pooling_type"] == "self_attention":
            src = {"tdnn4_relu": t4, "tdnn5_relu": t5}
            p = self.att_pool(src[self.cfg["att_key_input"]], src[self.cfg["att_value_input"]])
        else:
            raise NotImplementedError(self.cfg["pooling_type"])
        e = F.linear(p, self.w["tdnn/tdnn6_dense/kernel"].t(), self.w["tdnn/tdnn6_dense/bias"])
        e = e.numpy()
        return e[0] if squeeze else e


class TorchEtdnn(TorchTdnn):
    """etdnn() up to `tdnn12_dense` (model/tdnn.py:343-536), fp32, statistics pooling."""

    LAYERS = (("tdnn1", 5), ("tdnn2", 0), ("tdnn3", 5), ("tdnn4", 0), ("tdnn5", 7), ("tdnn6", 0), ("tdnn7", 9),
              ("tdnn8", 0), ("tdnn9", 0), ("tdnn10", 0))

    def __init__(self, weights, config):
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32))
        self.cfg = dict(config)
        self.w = {k: t(v) for k, v in weights.items()}
        self.kind = self.cfg.get("network_relu_type", None)

    @torch.no_grad()
    def embed(self, feats):
        x = torch.from_numpy(np.ascontiguousarray(feats, dtype=np.float32))
        squeeze = x.dim() == 2
        if squeeze:
            x = x.unsqueeze(0)
        x = x.transpose(1, 2)                                    # [B,C,T]
        for name, k in self.LAYERS:
            if k:
                # TF conv1d kernel [k,Cin,Cout] -> torch [Cout,Cin,k]
                x = F.conv1d(x, self.w["etdnn/%s_conv/kernel" % name].permute(2, 1, 0).contiguous(),
                             self.w["etdnn/%s_conv/bias" % name])
            else:
                x = F.linear(x.transpose(1, 2), self.w["etdnn/%s_dense/kernel" % name].t(),
                             self.w["etdnn/%s_dense/bias" % name]).transpose(1, 2)
            x = self._bn(x, "etdnn/%s_bn" % name)
            x = self._act(x, "etdnn/%s_relu" % name, 1)
        p = self.stat_pool(x)
        e = F.linear(p, self.w["etdnn/tdnn12_dense/kernel"].t(), self.w["etdnn/tdnn12_dense/bias"]).numpy()
        return e[0] if squeeze else e


class TorchResnet18(TorchTdnn):
    """resnet_18() up to `tdnn6_dense` (model/resnet.py:152-312), fp32, NCHW with H=time, W=freq.
    TF 'SAME' padding is applied explicitly (asymmetric for stride 2: SURVEY.md 7.3-i)."""

    def __init__(self, weights, config):
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32))
        self.cfg = dict(config)
        self.w = {k: t(v) for k, v in weights.items()}
        self.kind = self.cfg.get("network_relu_type", None)

    def _conv_same(self, x, name, strides=(1, 1)):
        k = self.w["resnet_18/%s/kernel" % name]                 # [kh,kw,Cin,Cout]
        kh, kw = k.shape[0], k.shape[1]
        pads = []
        for n_in, kk, s in ((x.shape[3], kw, strides[1]), (x.shape[2], kh, strides[0])):   # F.pad: last dim first
            n_out = -(-n_in // s)
            total = max((n_out - 1) * s + kk - n_in, 0)
            pads += [total // 2, total - total // 2]
        x = F.pad(x, pads)
        return F.conv2d(x, k.permute(3, 2, 0, 1).contiguous(), None, stride=strides)

    def _block(self, x, name, strides, project):
        s = "resnet_18/" + name
        y = self._conv_same(x, name + "_conv0", strides)
        y = self._act(self._bn(y, s + "_bn0"), s + "_relu0", 1)
        y = self._bn(self._conv_same(y, name + "_conv1"), s + "_bn1")
        sc = self._bn(self._conv_same(x, name + "_conv_short", strides), s + "_bn_short") if project else x
        return self._act(y + sc, s + "_relu_final", 1)

    @torch.no_grad()
    def embed(self, feats):
        x = torch.from_numpy(np.ascontiguousarray(feats, dtype=np.float32))
        squeeze = x.dim() == 2
        if squeeze:
            x = x.unsqueeze(0)
        x = x.unsqueeze(1)                                       # [B,1,T,F]
        x = self._act(self._bn(self._conv_same(x, "conv0_1"), "resnet_18/conv0_bn"), "resnet_18/conv0_relu", 1)
        if self.cfg.get("resnet_maxpooling", False):
            x = F.max_pool2d(x, 3, 1, 1)
        ts = 2 if self.cfg.get("resnet_time_stride", False) else 1
        blocks = self.cfg.get("resnet_blocks", [2, 2, 2, 2])
        for stage in range(4):
            x = self._block(x, "conv%da" % (stage + 1), (1, 1) if stage == 0 else (ts, 2), True)
            for i in range(blocks[stage] - 1):
                x = self._block(x, "conv%db_%d" % (stage + 1, i), (1, 1), False)
        k5 = self.w["resnet_18/conv5/kernel"]                    # [1,F',512,512]
        x = F.conv2d(x, k5.permute(3, 2, 0, 1).contiguous(), self.w["resnet_18/conv5/bias"])
        x = self._act(self._bn(x, "resnet_18/conv5_bn"), "resnet_18/conv5_relu", 1).squeeze(3)   # [B,512,T]
        for name in ("dense1", "dense2"):
            x = F.linear(x.transpose(1, 2), self.w["resnet_18/%s/kernel" % name].t(),
                         self.w["resnet_18/%s/bias" % name]).transpose(1, 2)
            x = self._act(self._bn(x, "resnet_18/%s_bn" % name), "resnet_18/%s_relu" % name, 1)
        p = self.stat_pool(x)
        e = F.linear(p, self.w["resnet_18/tdnn6_dense/kernel"].t(), self.w["resnet_18/tdnn6_dense/bias"]).numpy()
        return e[0] if squeeze else e


def make_net(weights, config):
    kind = config.get("network_type", "tdnn")
    return {"tdnn": TorchTdnn, "extended_tdnn": TorchEtdnn, "resnet_18": TorchResnet18}[kind](weights, config)


# ------------------------------------------------------------------- timed CPU baseline
def _worker(args):
    weights, config, feats, lengths, idx, budget_s = args
    torch.set_num_threads(1)
    try:
        torch.set_num_interop_threads(1)
    except RuntimeError:
        pass
    net = make_net(weights, config)
    offs = np.concatenate([[0], np.cumsum(lengths)])
    net.embed(feats[offs[idx[0]]:offs[idx[0] + 1]])            # warm-up, untimed
    frames = 0
    utts = 0
    t0 = time.perf_counter()
    for i in idx:
        net.embed(feats[offs[i]:offs[i + 1]])                  # one utterance per call (extract.py:89)
        frames += int(lengths[i])
        utts += 1
        if time.perf_counter() - t0 > budget_s:
            break
    return frames, utts, time.perf_counter() - t0


def time_cpu_extraction(weights, config, feats, lengths, processes=None, budget_s=15.0):
    """Reference-shaped CPU extraction: P single-thread processes, utterances dealt round-robin.

    Returns dict(frames_per_s, utts, frames, seconds, cores).  Each process stops after
    `budget_s` seconds; throughput = sum over processes of frames_i / seconds_i.
    """
    import multiprocessing as mp
    p = processes or (os.cpu_count() or 1)
    n = len(lengths)
    p = max(1, min(p, n))
    jobs = [(weights, config, feats, lengths, list(range(r, n, p)), budget_s) for r in range(p)]
    if p == 1:
        res = [_worker(jobs[0])]
    else:
        ctx = mp.get_context("fork")
        with ctx.Pool(p) as pool:
            res = pool.map(_worker, jobs)
    fps = sum(f / s for f, _, s in res)
    return {"frames_per_s": fps, "frames": sum(f for f, _, _ in res), "utts": sum(u for _, u, _ in res),
            "seconds": max(s for _, _, s in res), "cores": p}
