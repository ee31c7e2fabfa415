"""Engine: thin Python object over one xv_ctx (one GPU, one stream).  All compute goes through the C-ABI."""
from __future__ import annotations

import ctypes as C
import json

import numpy as np

from . import _lib
from ._lib import XVectorError, check


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _fptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _i32ptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


def _i64ptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_int64))


def config_json(params_dict, dim, precision="bf16", batch_frames=None):
    """nnet_conf dict (+ feature_dim / precision) -> JSON text for xv_create.  Non-JSON values (e.g. the TF
    global_step placeholder the reference stuffs into params.dict, model/trainer.py:154) are dropped."""
    out = {}
    for k, v in dict(params_dict).items():
        try:
            json.dumps(v)
        except (TypeError, ValueError):
            continue
        out[k] = v
    out["feature_dim"] = int(dim)
    out["precision"] = precision
    if batch_frames is not None:
        out["batch_frames"] = int(batch_frames)
    return json.dumps(out)


class PinnedArray(object):
    """numpy view over cudaMallocHost memory (true async H2D for xv_forward)."""

    def __init__(self, shape, dtype=np.float32):
        self._lib = _lib.load()
        self.nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        self._ptr = self._lib.xv_host_alloc(max(self.nbytes, 16))
        if not self._ptr:
            raise XVectorError(_lib.XV_ENOMEM, "xv_host_alloc(%d) failed (no CUDA device?)" % self.nbytes)
        buf = (C.c_char * max(self.nbytes, 1)).from_address(self._ptr)
        self.array = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    def close(self):
        if self._ptr:
            self.array = None
            self._lib.xv_host_free(self._ptr)
            self._ptr = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Engine(object):
    def __init__(self, params_dict, dim, precision="bf16", device=0, batch_frames=None):
        self.lib = _lib.load()
        self.dim = int(dim)
        self.precision = precision
        self._ctx = C.c_void_p()
        cfg = config_json(params_dict, dim, precision, batch_frames)
        rc = self.lib.xv_create(int(device), cfg.encode("utf-8"), C.byref(self._ctx))
        if rc != _lib.XV_OK:
            self._ctx = C.c_void_p()
            msg = self.lib.xv_last_error(None)
            raise XVectorError(rc, msg.decode("utf-8", "replace") if msg else "")
        self.embedding_node = dict(params_dict).get("embedding_node", "tdnn6_dense")

    # -- life cycle
    def close(self):
        if getattr(self, "_ctx", None) is not None and self._ctx:
            self.lib.xv_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        return check(rc, self._ctx)

    # -- weights
    def set_weights(self, weights):
        """weights: {tf variable name: ndarray}.  Copies everything, then folds/casts on the device."""
        self._weight_shapes = {}
        for name, arr in weights.items():
            a = _f32(arr)
            self._weight_shapes[name] = tuple(a.shape)
            if a.ndim == 0:
                a = a.reshape(1)
            shape = np.asarray(a.shape, dtype=np.int64)
            self._check(self.lib.xv_set_weight(self._ctx, name.encode("utf-8"), _fptr(a), _i64ptr(shape), a.ndim))
        self._check(self.lib.xv_finalize_weights(self._ctx))

    # -- nodes
    def node_info(self, node=""):
        w, fl, cx = C.c_int32(), C.c_int32(), C.c_int32()
        self._check(self.lib.xv_node_info(self._ctx, (node or "").encode("utf-8"), C.byref(w), C.byref(fl), C.byref(cx)))
        return int(w.value), bool(fl.value), int(cx.value)

    # -- forward
    @staticmethod
    def _pack(feats, lengths):
        lengths = np.ascontiguousarray(lengths, dtype=np.int32)
        if isinstance(feats, np.ndarray) and feats.dtype == np.float32 and feats.flags["C_CONTIGUOUS"]:
            f = feats
        else:
            f = _f32(feats)
        if f.ndim != 2 or f.shape[0] != int(lengths.sum()):
            raise ValueError("feats must be packed [sum(lengths), dim]; got %s for %d frames" % (f.shape, lengths.sum()))
        return f, lengths

    def forward(self, feats, lengths, node=""):
        """Packed features [sum(lengths), dim] fp32 -> [n_utts, width] fp32 for an utterance-level node."""
        f, lengths = self._pack(feats, lengths)
        if f.shape[1] != self.dim:
            raise ValueError("feature dim %d != engine dim %d" % (f.shape[1], self.dim))
        width, frame_level, _ = self.node_info(node)
        if frame_level:
            raise ValueError("node %r is frame level; use forward_frames" % node)
        out = np.empty((len(lengths), width), np.float32)
        self._check(self.lib.xv_forward(self._ctx, f.ctypes.data_as(C.c_void_p), _i32ptr(lengths), len(lengths),
                                        (node or "").encode("utf-8"), _fptr(out), out.size))
        return out

    def forward_views(self, feats, view_row0, lengths, node=""):
        """Utterances as views into one packed feature buffer (xv_forward_views): view i = rows
        [view_row0[i], view_row0[i] + lengths[i]) of feats [rows, dim]; views may overlap (the extractor's chunks),
        ascending view_row0.  -> [n_views, width] fp32."""
        f = feats if (isinstance(feats, np.ndarray) and feats.dtype == np.float32 and feats.flags["C_CONTIGUOUS"]) else _f32(feats)
        if f.ndim != 2 or f.shape[1] != self.dim:
            raise ValueError("feats must be [rows, %d]; got %s" % (self.dim, f.shape))
        row0 = np.ascontiguousarray(view_row0, dtype=np.int64)
        lengths = np.ascontiguousarray(lengths, dtype=np.int32)
        width, frame_level, _ = self.node_info(node)
        if frame_level:
            raise ValueError("node %r is frame level" % node)
        out = np.empty((len(lengths), width), np.float32)
        self._check(self.lib.xv_forward_views(self._ctx, f.ctypes.data_as(C.c_void_p), f.shape[0], _i64ptr(row0), _i32ptr(lengths),
                                              len(lengths), (node or "").encode("utf-8"), _fptr(out), out.size))
        return out

    def stage_views(self, feats, view_row0, lengths):
        f = feats if (isinstance(feats, np.ndarray) and feats.dtype == np.float32 and feats.flags["C_CONTIGUOUS"]) else _f32(feats)
        row0 = np.ascontiguousarray(view_row0, dtype=np.int64)
        lengths = np.ascontiguousarray(lengths, dtype=np.int32)
        self._n_staged = len(lengths)
        self._check(self.lib.xv_stage_views(self._ctx, f.ctypes.data_as(C.c_void_p), f.shape[0], _i64ptr(row0), _i32ptr(lengths),
                                            len(lengths)))

    def forward_frames(self, feats, lengths, node):
        """Frame-level node -> list of per-utterance arrays [T_u - context, width] ([H, T_u - context] for
        attention_weights)."""
        f, lengths = self._pack(feats, lengths)
        width, frame_level, ctx = self.node_info(node)
        if not frame_level:
            raise ValueError("node %r is utterance level; use forward" % node)
        valid = lengths.astype(np.int64) - ctx
        sizes = valid * width
        offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
        out = np.empty(int(offs[-1]), np.float32)
        self._check(self.lib.xv_forward_frames(self._ctx, f.ctypes.data_as(C.c_void_p), _i32ptr(lengths), len(lengths),
                                               node.encode("utf-8"), _fptr(out), _i64ptr(offs), out.size))
        res = []
        for u in range(len(lengths)):
            blk = out[offs[u]:offs[u + 1]]
            res.append(blk.reshape(width, valid[u]) if node == "attention_weights" else blk.reshape(valid[u], width))
        return res

    # -- staged API (bench)
    def stage(self, feats, lengths):
        f, lengths = self._pack(feats, lengths)
        self._n_staged = len(lengths)
        self._check(self.lib.xv_stage(self._ctx, f.ctypes.data_as(C.c_void_p), _i32ptr(lengths), len(lengths)))

    def run(self, node=""):
        self._check(self.lib.xv_run(self._ctx, (node or "").encode("utf-8")))

    def fetch(self, node=""):
        width, _, _ = self.node_info(node)
        out = np.empty((self._n_staged, width), np.float32)
        self._check(self.lib.xv_fetch(self._ctx, _fptr(out), out.size))
        return out

    def sync(self):
        self._check(self.lib.xv_sync(self._ctx))

    # -- standalone pooling
    def stat_pooling(self, x, cu_lens, dtype="fp32"):
        """x [rows, C] host fp32 (converted to bf16 on the device when dtype='bf16'); returns [n, 2C]."""
        x = _f32(x)
        cu = np.ascontiguousarray(cu_lens, dtype=np.int32)
        n, c = len(cu) - 1, x.shape[1]
        out = np.empty((n, 2 * c), np.float32)
        if dtype == "bf16":
            dev = self.device_alloc(x.size * 2)
            try:
                self.upload_bf16(dev, x)
                self._check(self.lib.xv_stat_pooling(self._ctx, dev, _i32ptr(cu), n, c, _lib.DTYPE_BF16, _fptr(out)))
            finally:
                self.device_free(dev)
        else:
            self._check(self.lib.xv_stat_pooling(self._ctx, x.ctypes.data_as(C.c_void_p), _i32ptr(cu), n, c, _lib.DTYPE_F32,
                                                 _fptr(out)))
        return out

    def attention_pooling(self, value, key, query, cu_lens, split_value, split_key, use_scale, want_weights=False):
        value, key, query = _f32(value), _f32(key), _f32(query)
        cu = np.ascontiguousarray(cu_lens, dtype=np.int32)
        n, cv, ck, h = len(cu) - 1, value.shape[1], key.shape[1], query.shape[0]
        dv = cv // h if split_value else cv
        out = np.empty((n, 2 * h * dv), np.float32)
        wts = np.empty(h * int(cu[-1]), np.float32) if want_weights else None
        self._check(self.lib.xv_attention_pooling(
            self._ctx, value.ctypes.data_as(C.c_void_p), key.ctypes.data_as(C.c_void_p), _fptr(query), _i32ptr(cu), n, cv, ck,
            h, int(bool(split_value)), int(bool(split_key)), int(bool(use_scale)), _lib.DTYPE_F32, _fptr(out),
            _fptr(wts) if want_weights else None))
        if not want_weights:
            return out
        per_utt = [wts[h * cu[u]:h * cu[u + 1]].reshape(h, cu[u + 1] - cu[u]) for u in range(n)]
        return out, per_utt

    def dense_rows(self, x, kernel, bias, bn_scale=None, bn_shift=None, slope=None, act_tanh=False, taps=1):
        x, kernel, bias = _f32(x), _f32(kernel), _f32(bias)
        rows, cin = x.shape
        n = kernel.shape[-1]
        out = np.empty((rows - taps + 1, n), np.float32)
        opt = lambda a: _fptr(_f32(a)) if a is not None else None
        keep = [None if a is None else _f32(a) for a in (bn_scale, bn_shift, slope)]
        self._check(self.lib.xv_dense_rows_f32(
            self._ctx, _fptr(x), rows, cin, taps, _fptr(kernel.reshape(-1, n)), n, _fptr(bias),
            _fptr(keep[0]) if keep[0] is not None else None, _fptr(keep[1]) if keep[1] is not None else None,
            _fptr(keep[2]) if keep[2] is not None else None, int(bool(act_tanh)), _fptr(out)))
        return out

    def prelu(self, x, alpha, shared=False):
        x = _f32(x)
        x2 = x.reshape(-1, x.shape[-1])
        alpha = _f32(alpha).reshape(-1)
        out = np.empty_like(x2)
        self._check(self.lib.xv_prelu_f32(self._ctx, _fptr(x2), x2.shape[0], x2.shape[1], _fptr(alpha), int(bool(shared)),
                                          _fptr(out)))
        return out.reshape(x.shape)

    def l2_scaling(self, x, factor, eps=1e-12):
        x = _f32(x)
        x2 = x.reshape(-1, x.shape[-1])
        out = np.empty_like(x2)
        self._check(self.lib.xv_l2_scaling_f32(self._ctx, _fptr(x2), x2.shape[0], x2.shape[1], float(factor), float(eps),
                                               _fptr(out)))
        return out.reshape(x.shape)

    # -- training step (config 5)
    def train_init(self, num_speakers):
        self._check(self.lib.xv_train_init(self._ctx, int(num_speakers)))
        self._train_shapes = {}

    def train_step(self, feats, labels, learning_rate, global_step=0):
        """feats [B, T, D] fp32, labels [B] int32 -> (loss, total_loss)."""
        x = _f32(feats)
        y = np.ascontiguousarray(labels, dtype=np.int32)
        assert x.ndim == 3 and y.shape == (x.shape[0],)
        out = np.zeros(2, np.float32)
        self._check(self.lib.xv_train_step(self._ctx, _fptr(x), _i32ptr(y), x.shape[0], x.shape[1], float(learning_rate),
                                           int(global_step), _fptr(out)))
        return float(out[0]), float(out[1])

    def train_submit(self, feats, labels, learning_rate, global_step=0):
        """First half of a pipelined step (xv_train_submit): copies the batch on the copy stream and enqueues the step behind
        it without waiting.  At most two submitted steps may be uncollected.  The arrays are kept alive until the step is
        collected; a caller that reuses its own (pinned) buffers must not overwrite them before that."""
        x = _f32(feats)
        y = np.ascontiguousarray(labels, dtype=np.int32)
        assert x.ndim == 3 and y.shape == (x.shape[0],)
        self._check(self.lib.xv_train_submit(self._ctx, _fptr(x), _i32ptr(y), x.shape[0], x.shape[1], float(learning_rate),
                                             int(global_step)))
        if not hasattr(self, "_submitted"):
            self._submitted = []
        self._submitted.append((x, y))

    def train_collect(self):
        """Second half: waits for the oldest submitted step -> (loss, total_loss)."""
        out = np.zeros(2, np.float32)
        rc = self.lib.xv_train_collect(self._ctx, _fptr(out))
        if getattr(self, "_submitted", None):
            self._submitted.pop(0)
        self._check(rc)
        return float(out[0]), float(out[1])

    def train_step_resident(self, feats_dev, labels_dev, batch, frames, learning_rate, global_step=0):
        """xv_train_step on a batch that already lives in device buffers of the caller (device_alloc + h2d); only enqueues
        the step (loss_out = NULL): read the loss with a later train_step / train_collect, or sync()."""
        self._check(self.lib.xv_train_step(self._ctx, feats_dev, labels_dev, int(batch), int(frames), float(learning_rate),
                                           int(global_step), None))

    def _train_read(self, fn, name, shape):
        out = np.empty(shape, np.float32)
        self._check(fn(self._ctx, name.encode("utf-8"), _fptr(out), out.size))
        return out

    def train_weight(self, name, shape):
        return self._train_read(self.lib.xv_train_get_weight, name, shape)

    def train_grad(self, name, shape):
        return self._train_read(self.lib.xv_train_get_grad, name, shape)

    def train_variable_names(self):
        """[(TF variable name, element count)] of everything the training state holds (what tf.train.Saver would save)."""
        n = self.lib.xv_train_num_variables(self._ctx)
        if n < 0:
            raise XVectorError(_lib.XV_ESTATE, "no training state (call train_init)")
        out = []
        buf = C.create_string_buffer(512)
        size = C.c_int64()
        for i in range(n):
            self._check(self.lib.xv_train_variable_name(self._ctx, i, buf, 512, C.byref(size)))
            out.append((buf.value.decode("utf-8"), int(size.value)))
        return out

    def train_variables(self):
        """{TF variable name: current value}, in the shapes the variables were given with set_weights."""
        out = {}
        shapes = getattr(self, "_weight_shapes", {})
        for name, size in self.train_variable_names():
            shape = shapes.get(name, (size,))
            if int(np.prod(shape, dtype=np.int64)) != size:
                shape = (size,)
            out[name] = self.train_weight(name, shape)
        return out

    def comm_info(self):
        rank, world, buckets = C.c_int32(), C.c_int32(), C.c_int32()
        elems = C.c_int64()
        self._check(self.lib.xv_comm_info(self._ctx, C.byref(rank), C.byref(world), C.byref(elems), C.byref(buckets)))
        dt = self.lib.xv_comm_dtype(self._ctx)
        return {"rank": int(rank.value), "world": int(world.value), "grad_elems": int(elems.value), "buckets": int(buckets.value),
                "dtype": "bf16" if dt == _lib.DTYPE_BF16 else "fp32"}

    def comm_unique_id(self):
        buf = C.create_string_buffer(128)
        self._check(self.lib.xv_comm_unique_id(self._ctx, buf))
        return buf.raw

    def comm_init(self, unique_id, rank, world):
        self._check(self.lib.xv_comm_init(self._ctx, C.create_string_buffer(bytes(unique_id), 128), int(rank), int(world)))

    # -- measurement / memory
    def event_record(self, slot):
        self._check(self.lib.xv_event_record(self._ctx, slot))

    def elapsed_ms(self, a, b):
        ms = C.c_float()
        self._check(self.lib.xv_event_elapsed_ms(self._ctx, a, b, C.byref(ms)))
        return float(ms.value)

    def launch_count(self):
        return int(self.lib.xv_launch_count(self._ctx))

    def launch_summary(self):
        buf = C.create_string_buffer(2048)
        self._check(self.lib.xv_launch_summary(self._ctx, buf, 2048))
        out = {}
        for part in buf.value.decode().split(";"):
            if "=" in part:
                k, v = part.split("=")
                out[k] = int(v)
        return out

    def profile_enable(self, on=True):
        self._check(self.lib.xv_profile_enable(self._ctx, int(on)))

    def profile_query(self, family):
        ms, cnt = C.c_float(), C.c_int64()
        self._check(self.lib.xv_profile_query(self._ctx, family.encode(), C.byref(ms), C.byref(cnt)))
        return float(ms.value), int(cnt.value)

    # -- post-extraction vector operations
    def normalize_length(self, x, scaleup=False):
        """ivector-normalize-length [--scaleup=false]: rows of x [n, dim] scaled to unit length (or to sqrt(dim))."""
        x = _f32(x)
        out = np.empty_like(x)
        self._check(self.lib.xv_vec_normalize_length(self._ctx, _fptr(x), x.shape[0], x.shape[1], int(bool(scaleup)), _fptr(out)))
        return out

    def group_mean(self, x, group, n_groups):
        """ivector-mean: -> (means [n_groups, dim], counts [n_groups]); group[i] < 0 leaves vector i out."""
        x = _f32(x)
        group = np.ascontiguousarray(group, np.int32)
        out = np.empty((n_groups, x.shape[1]), np.float32)
        counts = np.zeros(n_groups, np.int32)
        self._check(self.lib.xv_vec_group_mean(self._ctx, _fptr(x), x.shape[0], x.shape[1], group.ctypes.data_as(C.POINTER(C.c_int32)),
                                               n_groups, _fptr(out), counts.ctypes.data_as(C.POINTER(C.c_int32))))
        return out, counts

    def cosine_scores(self, a, b=None, eps=1e-12):
        """[na, nb] cosine similarities (misc/utils.py:318,326)."""
        a = _f32(a)
        b = a if b is None else _f32(b)
        out = np.empty((a.shape[0], b.shape[0]), np.float32)
        self._check(self.lib.xv_cosine_scores(self._ctx, _fptr(a), a.shape[0], _fptr(b), b.shape[0], a.shape[1], C.c_float(eps), _fptr(out)))
        return out

    def flush_l2(self):
        self._check(self.lib.xv_flush_l2(self._ctx))

    def device_alloc(self, nbytes):
        p = self.lib.xv_device_alloc(self._ctx, int(nbytes))
        if not p:
            self._check(_lib.XV_ENOMEM)
        return C.c_void_p(p)

    def device_free(self, p):
        self.lib.xv_device_free(self._ctx, p)

    def h2d(self, dst, arr):
        arr = np.ascontiguousarray(arr)
        self._check(self.lib.xv_memcpy_h2d(self._ctx, dst, arr.ctypes.data_as(C.c_void_p), arr.nbytes))

    def d2h(self, arr, src):
        self._check(self.lib.xv_memcpy_d2h(self._ctx, arr.ctypes.data_as(C.c_void_p), src, arr.nbytes))

    def upload_bf16(self, dst, arr):
        arr = _f32(arr)
        self._check(self.lib.xv_upload_bf16(self._ctx, dst, _fptr(arr), arr.size))
