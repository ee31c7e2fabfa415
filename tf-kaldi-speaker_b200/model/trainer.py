"""Trainer / TrainerMGPU: the reference's driver classes (model/trainer.py:87-219, 270-376, 379-405, 590-673, 886-913,
1107-1248) with the TF session replaced by one xv_ctx per GPU.  Signatures are kept; the bodies are ctypes calls into
libxvector.so.

  predict path : build("predict") / load / set_embedding / predict  (extract.py's use)
  training path: build("train") / train_step (the body of the reference's step loop, sess.run(train_op, ...) at
                 model/trainer.py:652) / save (tf.train.Saver bundle) / train (epoch loop over a Kaldi data directory);
                 TrainerMGPU = the same over `num_gpus` GPUs: the batch is cut into equal tower slices
                 (:1133-1150), every GPU runs xv_train_step on its slice from its own host thread, gradients are
                 averaged by the library's bucketed ncclAllReduce (== add_n/N, :1240-1244).
Validation (build("valid") / valid()), summaries and the insight/finetune variants are outside this build's scope."""
import concurrent.futures
import os
import sys

import numpy as np

from .. import tf_checkpoint
from .. import train_dp
from ..engine import Engine
from . import variables as V
from .tdnn import tdnn, etdnn, TDNN_RANK4_NODES  # noqa: F401  (kept importable from here like the reference)
from .resnet import resnet_18  # noqa: F401

_NETWORKS = ("tdnn", "extended_tdnn", "resnet_18")      # "tdnn-s" is dead code in the reference (quit() at model/tdnn.py:201)
_KNOWN_NETWORKS = ("tdnn", "tdnn-s", "extended_tdnn", "resnet_18")
_KNOWN_LOSSES = ("softmax", "asoftmax", "additive_margin_softmax", "additive_angular_margin_softmax",
                 "semihard_triplet_loss", "angular_triplet_loss", "generalized_angular_triplet_loss")
_TRAINABLE_LOSSES = ("softmax", "additive_margin_softmax")        # what xv_train_init covers (model/loss.py:9-48, 201-286)


class Trainer(object):
    """Handles network building, weight loading and prediction (the extraction half of the reference class)."""

    def __init__(self, params, model_dir, dim, num_speakers=None, single_cpu=False, num_gpus=1,
                 precision="bf16", device=0, batch_frames=None):
        """Same leading arguments as the reference (model/trainer.py:87).  `single_cpu`/`num_gpus` are accepted
        for call-site compatibility (extract.py:56 passes single_cpu=True) and ignored: the path runs on one B200
        (`device`).  `precision` is 'bf16' (tcgen05 path) or 'fp32' (validation mode)."""
        self.network_type = params.network_type
        if params.network_type not in _KNOWN_NETWORKS:
            raise NotImplementedError("Not implement %s network" % params.network_type)
        if params.network_type not in _NETWORKS:
            raise NotImplementedError("Not implement %s network (this build covers: %s)" %
                                      (params.network_type, ", ".join(_NETWORKS)))
        self.loss_func = params.dict.get("loss_func", "softmax")
        if self.loss_func not in _KNOWN_LOSSES:
            raise NotImplementedError("Not implement %s loss" % self.loss_func)
        self.params = params
        self.model = os.path.join(model_dir, "nnet")
        self.dim = dim
        self.num_speakers = num_speakers
        self.precision = precision
        self.device = device
        self.batch_frames = batch_frames
        self.is_built = False
        self.is_loaded = False
        self.first_feature_split_alert = True
        self.engine = None
        self.engines = []
        self._pool = None
        self.embeddings = None
        self.endpoints = None
        self._weights = None
        self.train_built = False
        self.curr_step = 0
        self._predict_stale = False

    def reset(self):
        """model/trainer.py:248-268: drop the graph."""
        self.close()
        self.is_built = False
        self.is_loaded = False

    def close(self):
        """model/trainer.py:270-275."""
        for eng in self.engines:
            if eng is not self.engine:
                eng.close()
        self.engines = []
        if self._pool is not None:
            self._pool.shutdown(wait=True)
            self._pool = None
        if self.engine is not None:
            self.engine.close()
            self.engine = None
        self.train_built = False

    def trainable(self):
        """None when build("train") can serve this configuration, else the reason (the training step covers network_type
        tdnn + statistics_pooling with softmax / additive_margin_softmax; see README "What's missing")."""
        if self.network_type != "tdnn":
            return "Not implement training of the %s network (the training step covers network_type tdnn)" % self.network_type
        if self.params.dict.get("pooling_type") != "statistics_pooling":
            return "Not implement training with %s pooling (the training step covers statistics_pooling)" % self.params.dict.get("pooling_type")
        if self.loss_func not in _TRAINABLE_LOSSES:
            return "Not implement %s loss (the training step covers: %s)" % (self.loss_func, ", ".join(_TRAINABLE_LOSSES))
        if self.params.dict.get("feature_norm", False):
            return "Not implement training with feature_norm"
        if self.params.dict.get("aux_loss_func"):
            return "Not implement auxiliary losses %s" % (self.params.dict.get("aux_loss_func"),)
        return None

    def _train_devices(self):
        return [self.device]

    def build(self, mode, noupdate_var_list=None):
        """model/trainer.py:309-376.  "predict" builds the embedding path; "train" builds the training step (network in
        training mode + loss + gradients + optimizer, :340-376); "valid" is accepted after "train" for call-site
        compatibility (train.py:105-106 builds both): valid() runs the embedding path in eval mode and the loss head on the host."""
        assert (mode == "train" or mode == "valid" or mode == "predict")
        if mode == "predict":
            if self.engine is None:
                self.engine = Engine(self.params.dict, self.dim, precision=self.precision, device=self.device,
                                     batch_frames=self.batch_frames)
            self.embeddings = self.params.embedding_node
            self.is_built = True
            self.is_loaded = False
            if self._weights is not None:
                self.set_weights(self._weights)
            return
        why = self.trainable()
        if why is not None:
            raise NotImplementedError(why)
        if noupdate_var_list:
            raise NotImplementedError("build(%r): noupdate_var_list (fine-tuning with frozen variables) is outside this build's scope" % mode)
        if mode == "valid":
            if not self.train_built:
                raise NotImplementedError("build('valid') without build('train'): the validation graph is outside this build's scope")
            return
        if self.num_speakers is None:
            raise ValueError("build('train') needs num_speakers (softmax-like losses, model/trainer.py:92)")
        if self.train_built:
            return
        self.close()
        weights, self.curr_step = self._initial_or_saved_variables()
        devices = self._train_devices()
        self.engines = []
        for dev in devices:
            eng = Engine(self.params.dict, self.dim, precision=self.precision, device=dev, batch_frames=self.batch_frames)
            eng.set_weights(weights)
            eng.train_init(self.num_speakers)
            self.engines.append(eng)
        self.engine = self.engines[0]
        world = len(devices)
        if world > 1:
            uid = self.engine.comm_unique_id()
            self._pool = concurrent.futures.ThreadPoolExecutor(max_workers=world)
            # ncclCommInitRank blocks until every rank has joined: one host thread per GPU
            list(self._pool.map(lambda r: self.engines[r].comm_init(uid, r, world), range(world)))
        self._weights = weights
        self.embeddings = self.params.dict.get("embedding_node", "tdnn6_dense")
        self.is_built = True
        self.is_loaded = True
        self.train_built = True
        self._predict_stale = False

    def _initial_or_saved_variables(self):
        """Variables at the start of training: injected ones (set_weights), else the checkpoint of the model directory
        (model/trainer.py:606-607), else fresh initial values (tf.global_variables_initializer, :601)."""
        if self._weights is not None:
            return dict(self._weights), 0
        if os.path.isfile(os.path.join(self.model, "checkpoint")) and self._has_tf_checkpoint():
            return tf_checkpoint.load_weights(self.model)
        return V.initial_variables(self.params, self.dim, self.num_speakers, seed=self.params.dict.get("seed", 0)), 0

    def train_step(self, features, labels, learning_rate, global_step=None):
        """One optimizer step on a batch: features [B, T, D] fp32, labels [B] int32 -> (loss, total_loss) -- the
        sess.run([train_ops, train_op], {features, labels, is_training: True, global_step, learning_rate}) of
        model/trainer.py:634-656.  global_step defaults to the trainer's own counter, which the call advances."""
        if not self.train_built:
            sys.exit("The training graph has not been build. Call build('train') first.")
        features = np.asarray(features, np.float32)
        if features.shape[-1] != self.dim:
            features = features[:, :, :self.dim]            # model/trainer.py:626-627
        labels = np.ascontiguousarray(labels, np.int32)
        step = self.curr_step if global_step is None else int(global_step)
        world = len(self.engines)
        if world == 1:
            out = self.engine.train_step(features, labels, learning_rate, step)
        else:
            slices = [train_dp.slice_batch(features, labels, r, world) for r in range(world)]
            res = list(self._pool.map(lambda r: self.engines[r].train_step(slices[r][0], slices[r][1], learning_rate, step),
                                      range(world)))
            out = res[0]                                    # every rank reports the mean of the tower losses (:1218-1219)
        self.curr_step = step + 1
        self._predict_stale = True
        return out

    def train_submit(self, features, labels, learning_rate, global_step=None):
        """train_step as a two-stage pipeline (xv_train_submit / xv_train_collect): copies the batch and enqueues the step
        without waiting for it, so that the host prepares batch k+1 while step k runs.  At most two submitted steps may be
        uncollected; train_collect() returns their (loss, total_loss) in order."""
        if not self.train_built:
            sys.exit("The training graph has not been build. Call build('train') first.")
        features = np.asarray(features, np.float32)
        if features.shape[-1] != self.dim:
            features = features[:, :, :self.dim]
        labels = np.ascontiguousarray(labels, np.int32)
        step = self.curr_step if global_step is None else int(global_step)
        world = len(self.engines)
        if world == 1:
            self.engine.train_submit(features, labels, learning_rate, step)
        else:
            slices = [train_dp.slice_batch(features, labels, r, world) for r in range(world)]
            list(self._pool.map(lambda r: self.engines[r].train_submit(slices[r][0], slices[r][1], learning_rate, step),
                                range(world)))
        self.curr_step = step + 1
        self._predict_stale = True

    def train_collect(self):
        if len(self.engines) == 1:
            return self.engine.train_collect()
        return list(self._pool.map(lambda e: e.train_collect(), self.engines))[0]

    def variables(self):
        """{TF variable name: current value} of the training state (replica 0; replicas are bit-identical)."""
        return self.engine.train_variables()

    def save(self, step):
        """model/trainer.py:296-303: saver.save(sess, <model>/model, global_step=step) -- a tensor bundle + the
        `checkpoint` state file, keep_checkpoint_max bundles kept (:382)."""
        return tf_checkpoint.save_checkpoint(self.model, step, self.variables(),
                                             max_to_keep=self.params.dict.get("keep_checkpoint_max", None))

    def train(self, data, spklist, learning_rate, aux_data=None):
        """model/trainer.py:590-673: one epoch of steps over a Kaldi data directory (feats.scp + utt2spk/spk2utt +
        spklist), checkpoints every save_checkpoints_steps and at the end.  The batches come from
        data_loader.KaldiDataRandomQueue's sampling rule restated over the native scp reader (data_loader.py)."""
        from ..data_loader import RandomSegmentLoader
        if not self.train_built:
            sys.exit("The training graph has not been build. Call build('train') first.")
        p = self.params.dict
        loader = RandomSegmentLoader(data, spklist, num_speakers=p["num_speakers_per_batch"],
                                     num_segments=p["num_segments_per_speaker"], min_len=p["min_segment_len"],
                                     max_len=p["max_segment_len"], dim=self.dim, seed=p.get("seed", 0) + self.curr_step,
                                     sample_with_prob=p.get("sample_with_prob", False))
        steps = p["num_steps_per_epoch"]
        last = [(0.0, 0.0)]
        pending = []                                     # (global step, step of the epoch) of the submitted, uncollected steps

        def collect_one():
            curr, step = pending.pop(0)
            last[0] = self.train_collect()
            if step % p.get("show_training_progress", 100) == 0:
                print("Epoch: [%2d] step: [%2d/%2d] raw loss: %f, total loss: %f" % (curr // steps, step, steps, last[0][0], last[0][1]),
                      file=sys.stderr)

        # two-stage pipeline: while step k runs on the GPU the loader assembles batch k+1 and its copy is already queued
        for step in range(self.curr_step % steps, steps):
            features, labels = loader.fetch()
            curr = self.curr_step
            self.train_submit(features, labels, learning_rate)
            pending.append((curr, step))
            if len(pending) == 2:
                collect_one()
            if step % p.get("save_checkpoints_steps", steps) == 0 and curr != 0:
                while pending:                           # a checkpoint holds exactly the steps up to and including `curr`
                    collect_one()
                self.save(curr)
        while pending:
            collect_one()
        self.save(self.curr_step)
        return last[0]

    def valid(self, data, spklist, batch_type="softmax", output_embeddings=False, aux_data=None):
        """model/trainer.py:756-884: the loss of the CURRENT variables over one sequential pass of the validation set, in
        eval mode (batch norm on its moving statistics) -- batches of num_speakers_per_batch * num_segments_per_speaker
        utterances cut to (min_segment_len + max_segment_len) / 2 frames from frame 0, at most valid_max_iterations of
        them, the mean of the batch losses (tf.metrics.mean).  -> (valid_loss, embeddings, labels); the last two are None
        unless output_embeddings (the caller computes EER from them, train.py:119-123).  The network output comes from
        the device path (node "output"), the loss head runs on the host (model/loss.py here).  batch_type "end2end"
        belongs to the end-to-end losses, which this build does not train."""
        from ..data_loader import DataOutOfRange, SequentialSegmentLoader
        from . import loss as host_loss
        assert batch_type == "softmax" or batch_type == "end2end", "The batch_type can only be softmax or end2end"
        if batch_type == "end2end":
            raise NotImplementedError("valid(batch_type='end2end'): the end-to-end losses are outside this build's scope")
        loss_func = self.params.dict.get("loss_func", "softmax")
        if loss_func not in _TRAINABLE_LOSSES:
            raise NotImplementedError("Not implement %s loss" % loss_func)
        curr_step = self.curr_step if self.train_built else 0
        if self.train_built:
            self._refresh_predict_weights()
        else:
            if not self.is_built:
                self.build("predict")
            if self._has_model():
                curr_step = self.load()
            elif self._weights is None:
                sys.exit("Cannot find model in %s" % self.model)
        variables = self._weights
        p = self.params.dict
        length = (int(p["min_segment_len"]) + int(p["max_segment_len"])) // 2
        loader = SequentialSegmentLoader(data, spklist, batch_size=int(p["num_speakers_per_batch"]) * int(p["num_segments_per_speaker"]),
                                         min_len=length, max_len=length, dim=self.dim, shuffle=False)
        losses, embeddings, all_labels = [], [], []
        try:
            for _ in range(int(p.get("valid_max_iterations", 1 << 30))):
                try:
                    features, labels = loader.fetch()
                except DataOutOfRange:
                    break
                b, t, d = features.shape
                packed = features.reshape(b * t, d)
                lengths = np.full(b, t, np.int32)
                out = self.engine.forward(packed, lengths, "output")
                losses.append(host_loss.valid_loss(loss_func, out, labels, variables, p, curr_step))
                if output_embeddings:
                    embeddings.append(np.asarray(self.engine.forward(packed, lengths, self.params.embedding_node), np.float32))
                    all_labels.append(labels)
        finally:
            loader.stop()
        valid_loss = float(np.mean(losses)) if losses else 0.0
        print("[Validation %d batches] valid loss: %f" % (len(losses), valid_loss), file=sys.stderr)
        if output_embeddings and embeddings:
            return valid_loss, np.concatenate(embeddings, axis=0), np.concatenate(all_labels, axis=0)
        return valid_loss, None, None

    def set_embedding(self, embedding_node):
        """model/trainer.py:305-307."""
        self.params.embedding_node = embedding_node
        self.embeddings = embedding_node

    def set_weights(self, weights):
        """Inject variables ({TF variable name: array}) instead of reading <model_dir>/nnet/weights.npz."""
        self._weights = weights
        if self.engine is not None:
            self.engine.set_weights(weights)
            self.is_loaded = True

    def load(self):
        """model/trainer.py:277-295.  Reads <model_dir>/nnet/weights.npz (variables keyed by TF name); returns the
        step parsed from the TF `checkpoint` state file when one is present, else 0."""
        if not self.is_built:
            sys.exit("The graph has not been build. Cannot load the models.")
        path = os.path.join(self.model, "weights.npz")
        step = 0
        if os.path.isfile(path):
            with np.load(path) as z:
                weights = {k: z[k] for k in z.files}
            state = tf_checkpoint.get_checkpoint_state(self.model)
            if state:
                try:
                    step = tf_checkpoint.checkpoint_step(state[0])
                except StopIteration:
                    step = 0
        elif self._has_tf_checkpoint():
            # the reference's own format: tf.train.Saver tensor bundle named by the `checkpoint` state file
            weights, step = tf_checkpoint.load_weights(self.model)
        else:
            sys.exit("Failed to find a checkpoint in {}".format(self.model))
        self.set_weights(weights)
        self.is_loaded = True
        return step

    def _has_model(self):
        return os.path.isfile(os.path.join(self.model, "weights.npz")) or self._has_tf_checkpoint()

    def _has_tf_checkpoint(self):
        state = tf_checkpoint.get_checkpoint_state(self.model)
        return bool(state) and os.path.isfile(os.path.join(self.model, os.path.basename(state[0])) + ".index")

    def predict(self, features):
        """Output the embeddings (model/trainer.py:886-913): rank-2 [T, D] -> [E]; rank-3 [B, T, D] -> [B, E];
        frame-level nodes give [T', N] / [B, T', N] -- [1, T', N] / [B, 1, T', N] for tdnn()'s conv2d endpoints, exactly the
        reference's shapes (model/tdnn.py:38-96).  Features wider than `dim` are sliced."""
        if not self.is_loaded:
            if self._has_model():
                self.load()
            else:
                sys.exit("Cannot find model in %s" % self.model)
        self._refresh_predict_weights()
        features = np.asarray(features)
        rank = len(features.shape)
        assert (rank == 2 or rank == 3)
        if rank == 2:
            features = np.expand_dims(features, axis=0)
        if self.first_feature_split_alert and features.shape[-1] != self.dim:
            self.first_feature_split_alert = False
        if features.shape[-1] != self.dim:
            features = features[:, :, :self.dim]
        b, t, d = features.shape
        packed = np.ascontiguousarray(features, dtype=np.float32).reshape(b * t, d)
        lengths = np.full(b, t, np.int32)
        embeddings = self._run(packed, lengths)
        if rank == 2:
            embeddings = np.squeeze(embeddings, axis=0)
        return embeddings

    def _refresh_predict_weights(self):
        """After train_step the embedding path must see the updated variables (one TF graph shares them; here the
        inference plan folds BN into its own copies): re-install the training state's variables."""
        if self._predict_stale and self.train_built:
            self._weights = self.variables()
            self.engine.set_weights(self._weights)
            self._predict_stale = False

    def _run(self, packed, lengths):
        node = self.params.embedding_node
        _, frame_level, _ = self.engine.node_info(node)
        if frame_level:
            out = np.stack(self.engine.forward_frames(packed, lengths, node), axis=0)
            if self.network_type == "tdnn" and node in TDNN_RANK4_NODES:
                out = np.expand_dims(out, 1)       # the reference's conv2d endpoints are [B, 1, T', 512] (model/tdnn.py:38-91)
            return out
        return self.engine.forward(packed, lengths, node)

    def predict_packed(self, packed, lengths):
        """Utterances already packed back to back ([sum T_u, D] float32, e.g. a pinned buffer filled by
        kaldi_io.BatchReader) + their lengths: the layout xv_forward takes, no host-side copy."""
        if not self.is_loaded:
            if self._has_model():
                self.load()
            else:
                sys.exit("Cannot find model in %s" % self.model)
        return self.engine.forward(packed, np.ascontiguousarray(lengths, np.int32), self.params.embedding_node)

    def predict_views(self, packed, view_row0, lengths):
        """Utterances (or chunks) as views into packed features: view i = rows [view_row0[i], view_row0[i] + lengths[i]);
        views may overlap (extract.py's chunk hop is chunk_size / 2).  One xv_forward_views call."""
        if not self.is_loaded:
            if self._has_model():
                self.load()
            else:
                sys.exit("Cannot find model in %s" % self.model)
        self._refresh_predict_weights()
        return self.engine.forward_views(packed, view_row0, lengths, self.params.embedding_node)

    def predict_ragged(self, features_list):
        """Extension of predict(): a list of [T_u, D] utterances of different lengths in ONE packed call.
        Row u of the result equals predict(features_list[u]) (each utterance is computed as a batch of one)."""
        if not self.is_loaded:
            if self._has_model():
                self.load()
            else:
                sys.exit("Cannot find model in %s" % self.model)
        lengths = np.array([f.shape[0] for f in features_list], np.int32)
        packed = np.concatenate([np.asarray(f, np.float32)[:, :self.dim] for f in features_list], axis=0)
        return self.engine.forward(packed, lengths, self.params.embedding_node)


class TrainerMGPU(Trainer):
    """Used for multi-gpu training (model/trainer.py:1107-1297): same constructor; build("train") creates one training
    context per GPU (devices 0..num_gpus-1, or `devices`), each driven by its own host thread; train_step takes the
    GLOBAL batch and cuts it into num_gpus equal tower slices (:1133-1150); BN statistics stay per tower and the moving
    averages come from tower 0 (:1183-1190); gradients are averaged by ncclAllReduce(sum) * 1/N over NVLink in three
    buckets overlapped with the backward pass (csrc/xv_train.cu) in place of the per-variable add_n/N on parameter-server
    devices (:1225-1248).  predict()/save() use replica 0 (replicas stay bit-identical)."""

    def __init__(self, params, model_dir, dim, num_speakers=None, single_cpu=False, num_gpus=1,
                 precision="bf16", device=0, batch_frames=None, devices=None):
        super(TrainerMGPU, self).__init__(params, model_dir, dim, num_speakers, single_cpu, num_gpus,
                                          precision=precision, device=device, batch_frames=batch_frames)
        self.num_gpus = num_gpus
        self.devices = list(devices) if devices is not None else list(range(num_gpus))
        if len(self.devices) != num_gpus:
            raise ValueError("devices must list num_gpus = %d GPUs" % num_gpus)
        self.device = self.devices[0]

    def _train_devices(self):
        return self.devices

    def build(self, mode, noupdate_var_list=None):
        if mode == "train":
            p = self.params.dict
            if "num_speakers_per_batch" in p and "num_segments_per_speaker" in p:
                batch = p["num_speakers_per_batch"] * p["num_segments_per_speaker"]
                if batch % self.num_gpus != 0:          # model/trainer.py:1133-1134 (Python-2 integer division there)
                    raise ValueError("the batch (%d) must be divisible by num_gpus (%d)" % (batch, self.num_gpus))
        return super(TrainerMGPU, self).build(mode, noupdate_var_list)
