"""Trainer: the reference's predict/extract driver (model/trainer.py:87-219, 270-338, 379-405, 886-913) with the
TF session replaced by one xv_ctx.  Signatures are kept; the body is a ctypes call into libxvector.so."""
import os
import re
import sys

import numpy as np

from .. import tf_checkpoint
from ..engine import Engine
from .tdnn import tdnn, etdnn, TDNN_RANK4_NODES  # noqa: F401  (kept importable from here like the reference)
from .resnet import resnet_18  # noqa: F401

_NETWORKS = ("tdnn", "extended_tdnn", "resnet_18")      # "tdnn-s" is dead code in the reference (quit() at model/tdnn.py:201)
_KNOWN_NETWORKS = ("tdnn", "tdnn-s", "extended_tdnn", "resnet_18")
_KNOWN_LOSSES = ("softmax", "asoftmax", "additive_margin_softmax", "additive_angular_margin_softmax",
                 "semihard_triplet_loss", "angular_triplet_loss", "generalized_angular_triplet_loss")


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
        self.embeddings = None
        self.endpoints = None
        self._weights = None

    def reset(self):
        """model/trainer.py:248-268: drop the graph."""
        self.close()
        self.is_built = False
        self.is_loaded = False

    def close(self):
        """model/trainer.py:270-275."""
        if self.engine is not None:
            self.engine.close()
            self.engine = None

    def build(self, mode, noupdate_var_list=None):
        """model/trainer.py:309-338.  Only mode == "predict" is implemented in this build."""
        assert (mode == "train" or mode == "valid" or mode == "predict")
        if mode != "predict":
            raise NotImplementedError("build(%r): the training/validation graph is outside this build's scope" % mode)
        if self.engine is None:
            self.engine = Engine(self.params.dict, self.dim, precision=self.precision, device=self.device,
                                 batch_frames=self.batch_frames)
        self.embeddings = self.params.embedding_node
        self.is_built = True
        self.is_loaded = False
        if self._weights is not None:
            self.set_weights(self._weights)

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
