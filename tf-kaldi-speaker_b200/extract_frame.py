"""Frame-level extraction with the reference script's semantics (egs/voxceleb/v1/nnet/lib/extract_frame.py:67-96):
the network output at a frame-level node is shorter than the input by the total context of the layers below it; it is
padded back to the input length by repeating its first and last frame (pad = (L - L_out) / 2 on each side), and
written as a matrix per utterance.  Long utterances are cut into consecutive chunks of `chunk_size` frames, each
embedded and padded on its own.

The reference's chunked branch is marked "TODO: has not been checked" (:82) and cannot run as written (np.tile with a
rank-3 repetition on a rank-2 slice); this module implements what it describes: per-chunk edge padding, then
concatenation, so that the output always has one row per input frame (the assert at :94)."""
import argparse
import os
import sys

import numpy as np


def pad_edges(frames, length):
    """[L_out, C] -> [length, C] by repeating the first / last row (extract_frame.py:90-92)."""
    pad = (length - frames.shape[0]) // 2
    out = np.concatenate([np.tile(frames[0, :], [pad, 1]), frames, np.tile(frames[-1, :], [pad, 1])], axis=0)
    if out.shape[0] != length:
        raise ValueError("cannot pad %d frames to %d symmetrically (odd total context)" % (frames.shape[0], length))
    return out


def chunk_plan_frames(length, chunk_size):
    """Consecutive chunks (no overlap): extract_frame.py:72-79."""
    n = int(np.ceil(float(length) / chunk_size))
    return [(i * chunk_size, min(chunk_size, length - i * chunk_size)) for i in range(n)]


def extract_frames_one(trainer, feature, chunk_size=10000, min_chunk_size=25):
    """One utterance [L, D] -> [L, C] frame-level outputs of params.embedding_node, or None if skipped."""
    length = feature.shape[0]
    if length < min_chunk_size:
        return None
    if length > chunk_size:
        plan = chunk_plan_frames(length, chunk_size)
        pieces = [feature[s:s + z] for s, z in plan]
        full = trainer.predict(np.array(pieces[:-1], dtype=np.float32))            # [n-1, chunk - context, C]
        context = chunk_size - full.shape[1]
        out = [pad_edges(full[i], chunk_size) for i in range(full.shape[0])]
        if pieces[-1].shape[0] <= context:
            # a tail shorter than the receptive field has no output frame (the reference would fail here): embed it
            # together with the chunk before it
            tail = np.concatenate(pieces[-2:], axis=0)
            out[-1] = pad_edges(trainer.predict(tail), tail.shape[0])
        else:
            out.append(pad_edges(trainer.predict(pieces[-1]), pieces[-1].shape[0]))
        embedding = np.concatenate(out, axis=0)
    else:
        embedding = pad_edges(trainer.predict(feature), length)
    assert embedding.shape[0] == length
    return embedding


def extract_attention_one(trainer, feature, chunk_size=10000, min_chunk_size=25):
    """Attention weights of one utterance (egs/voxceleb/v1/nnet/lib/extract_attention.py:52-60): only the first
    `chunk_size` frames of a long utterance are used; params.embedding_node must be "attention_weights"."""
    if feature.shape[0] < min_chunk_size:
        return None
    if feature.shape[0] > chunk_size:
        feature = feature[:chunk_size]
    return trainer.predict(feature)


def _main(argv, attention):
    parser = argparse.ArgumentParser()
    parser.add_argument("-g", "--gpu", type=int, default=0, help="The GPU id.")
    parser.add_argument("-m", "--min-chunk-size", type=int, default=25)
    parser.add_argument("-s", "--chunk-size", type=int, default=10000)
    if not attention:
        parser.add_argument("--node", type=str, default="")
    parser.add_argument("--precision", type=str, default="bf16", choices=["bf16", "fp32"])
    parser.add_argument("model_dir", type=str)
    parser.add_argument("rspecifier", type=str)
    parser.add_argument("wspecifier", type=str)
    args = parser.parse_args(argv)

    from . import kaldi_io
    from .misc.utils import Params
    from .model.trainer import Trainer
    config_json = os.path.join(args.model_dir, "nnet/config.json")
    if not os.path.isfile(config_json):
        sys.exit("Cannot find params.json in %s" % config_json)
    params = Params(config_json)
    if attention:
        params.embedding_node = "attention_weights"                              # extract_attention.py:40
    elif len(args.node) != 0:
        params.embedding_node = args.node
    with open(os.path.join(args.model_dir, "nnet", "feature_dim"), "r") as f:
        dim = int(f.readline().strip())
    trainer = Trainer(params, args.model_dir, dim, single_cpu=True, precision=args.precision, device=max(args.gpu, 0))
    trainer.build("predict")
    if args.rspecifier.rsplit(".", 1)[-1] == "scp":
        sys.exit("The rspecifier must be ark or input pipe")
    one = extract_attention_one if attention else extract_frames_one
    with kaldi_io.open_or_fd(args.wspecifier, "wb") as fp_out:
        for key, feature in kaldi_io.read_mat_ark(args.rspecifier):
            out = one(trainer, feature, args.chunk_size, args.min_chunk_size)
            if out is not None:
                kaldi_io.write_mat(fp_out, np.ascontiguousarray(out, np.float32), key=key)
    trainer.close()


def main(argv=None):
    """CLI of extract_frame.py (flags :11-21)."""
    _main(argv, attention=False)


def main_attention(argv=None):
    """CLI of extract_attention.py."""
    _main(argv, attention=True)


if __name__ == "__main__":
    main()
