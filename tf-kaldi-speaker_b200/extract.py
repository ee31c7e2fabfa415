"""Embedding extraction with the reference extractor's semantics (egs/voxceleb/v1/nnet/lib/extract.py:64-93):
skip short utterances, split long ones into overlapping chunks (hop = chunk_size/2), embed, optionally
L2-normalise, average the chunk embeddings weighted by chunk length.

`extract_one` is the literal per-utterance loop body (two predict calls for a chunked utterance).
`extract_many` is the B200 form: every utterance and every chunk of the whole list goes down in ONE packed
xv_forward call (utterances stay independent, so results are identical to the per-utterance loop)."""
import argparse
import os
import sys

import numpy as np

from .scheduler import chunk_plan


def extract_one(trainer, feature, chunk_size=10000, min_chunk_size=25, normalize=False):
    """One utterance [L, D] -> embedding, or None if skipped (L < min_chunk_size)."""
    if feature.shape[0] < min_chunk_size:
        return None
    if feature.shape[0] > chunk_size:
        plan = chunk_plan(feature.shape[0], chunk_size)
        feature_array = [feature[s:s + z] for s, z in plan]
        feature_length = np.expand_dims(np.array([z for _, z in plan]), axis=1)
        embeddings = trainer.predict(np.array(feature_array[:-1], dtype=np.float32))
        embedding_last = trainer.predict(feature_array[-1])
        embeddings = np.concatenate([embeddings, np.expand_dims(embedding_last, axis=0)], axis=0)
        if normalize:
            embeddings /= np.sqrt(np.sum(np.square(embeddings), axis=1, keepdims=True))
        embedding = np.sum(embeddings * feature_length, axis=0) / np.sum(feature_length)
    else:
        embedding = trainer.predict(feature)
    if normalize:
        embedding /= np.sqrt(np.sum(np.square(embedding)))
    return embedding


def plan_many(lengths, chunk_size=10000, min_chunk_size=25):
    """Work list for a set of utterances: (utt index, start, size) per chunk, skipping short utterances."""
    work = []
    for u, n in enumerate(lengths):
        n = int(n)
        if n < min_chunk_size:
            continue
        if n > chunk_size:
            work += [(u, s, z) for s, z in chunk_plan(n, chunk_size)]
        else:
            work.append((u, 0, n))
    return work


def extract_many(trainer, features_list, chunk_size=10000, min_chunk_size=25, normalize=False):
    """List of [L_u, D] arrays -> list of embeddings (None where skipped), one packed device call."""
    lengths = [f.shape[0] for f in features_list]
    work = plan_many(lengths, chunk_size, min_chunk_size)
    out = [None] * len(features_list)
    if not work:
        return out
    pieces = [features_list[u][s:s + z] for u, s, z in work]
    emb = trainer.predict_ragged(pieces).astype(np.float32)
    sizes = np.array([z for _, _, z in work], np.float64)
    owners = np.array([u for u, _, _ in work])
    for u in np.unique(owners):
        idx = np.nonzero(owners == u)[0]
        e = emb[idx]
        if len(idx) > 1:
            if normalize:
                e = e / np.sqrt(np.sum(np.square(e), axis=1, keepdims=True))
            e = np.sum(e * sizes[idx, None], axis=0) / np.sum(sizes[idx])
        else:
            e = e[0]
        if normalize:
            e = e / np.sqrt(np.sum(np.square(e)))
        out[u] = e
    return out


def extract_stream(trainer, rspecifier, wspecifier, dim, chunk_size=10000, min_chunk_size=25, normalize=False,
                   max_rows=1 << 20, max_utts=4096, pinned=True):
    """The reference's extraction loop (extract.py:63-93) as a stream of packed batches: the native ark reader
    (kaldi_io.BatchReader, lib/libxvkaldi.so) decodes whole batches into one pinned buffer, xv_forward embeds the
    batch, the native writer emits the vectors.  Batches whose utterances all lie in [min_chunk_size, chunk_size]
    go down without any host-side copy; others take extract_many's chunk plan.  Returns (written, skipped)."""
    from . import kaldi_io
    pin = None
    buffer = None
    if pinned:
        from .engine import PinnedArray
        pin = PinnedArray((max_rows, dim))
        buffer = pin.array
    written = skipped = 0
    try:
        with kaldi_io.open_or_fd(wspecifier, "wb") as fp_out:
            for keys, feats, lengths in kaldi_io.BatchReader(rspecifier, dim, max_rows=max_rows, max_utts=max_utts, buffer=buffer):
                if lengths.min() >= min_chunk_size and lengths.max() <= chunk_size:
                    emb = np.asarray(trainer.predict_packed(feats, lengths), np.float32)
                    if normalize:
                        emb = emb / np.sqrt(np.sum(np.square(emb), axis=1, keepdims=True))
                    kaldi_io.write_vecs_flt(fp_out, keys, emb)
                    written += len(keys)
                    continue
                offs = np.concatenate([[0], np.cumsum(lengths)])
                embs = extract_many(trainer, [feats[offs[i]:offs[i + 1]] for i in range(len(keys))], chunk_size,
                                    min_chunk_size, normalize)
                keep = [i for i, e in enumerate(embs) if e is not None]
                skipped += len(keys) - len(keep)
                if keep:
                    kaldi_io.write_vecs_flt(fp_out, [keys[i] for i in keep], np.stack([embs[i] for i in keep]).astype(np.float32))
                    written += len(keep)
    finally:
        if pin is not None:
            pin.close()
    return written, skipped


def main(argv=None):
    """CLI with the reference's flags (extract.py:11-22)."""
    parser = argparse.ArgumentParser()
    parser.add_argument("-g", "--gpu", type=int, default=0, help="The GPU id.")
    parser.add_argument("-m", "--min-chunk-size", type=int, default=25)
    parser.add_argument("-s", "--chunk-size", type=int, default=10000)
    parser.add_argument("-n", "--normalize", action="store_true")
    parser.add_argument("--node", type=str, default="")
    parser.add_argument("--precision", type=str, default="bf16", choices=["bf16", "fp32"])
    parser.add_argument("model_dir", type=str)
    parser.add_argument("rspecifier", type=str)
    parser.add_argument("wspecifier", type=str)
    args = parser.parse_args(argv)

    from .misc.utils import Params
    from .model.trainer import Trainer
    config_json = os.path.join(args.model_dir, "nnet/config.json")
    if not os.path.isfile(config_json):
        sys.exit("Cannot find params.json in %s" % config_json)
    params = Params(config_json)
    if len(args.node) != 0:
        params.embedding_node = args.node
    with open(os.path.join(args.model_dir, "nnet", "feature_dim"), "r") as f:
        dim = int(f.readline().strip())
    trainer = Trainer(params, args.model_dir, dim, single_cpu=True, precision=args.precision, device=max(args.gpu, 0))
    trainer.build("predict")
    if args.rspecifier.rsplit(".", 1)[-1] == "scp":
        sys.exit("The rspecifier must be ark or input pipe")
    written, skipped = extract_stream(trainer, args.rspecifier, args.wspecifier, dim, args.chunk_size, args.min_chunk_size,
                                      args.normalize)
    print("[INFO] wrote %d embeddings, skipped %d short utterances" % (written, skipped))
    trainer.close()


if __name__ == "__main__":
    main()
