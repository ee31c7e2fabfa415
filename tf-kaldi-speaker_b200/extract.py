"""Embedding extraction with the reference extractor's semantics (egs/voxceleb/v1/nnet/lib/extract.py:64-93):
skip short utterances, split long ones into overlapping chunks (hop = chunk_size/2), embed, optionally
L2-normalise, average the chunk embeddings weighted by chunk length.

`extract_one` is the per-utterance loop body (two predict calls for a chunked utterance).
`extract_many` is the B200 form for a list of arrays: every utterance and every chunk of the whole list goes down in
ONE packed xv_forward call (utterances stay independent, so results are identical to the per-utterance loop).
`extract_packed` is the same for utterances already packed back to back: chunks become overlapping views of the packed
rows (xv_forward_views), no host-side slicing or copying at all; the ark-to-ark stream uses it."""
import argparse
import os
import sys

import numpy as np

from .scheduler import chunk_plan, plan_arrays


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


def extract_packed(trainer, feats, lengths, chunk_size=10000, min_chunk_size=25, normalize=False):
    """extract_many for utterances that are already packed back to back (feats [sum(lengths), D] float32, e.g. the pinned
    buffer the native ark reader fills): the chunks of a long utterance (extract.py:68-78: chunk_size frames every
    chunk_size/2) are handed to the library as overlapping VIEWS of the packed rows (xv_forward_views), so nothing is
    sliced, copied or concatenated on the host and the features cross the bus once; the chunk embeddings are averaged
    weighted by chunk length (:86).  -> (embeddings [n_kept, E] float32, kept utterance indices); utterances shorter than
    min_chunk_size are left out (:65-67)."""
    lengths = np.asarray(lengths)
    offs = np.concatenate([[0], np.cumsum(lengths, dtype=np.int64)])
    owners, starts, sizes = plan_arrays(lengths, chunk_size, min_chunk_size)
    if owners.size == 0:
        return np.zeros((0, 0), np.float32), np.zeros(0, np.int64)
    sizes = sizes.astype(np.int32)
    row0 = offs[owners] + starts
    emb = trainer.predict_views(feats, row0, sizes)
    first = np.concatenate([[0], np.nonzero(np.diff(owners))[0] + 1])             # pieces of an utterance are consecutive
    kept = owners[first]
    counts = np.diff(np.concatenate([first, [owners.size]]))
    out = np.array(emb[first], np.float32)                                        # unchunked utterances: the embedding itself, bit for bit
    if counts.max() > 1:
        # chunked utterances only, grouped by their number of pieces (a handful of groups): [n, k, E] gathers summed over k in
        # fp64.  (np.add.reduceat over all 3432 x 512 piece rows of a 2048-utterance batch took 50 ms -- its axis-0 path is
        # a strided scalar loop -- against 3 ms here.)
        multi = np.nonzero(counts > 1)[0]
        for k in np.unique(counts[multi]):
            sel = multi[counts[multi] == k]
            rows = first[sel][:, None] + np.arange(int(k))[None, :]
            e = emb[rows].astype(np.float64)
            if normalize:
                e = e / np.sqrt(np.sum(np.square(e), axis=2, keepdims=True))      # per-chunk normalisation (:80-81) of chunked utterances only
            w = sizes[rows].astype(np.float64)[:, :, None]
            out[sel] = (np.sum(e * w, axis=1) / np.sum(w, axis=1)).astype(np.float32)
    if normalize:
        out = out / np.sqrt(np.sum(np.square(out), axis=1, keepdims=True))
    return out.astype(np.float32), kept


def extract_stream(trainer, rspecifier, wspecifier, dim, chunk_size=10000, min_chunk_size=25, normalize=False,
                   max_rows=1 << 20, max_utts=4096, pinned=True, threads=None, batch_rows=None):
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
            for keys, feats, lengths in kaldi_io.BatchReader(rspecifier, dim, max_rows=max_rows, max_utts=max_utts, buffer=buffer,
                                                             threads=threads or default_reader_threads(), batch_rows=batch_rows):
                if lengths.min() >= min_chunk_size and lengths.max() <= chunk_size:
                    emb = np.asarray(trainer.predict_packed(feats, lengths), np.float32)
                    if normalize:
                        emb = emb / np.sqrt(np.sum(np.square(emb), axis=1, keepdims=True))
                    kaldi_io.write_vecs_flt(fp_out, keys, emb)
                    written += len(keys)
                    continue
                emb, keep = extract_packed(trainer, feats, lengths, chunk_size, min_chunk_size, normalize)
                skipped += len(keys) - len(keep)
                if len(keep):
                    kaldi_io.write_vecs_flt(fp_out, [keys[i] for i in keep], emb)
                    written += len(keep)
    finally:
        if pin is not None:
            pin.close()
    return written, skipped


def default_reader_threads():
    """Decode threads of the native ark reader: every host core up to 32 (one B200 consumes ~150 M frames/s = 18 GB/s of
    fp32 features; 8 threads deliver ~10 GB/s out of the page cache)."""
    return max(1, min(32, os.cpu_count() or 8))


# Where the CLI cuts its batches (the pinned buffers keep their 1 M-row capacity for long recordings).  Measured on the
# ark -> x-vector ark leg (7 M frames, 1 / 2 GPUs): 2^19 rows 125 / 202 M frames/s, 2^18 rows 115 / 166 M -- a batch costs
# ~0.45 ms on the GPU side (the first slice's H2D before any kernel can start, the last D2H), which outweighs the shorter
# fill and drain of the reader -> GPU -> writer pipeline.
CLI_BATCH_ROWS = 1 << 19

_PIN_RING = {}


def _pinned_ring(n_buf, max_rows, dim):
    """Ring of pinned [max_rows, dim] buffers, kept for the life of the process: cudaHostAlloc of a 63 MB buffer costs tens
    of milliseconds, a whole extraction call of a few million frames not much more."""
    from .engine import PinnedArray
    ring = _PIN_RING.setdefault((int(max_rows), int(dim)), [])
    while len(ring) < n_buf:
        ring.append(PinnedArray((max_rows, dim)))
    return ring[:n_buf]


def release_pinned_rings():
    for ring in _PIN_RING.values():
        for p in ring:
            p.close()
    _PIN_RING.clear()


def extract_stream_mgpu(make_trainer, devices, rspecifier, wspecifier, dim, chunk_size=10000, min_chunk_size=25,
                        normalize=False, max_rows=1 << 20, max_utts=4096, threads=None, trainers=None, batch_rows=None):
    """extract_stream over several GPUs of one box: what `run.pl JOB=1:nj extract_wrapper.sh` does with nj processes
    (egs/voxceleb/v1/nnet/run_extract_embeddings.sh:68-71), as ONE process -- one reader (native ark decoder into a ring of
    pinned buffers), one host thread + one xv_ctx per GPU pulling batches from a queue (utterances are independent: no
    collective), one writer that emits the vectors in input order.  Batches are cut exactly as the
    single-GPU stream cuts them and each is embedded by one GPU, so the output file is byte-identical to the single-GPU
    CLI's.  make_trainer(device) -> a built, loaded Trainer (or pass `trainers`, one per device, to reuse contexts across
    calls; they are then left open).  Returns (written, skipped)."""
    import queue
    import threading
    from . import kaldi_io
    from .engine import PinnedArray
    threads = threads or default_reader_threads()
    n_dev = len(devices)
    # two batches per GPU (one running, one queued) + the one being read; XV_EXTRACT_RING overrides (measurement)
    n_buf = int(os.environ.get("XV_EXTRACT_RING", 0)) or 2 * n_dev + 1
    pins = _pinned_ring(n_buf, max_rows, dim)
    free_bufs = queue.Queue()
    for i in range(n_buf):
        free_bufs.put(i)
    work = queue.Queue(maxsize=n_buf)
    done = {}
    done_cv = threading.Condition()
    errors = []
    n_batches = [None]

    def embed(trainer, feats, lengths, keys):
        if lengths.min() >= min_chunk_size and lengths.max() <= chunk_size:
            emb = np.asarray(trainer.predict_packed(feats, lengths), np.float32)
            if normalize:
                emb = emb / np.sqrt(np.sum(np.square(emb), axis=1, keepdims=True))
            return keys, emb, 0
        emb, keep = extract_packed(trainer, feats, lengths, chunk_size, min_chunk_size, normalize)
        return [keys[i] for i in keep], emb, len(keys) - len(keep)

    def worker(slot):
        trainer = None
        try:
            trainer = trainers[slot] if trainers is not None else make_trainer(devices[slot])
            while True:
                item = work.get()
                if item is None:
                    break
                seq, buf, keys, rows, lengths = item
                res = embed(trainer, pins[buf].array[:rows], lengths, keys)
                free_bufs.put(buf)
                with done_cv:
                    done[seq] = res
                    done_cv.notify_all()
        except BaseException as e:                  # surface in the main thread
            with done_cv:
                errors.append(e)
                done_cv.notify_all()
        finally:
            if trainer is not None and trainers is None:
                trainer.close()

    # Both specifiers are opened HERE, before any worker thread exists: "ark:cmd |" and "ark:| cmd" start a child process, and
    # that must not happen while another thread is inside CUDA context creation (make_trainer in the workers).
    first_buf = free_bufs.get()
    rd = kaldi_io.BatchReader(rspecifier, dim, max_rows=max_rows, max_utts=max_utts, buffer=pins[first_buf].array, threads=threads,
                              batch_rows=batch_rows)
    try:
        fp_out = kaldi_io.open_or_fd(wspecifier, "wb")
    except BaseException:
        rd.close()
        raise

    def reader():
        try:
            buf = first_buf
            seq = 0
            for keys, feats, lengths in rd:
                work.put((seq, buf, keys, int(feats.shape[0]), lengths))
                seq += 1
                buf = free_bufs.get()               # the generator reads self.buffer when it resumes
                rd.buffer = pins[buf].array
            n_batches[0] = seq
        except BaseException as e:
            with done_cv:
                errors.append(e)
        finally:
            for _ in devices:
                work.put(None)
            with done_cv:
                if n_batches[0] is None:
                    n_batches[0] = -1
                done_cv.notify_all()

    workers = [threading.Thread(target=worker, args=(i,), daemon=True) for i in range(n_dev)]
    for t in workers:
        t.start()
    rt = threading.Thread(target=reader, daemon=True)
    rt.start()
    written = skipped = 0
    try:
        with fp_out:
            seq = 0
            while True:
                with done_cv:
                    while seq not in done and not errors and not (n_batches[0] is not None and seq >= n_batches[0]):
                        done_cv.wait(0.5)
                    if errors:
                        raise errors[0]
                    if seq not in done:
                        break
                    keys, emb, sk = done.pop(seq)
                if len(keys):
                    kaldi_io.write_vecs_flt(fp_out, keys, emb)
                written += len(keys)
                skipped += sk
                seq += 1
    finally:
        rt.join(timeout=30)
        for t in workers:
            t.join(timeout=30)
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
    parser.add_argument("--gpus", type=int, default=1, help="extract on this many GPUs of the box (devices gpu .. gpu+gpus-1): "
                        "one reader, one host thread + context per GPU, one writer; replaces the run.pl JOB=1:nj fan-out")
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
    if args.rspecifier.rsplit(".", 1)[-1] == "scp":
        sys.exit("The rspecifier must be ark or input pipe")

    def make_trainer(device):
        p = Params(config_json)
        if len(args.node) != 0:
            p.embedding_node = args.node
        t = Trainer(p, args.model_dir, dim, single_cpu=True, precision=args.precision, device=device)
        t.build("predict")
        if not t._has_model():
            sys.exit("Cannot find model in %s" % t.model)
        t.load()
        return t

    # one pipeline for any number of GPUs: the reader runs ahead of the device(s) through a ring of pinned buffers
    first = max(args.gpu, 0)
    written, skipped = extract_stream_mgpu(make_trainer, list(range(first, first + max(args.gpus, 1))), args.rspecifier,
                                           args.wspecifier, dim, args.chunk_size, args.min_chunk_size, args.normalize,
                                           batch_rows=CLI_BATCH_ROWS)
    release_pinned_rings()
    print("[INFO] wrote %d embeddings, skipped %d short utterances" % (written, skipped))


if __name__ == "__main__":
    main()
