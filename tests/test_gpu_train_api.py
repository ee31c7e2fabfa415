"""GPU: the Python training boundary -- Trainer.build("train") / train_step / save and TrainerMGPU with the reference's
constructor and build signatures (model/trainer.py:87, 309, 590-673, 1107-1134) over xv_train_init / xv_train_step /
xv_comm_init."""
import json
import os

import numpy as np
import pytest

from util import TOL, max_rel

pytestmark = pytest.mark.gpu


def _n_gpus():
    try:
        out = os.popen("nvidia-smi -L 2>/dev/null").read()
        return sum(1 for l in out.splitlines() if l.startswith("GPU "))
    except Exception:
        return 0


def _params(tmp_path, cfg):
    from xvector_b200.misc.utils import Params
    nnet = tmp_path / "nnet"
    nnet.mkdir(exist_ok=True)
    (nnet / "config.json").write_text(json.dumps(cfg))
    return Params(str(nnet / "config.json"))


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_trainer_build_train_step_save_and_predict(tmp_path, precision):
    from oracle import numpy_ref, synth, train_ref
    from xvector_b200.model.trainer import Trainer
    cfg = dict(synth.TDNN_AMSOFTMAX_CONFIG, keep_checkpoint_max=2)
    dim, speakers, batch, frames, lr = 24, 101, 8, 50, 0.05
    w = synth.add_softmax_weights(synth.make_tdnn_weights(dim, cfg, seed=31), speakers)
    x, y = synth.config5_batch(batch, frames, dim + 3, speakers, seed=32, speaker_signal=1.0)     # wider than dim: sliced
    trainer = Trainer(_params(tmp_path, cfg), str(tmp_path), dim, num_speakers=speakers, precision=precision)
    trainer.set_weights(w)
    trainer.build("train")
    trainer.build("valid")                                     # accepted after "train" (train.py:105-106)
    loss, total = trainer.train_step(x, y, lr)
    new_ref, info = train_ref.train_step(x[:, :, :dim], y, w, numpy_ref.OracleParams(**cfg), lr, 0,
                                         bf16_gemm=(precision == "bf16"))
    tol = 1e-5 if precision == "fp32" else 2e-3
    assert abs(loss - info["loss"]) <= tol * abs(info["loss"]) and abs(total - info["total"]) <= tol * abs(info["total"])
    assert trainer.curr_step == 1
    got = trainer.variables()
    assert set(got) == set(new_ref) - {k for k in new_ref if k not in w} | set(w)
    k = "tdnn/tdnn6_dense/kernel"
    assert got[k].shape == w[k].shape
    assert np.linalg.norm(got[k] - new_ref[k].reshape(got[k].shape)) <= (1e-4 if precision == "fp32" else 2e-2) * np.linalg.norm(new_ref[k])
    # the embedding path sees the UPDATED variables (one graph shares them in the reference)
    feats = synth.config1_features(2, 60, dim, seed=5)
    emb = trainer.predict(feats)
    upd = {k2: np.asarray(v).reshape(np.asarray(w[k2]).shape) for k2, v in got.items()}
    ref = numpy_ref.predict(feats, numpy_ref.OracleParams(**cfg), upd, dim)
    assert emb.shape == (2, 512) and max_rel(emb, ref) <= TOL[precision]
    # checkpoints: the reference's on-disk format, restored by a fresh Trainer
    trainer.train_step(x, y, lr)
    prefix = trainer.save(trainer.curr_step)
    assert os.path.isfile(prefix + ".index") and os.path.isfile(prefix + ".data-00000-of-00001")
    after = trainer.variables()
    trainer.close()
    t2 = Trainer(_params(tmp_path, cfg), str(tmp_path), dim, num_speakers=speakers, precision=precision)
    t2.build("predict")
    assert t2.load() == 2
    emb2 = t2.predict(feats)
    ref2 = numpy_ref.predict(feats, numpy_ref.OracleParams(**cfg), {k2: np.asarray(v).reshape(np.asarray(w[k2]).shape) for k2, v in after.items()}, dim)
    assert max_rel(emb2, ref2) <= TOL[precision]
    t2.close()
    # resuming: build("train") without injected weights picks the checkpoint up (model/trainer.py:606-607)
    t3 = Trainer(_params(tmp_path, cfg), str(tmp_path), dim, num_speakers=speakers, precision=precision)
    t3.build("train")
    assert t3.curr_step == 2
    assert np.array_equal(t3.variables()[k], after[k])
    t3.close()


def test_trainer_rejects_what_the_training_step_does_not_cover(tmp_path):
    from oracle import synth
    from xvector_b200.model.trainer import Trainer
    for cfg, msg in ((dict(synth.TDNN_AMSOFTMAX_CONFIG, loss_func="asoftmax"), "Not implement asoftmax loss"),
                     (dict(synth.TDNN_ATT_CONFIG), "self_attention pooling"),
                     (dict(synth.RESNET_CONFIG), "resnet_18 network")):
        t = Trainer(_params(tmp_path, cfg), str(tmp_path), 40 if cfg["network_type"] == "resnet_18" else 24, num_speakers=10)
        assert t.trainable() is not None
        with pytest.raises(NotImplementedError, match=msg):
            t.build("train")
        t.build("predict")                                      # extraction is served for all of them
        t.close()
    t = Trainer(_params(tmp_path, dict(synth.TDNN_AMSOFTMAX_CONFIG)), str(tmp_path), 24, num_speakers=10)
    with pytest.raises(NotImplementedError, match="noupdate_var_list"):
        t.build("train", noupdate_var_list=["tdnn1"])
    t.close()


def test_trainer_train_epoch_over_a_kaldi_directory(tmp_path):
    """Trainer.train (model/trainer.py:590-673): an epoch of steps from feats.scp / spk2utt / spklist, checkpoint at the end."""
    from oracle import synth
    from xvector_b200 import kaldi_io
    from xvector_b200.model.trainer import Trainer
    rng = np.random.default_rng(0)
    data = tmp_path / "data"
    data.mkdir()
    dim, n_spk = 24, 12
    with kaldi_io.open_or_fd("ark,scp:%s,%s" % (data / "feats.ark", data / "feats.scp"), "wb") as fd:
        for s in range(n_spk):
            for u in range(2):
                m = (rng.standard_normal((int(rng.integers(90, 140)), dim)) + 0.5 * np.sin(s + np.arange(dim))).astype(np.float32)
                kaldi_io.write_mat(fd, m, "s%02d-u%d" % (s, u))
    (data / "spk2utt").write_text("".join("s%02d s%02d-u0 s%02d-u1\n" % (s, s, s) for s in range(n_spk)))
    (tmp_path / "spklist").write_text("".join("s%02d %d\n" % (s, s) for s in range(n_spk)))
    cfg = dict(synth.TDNN_AMSOFTMAX_CONFIG, num_speakers_per_batch=8, num_segments_per_speaker=1, min_segment_len=40,
               max_segment_len=60, num_steps_per_epoch=12, save_checkpoints_steps=5, show_training_progress=100, seed=1,
               loss_func="softmax", last_layer_linear=False)
    trainer = Trainer(_params(tmp_path, cfg), str(tmp_path), dim, num_speakers=n_spk, precision="bf16")
    trainer.build("train")                                     # fresh variables (tf.global_variables_initializer)
    first = trainer.train_step(*__import__("xvector_b200.data_loader", fromlist=["x"]).RandomSegmentLoader(
        str(data), str(tmp_path / "spklist"), num_speakers=8, num_segments=1, min_len=40, max_len=60, seed=9).fetch(), 0.05)
    last = trainer.train(str(data), str(tmp_path / "spklist"), 0.05)
    assert trainer.curr_step == 12 and np.isfinite(last[0]) and last[0] < first[0]
    assert abs(first[0] - np.log(n_spk)) < 1.5           # fresh variables: cross entropy in the neighbourhood of ln(S)
    state = __import__("xvector_b200.tf_checkpoint", fromlist=["x"]).get_checkpoint_state(str(tmp_path / "nnet"))
    assert state[0] == "model-12" and "model-5" in state[1] and "model-10" in state[1]
    trainer.close()


def test_trainer_valid_matches_oracle(tmp_path):
    """Trainer.valid (model/trainer.py:756-884) after two training steps: the eval-mode network output of the CURRENT
    variables from the device, the loss head on the host -- against the oracle's eval-mode forward of the same variables
    and the oracle's loss, batch by batch in the sequential loader's order."""
    from oracle import numpy_ref, synth, train_ref
    from xvector_b200 import kaldi_io
    from xvector_b200.data_loader import DataOutOfRange, SequentialSegmentLoader
    from xvector_b200.model.trainer import Trainer
    rng = np.random.default_rng(1)
    data = tmp_path / "data"
    data.mkdir()
    dim, n_spk = 24, 6
    with kaldi_io.open_or_fd("ark,scp:%s,%s" % (data / "feats.ark", data / "feats.scp"), "wb") as fd:
        for s in range(n_spk):
            for u in range(3):
                m = (rng.standard_normal((int(rng.integers(60, 100)), dim)) + 0.5 * np.sin(s + np.arange(dim))).astype(np.float32)
                kaldi_io.write_mat(fd, m, "s%02d-u%d" % (s, u))
    (data / "spk2utt").write_text("".join("s%02d s%02d-u0 s%02d-u1 s%02d-u2\n" % (s, s, s, s) for s in range(n_spk)))
    spklist = tmp_path / "spklist"
    spklist.write_text("".join("s%02d %d\n" % (s, s) for s in range(n_spk)))
    cfg = dict(synth.TDNN_AMSOFTMAX_CONFIG, num_speakers_per_batch=4, num_segments_per_speaker=2, min_segment_len=40,
               max_segment_len=60, valid_max_iterations=100, seed=2)
    w = synth.add_softmax_weights(synth.make_tdnn_weights(dim, cfg, seed=33), n_spk)
    trainer = Trainer(_params(tmp_path, cfg), str(tmp_path), dim, num_speakers=n_spk, precision="fp32")
    trainer.set_weights(w)
    trainer.build("train")
    trainer.build("valid")
    x, y = synth.config5_batch(8, 50, dim, n_spk, seed=34, speaker_signal=1.0)
    for _ in range(2):
        trainer.train_step(x, y, 0.05)
    loss, emb, labels = trainer.valid(str(data), str(spklist), output_embeddings=True)
    got = trainer.variables()
    upd = {k: np.asarray(v).reshape(np.asarray(w[k]).shape) for k, v in got.items()}
    oparams = numpy_ref.OracleParams(**cfg)
    ld = SequentialSegmentLoader(str(data), str(spklist), batch_size=8, min_len=50, max_len=50, dim=dim)
    ref_losses, ref_emb, ref_labels = [], [], []
    while True:
        try:
            f, lab = ld.fetch()
        except DataOutOfRange:
            break
        out = numpy_ref.predict(f, oparams, upd, dim, node="output")
        ref_losses.append(train_ref.loss_forward_backward(out, lab, upd, oparams, trainer.curr_step, want_grad=False)[0])
        ref_emb.append(numpy_ref.predict(f, oparams, upd, dim))
        ref_labels.append(lab)
    assert len(ref_losses) == 18 // 8 + 1
    assert abs(loss - np.mean(ref_losses)) <= 1e-4 * abs(np.mean(ref_losses)), (loss, ref_losses)
    assert emb.shape == (24, 512) and np.array_equal(labels, np.concatenate(ref_labels))
    assert max_rel(emb, np.concatenate(ref_emb)) <= TOL["fp32"]
    trainer.close()


@pytest.mark.skipif(_n_gpus() < 2, reason="needs 2 GPUs (run with gpurun --gpus 2)")
def test_trainer_mgpu_matches_two_tower_oracle(tmp_path):
    from oracle import numpy_ref, synth, train_ref
    from xvector_b200.model.trainer import TrainerMGPU
    cfg = dict(synth.TDNN_AMSOFTMAX_CONFIG, num_speakers_per_batch=8, num_segments_per_speaker=1)
    dim, speakers = 24, 101
    w = synth.add_softmax_weights(synth.make_tdnn_weights(dim, cfg, seed=31), speakers)
    x, y = synth.config5_batch(8, 50, dim, speakers, seed=32)
    trainer = TrainerMGPU(_params(tmp_path, cfg), str(tmp_path), dim, num_speakers=speakers, num_gpus=2, precision="fp32")
    trainer.set_weights(w)
    trainer.build("train")
    assert trainer.engine.comm_info() == {"rank": 0, "world": 2, "grad_elems": trainer.engine.comm_info()["grad_elems"], "buckets": 3,
                                          "dtype": "fp32"}
    loss, total = trainer.train_step(x, y, 0.05)
    new_ref, info = train_ref.train_step(x, y, w, numpy_ref.OracleParams(**cfg), 0.05, 0, num_towers=2)
    assert abs(loss - info["loss"]) <= 1e-5 * abs(info["loss"])
    v0, v1 = trainer.engines[0].train_variables(), trainer.engines[1].train_variables()
    for k in ("tdnn/tdnn3_conv/kernel", "tdnn/tdnn6_dense/kernel", "softmax/output/kernel", "tdnn/tdnn1_bn/moving_mean"):
        np.testing.assert_array_equal(v0[k], v1[k])            # replicas stay bit-identical
        assert np.linalg.norm(v0[k] - new_ref[k].reshape(v0[k].shape)) <= 5e-3 * np.linalg.norm(new_ref[k]), k
    trainer.close()


@pytest.mark.skipif(_n_gpus() < 2, reason="needs 2 GPUs (run with gpurun --gpus 2)")
def test_trainer_mgpu_bf16_exchange_tracks_fp32_exchange(tmp_path):
    """bf16 precision, two towers: gradients cross NVLink as bf16 by default (fp32 masters; "grad_allreduce_dtype"), the
    replicas stay bit-identical, and the step stays within bf16 rounding of the same step with an fp32 exchange."""
    from oracle import synth
    from xvector_b200.model.trainer import TrainerMGPU
    dim, speakers = 24, 101
    x, y = synth.config5_batch(16, 50, dim, speakers, seed=32, speaker_signal=1.0)
    out = {}
    for comm in ("bf16", "fp32"):
        cfg = dict(synth.TDNN_AMSOFTMAX_CONFIG, num_speakers_per_batch=16, num_segments_per_speaker=1, grad_allreduce_dtype=comm)
        w = synth.add_softmax_weights(synth.make_tdnn_weights(dim, cfg, seed=31), speakers)
        trainer = TrainerMGPU(_params(tmp_path, cfg), str(tmp_path), dim, num_speakers=speakers, num_gpus=2, precision="bf16")
        trainer.set_weights(w)
        trainer.build("train")
        assert trainer.engine.comm_info()["dtype"] == comm
        loss, _ = trainer.train_step(x, y, 0.05)
        v0, v1 = trainer.engines[0].train_variables(), trainer.engines[1].train_variables()
        for k in v0:
            np.testing.assert_array_equal(v0[k], v1[k])        # replicas stay bit-identical
        out[comm] = (loss, {k: trainer.engine.train_grad(k, np.asarray(w[k]).shape) for k in ("tdnn/tdnn3_conv/kernel", "tdnn/tdnn6_dense/kernel", "softmax/output/kernel")},
                     v0["tdnn/tdnn1_bn/moving_mean"])
        trainer.close()
    assert abs(out["bf16"][0] - out["fp32"][0]) <= 1e-6 * abs(out["fp32"][0])      # the loss travels in the fp32 tail
    np.testing.assert_array_equal(out["bf16"][2], out["fp32"][2])                  # so do the moving statistics
    for k, g in out["bf16"][1].items():
        ref = out["fp32"][1][k]
        assert np.linalg.norm(g - ref) <= 6e-3 * np.linalg.norm(ref), k            # one bf16 rounding per tower gradient + one of the sum


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_pipelined_steps_match_synchronous_steps(precision):
    """xv_train_submit / xv_train_collect (batch k+1 copied while step k runs, loss collected one step late) against
    xv_train_step on the same batches: same losses, same weights; also a step on device-resident buffers, and the state checks."""
    import xvector_b200
    from oracle import synth
    from xvector_b200 import XVectorError
    cfg = dict(synth.TDNN_AMSOFTMAX_CONFIG)
    dim, speakers, lr, n_steps = 24, 101, 0.01, 6
    w = synth.add_softmax_weights(synth.make_tdnn_weights(dim, cfg, seed=41), speakers)
    # batches of DIFFERENT shapes: the loader draws one segment length per batch (dataset/data_loader.py:322-349)
    batches = [synth.config5_batch(16 + 4 * (i % 2), 50 + 7 * i, dim, speakers, seed=60 + i, speaker_signal=1.0) for i in range(n_steps)]

    def engine():
        e = xvector_b200.Engine(cfg, dim, precision=precision)
        e.set_weights(w)
        e.train_init(speakers)
        return e

    a = engine()
    sync_losses = [a.train_step(x, y, lr, i) for i, (x, y) in enumerate(batches)]
    a2 = engine()                                               # the run-to-run spread of the synchronous call itself
    again = [a2.train_step(x, y, lr, i) for i, (x, y) in enumerate(batches)]
    a2.close()
    b = engine()
    with pytest.raises(XVectorError):
        b.train_collect()                                       # nothing submitted
    pipe_losses = []
    b.train_submit(batches[0][0], batches[0][1], lr, 0)
    for i in range(1, n_steps):
        b.train_submit(batches[i][0], batches[i][1], lr, i)
        if i == 1:
            with pytest.raises(XVectorError):                   # two steps in flight: a third submit is refused ...
                b.train_submit(batches[i][0], batches[i][1], lr, i)
            with pytest.raises(XVectorError):                   # ... and so is the synchronous call
                b.train_step(batches[i][0], batches[i][1], lr, i)
        pipe_losses.append(b.train_collect())
    pipe_losses.append(b.train_collect())
    # Reductions use fp32 atomics, so two runs of the same step agree to rounding only, and training amplifies that from step
    # to step: measured run-to-run spread of the SYNCHRONOUS call on these batches (printed below as the second run) is 2e-7 /
    # 1.5e-4 (fp32 / bf16) at step 1 and 1.4e-3 / 1e-2 by step 5.  Step 1 is the first one submitted while another is in
    # flight: a batch that went to the wrong step, or a copy that raced a running step, shows there at full precision.
    report = "sync %s\nsync %s (second run)\npipe %s" % tuple(" ".join("%.6f" % l[0] for l in ls) for ls in (sync_losses, again, pipe_losses))
    print(report)
    early = 1e-5 if precision == "fp32" else 1e-3
    late = 1e-2 if precision == "fp32" else 5e-2
    for i, ((l0, t0), (l1, t1)) in enumerate(zip(sync_losses, pipe_losses)):
        tol = 1e-5 if i == 0 else (early if i == 1 else late)
        assert abs(l0 - l1) <= tol * abs(l0) and abs(t0 - t1) <= tol * abs(t0), report
    k = "tdnn/tdnn3_conv/kernel"
    wa, wb = a.train_weight(k, w[k].shape), b.train_weight(k, w[k].shape)
    assert np.linalg.norm(wa - wb) <= 0.5 * np.linalg.norm(wa - w[k]), report
    tol = late
    # a step on buffers that already live on the device (loss_out = NULL: enqueued only), then a synchronous one
    x, y = batches[0]
    d_x, d_y = b.device_alloc(x.nbytes), b.device_alloc(y.nbytes)
    try:
        b.h2d(d_x, x)
        b.h2d(d_y, np.ascontiguousarray(y, np.int32))
        a_loss = a.train_step(x, y, lr, n_steps)
        b.train_step_resident(d_x, d_y, x.shape[0], x.shape[1], lr, n_steps)
        b.sync()
        nxt_a, nxt_b = a.train_step(x, y, lr, n_steps + 1), b.train_step(x, y, lr, n_steps + 1)
        # third and fourth pass over the same 16 utterances: the loss has collapsed to ~0.1 and is chaotic in relative terms
        # (0.080 vs 0.083 between two runs); what is checked is that the resident step trained at all and stayed on track
        assert np.isfinite(a_loss).all() and np.isfinite(nxt_b).all()
        assert abs(nxt_a[0] - nxt_b[0]) <= 4 * tol * max(abs(nxt_a[0]), 1.0), (nxt_a, nxt_b)
        assert nxt_b[0] < 0.5 * pipe_losses[0][0]
    finally:
        b.device_free(d_x)
        b.device_free(d_y)
    a.close()
    b.close()
