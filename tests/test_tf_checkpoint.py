"""TF checkpoint reader without TensorFlow, CPU.  parity unpinned (no TensorFlow, no saved model here):
round trips through the independent writer oracle/tf_bundle_ref.py, CRC-32C known answers, checkpoint selection."""
import json
import os

import numpy as np
import pytest


def test_crc32c_known_answers():
    from oracle import tf_bundle_ref
    from xvector_b200 import tf_checkpoint as T
    # RFC 3720 appendix B.4 + the classic check value
    for data, want in ((b"123456789", 0xE3069283), (b"\0" * 32, 0x8A9136AA), (b"\xff" * 32, 0x62A8AB43),
                       (bytes(range(32)), 0x46DD794E), (bytes(range(31, -1, -1)), 0x113FDB5C)):
        assert T.crc32c(data) == want == tf_bundle_ref.crc32c(data)
    rng = np.random.default_rng(0)
    blob = rng.integers(0, 256, 100003).astype(np.uint8).tobytes()
    assert T.crc32c(blob) == tf_bundle_ref.crc32c(blob)
    assert T.crc32c(blob[5000:], T.crc32c(blob[:5000])) == T.crc32c(blob)          # streaming


def test_bundle_round_trip_and_trainer_load(tmp_path):
    from oracle import synth, tf_bundle_ref
    from xvector_b200 import tf_checkpoint as T
    cfg = dict(synth.TDNN_STAT_CONFIG, network_relu_type="prelu", num_nodes_pooling_layer=96)
    w = synth.make_tdnn_weights(30, cfg, seed=5, hidden=64)          # narrow layers: the pure-python CRC of the writer is slow
    saved = dict(w)
    saved["global_step"] = np.array(1230000, np.int64)                               # scalar, int64
    saved["tdnn/tdnn1_conv/kernel/Momentum"] = np.zeros_like(w["tdnn/tdnn1_conv/kernel"])   # optimizer slot: not a weight
    saved["softmax/output/kernel"] = np.random.default_rng(1).standard_normal((512, 37)).astype(np.float32)
    saved["big/matrix"] = np.random.default_rng(2).standard_normal((300, 1000))     # 2.4 MB float64
    nnet = tmp_path / "nnet"
    nnet.mkdir()
    prefix = tf_bundle_ref.write_checkpoint(str(nnet), 1230000, saved, keep=(410000, 820000))
    reader = T.BundleReader(prefix)
    assert set(reader.entries) == set(saved) and len(saved) > 40                     # several data blocks
    for name, ref in saved.items():
        got = reader.get_tensor(name)
        assert got.dtype == ref.dtype and got.shape == ref.shape and np.array_equal(got, ref), name
    assert reader.variables()["tdnn/tdnn5_dense/kernel"] == ((64, 96), np.float32)
    weights, step = T.load_weights(str(nnet))
    assert step == 1230000 and "global_step" not in weights and not any(k.endswith("/Momentum") for k in weights)
    assert set(weights) == set(w) | {"softmax/output/kernel", "big/matrix"}
    T.main([str(nnet)])
    npz = np.load(str(nnet / "weights.npz"))
    assert all(np.array_equal(npz[k], v) for k, v in weights.items())
    # corruption is detected by the checksums TensorFlow would check
    data = prefix + ".data-00000-of-00001"
    blob = bytearray(open(data, "rb").read())
    blob[reader.entries["tdnn/tdnn2_bn/gamma"]["offset"] + 3] ^= 0x40
    open(data, "wb").write(bytes(blob))
    with pytest.raises(IOError, match="data checksum mismatch"):
        T.BundleReader(prefix).get_tensor("tdnn/tdnn2_bn/gamma")
    idx = bytearray(open(prefix + ".index", "rb").read())
    idx[10] ^= 1
    open(prefix + ".index", "wb").write(bytes(idx))
    with pytest.raises(IOError, match="block checksum mismatch"):
        T.BundleReader(prefix)
    with pytest.raises(IOError, match="bad table magic"):
        T.read_table(data)


def test_get_checkpoint_selects_like_the_reference(tmp_path):
    """misc/utils.py:251-304: "last", an explicit step, -1 = the epoch with the lowest validation loss."""
    from xvector_b200 import tf_checkpoint as T
    model = tmp_path / "nnet"
    model.mkdir()
    steps = [1000, 2000, 3000, 4000]
    with open(str(model / "checkpoint"), "w") as f:
        f.write('model_checkpoint_path: "/old/place/nnet/model-4000"\n')
        for s in steps:
            f.write('all_model_checkpoint_paths: "/old/place/nnet/model-%d"\n' % s)
    (model / "config.json").write_text(json.dumps({"num_steps_per_epoch": 1000}))
    (model / "valid_loss").write_text("0 2.5 0.11\n1 1.25 0.08\n2 1.75 0.09\n3 1.5 0.10\n")
    assert T.get_checkpoint_state(str(model)) == ("/old/place/nnet/model-4000", ["/old/place/nnet/model-%d" % s for s in steps])
    assert T.get_checkpoint(str(model), "last") == str(model / "model-4000")
    assert T.get_checkpoint(str(model), "3000") == str(model / "model-3000")
    assert T.get_checkpoint(str(model), -1) == str(model / "model-2000")             # epoch 1 (0-based) is the best
    state = T.get_checkpoint_state(str(model))
    assert state[0] == str(model / "model-2000") and state[1] == [str(model / ("model-%d" % s)) for s in steps]
    assert T.checkpoint_step(state[0]) == 2000
    with pytest.raises(AssertionError, match="not in the model directory"):
        T.get_checkpoint(str(model), "2500")
    with pytest.raises(SystemExit):
        T.get_checkpoint(str(tmp_path / "nowhere"))


def test_product_writer_matches_the_independent_writer(tmp_path):
    """tf_checkpoint.save_checkpoint (what Trainer.save writes, model/trainer.py:296-303) against the independent writer
    of the same published format: identical data shard, identical index ENTRIES (the block split points are a free
    choice of a table builder), optimizer slots named the reference's way ("/optimizer", model/trainer.py:234-243) are
    left out by load_weights, max_to_keep prunes old bundles."""
    from oracle import synth, tf_bundle_ref
    from xvector_b200 import tf_checkpoint as T
    cfg = dict(synth.TDNN_STAT_CONFIG, num_nodes_pooling_layer=96)
    w = synth.make_tdnn_weights(24, cfg, seed=3, hidden=64)
    saved = dict(w)
    saved["tdnn/tdnn1_conv/kernel/optimizer"] = np.ones_like(w["tdnn/tdnn1_conv/kernel"])      # momentum slot
    saved["tdnn/tdnn1_conv/kernel/optimizer_1"] = np.ones_like(w["tdnn/tdnn1_conv/kernel"])    # Adam second moment
    saved["global_step"] = np.array(7, np.int64)
    a, b = tmp_path / "a", tmp_path / "b"
    a.mkdir()
    b.mkdir()
    pa = T.save_checkpoint(str(a), 90000, saved)
    pb = tf_bundle_ref.write_checkpoint(str(b), 90000, saved)
    assert open(pa + ".data-00000-of-00001", "rb").read() == open(pb + ".data-00000-of-00001", "rb").read()
    assert T.read_table(pa + ".index") == T.read_table(pb + ".index")
    got, step = T.load_weights(str(a))
    assert step == 90000 and set(got) == set(w)
    for k in w:
        assert np.array_equal(got[k], w[k])
    # a flipped byte in the data shard is caught by the per-tensor CRC
    raw = bytearray(open(pa + ".data-00000-of-00001", "rb").read())
    raw[1000] ^= 0x40
    open(pa + ".data-00000-of-00001", "wb").write(bytes(raw))
    with pytest.raises(IOError, match="checksum"):
        T.load_weights(str(a))
    # keep_checkpoint_max (model/trainer.py:382)
    c = tmp_path / "c"
    for step in (100, 200, 300):
        T.save_checkpoint(str(c), step, {"v": np.arange(4, dtype=np.float32)}, max_to_keep=2)
    cur, every = T.get_checkpoint_state(str(c))
    assert cur == "model-300" and every == ["model-200", "model-300"]
    assert not os.path.exists(str(c / "model-100.index")) and os.path.exists(str(c / "model-200.index"))
