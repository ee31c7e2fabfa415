"""GPU: the Python drop-ins keep the reference's signatures and return the oracle's arrays."""
import numpy as np
import pytest

from util import TOL, max_rel, tensor_rel

pytestmark = pytest.mark.gpu


class P(object):
    def __init__(self, **kw):
        self.__dict__.update(kw)

    @property
    def dict(self):
        return self.__dict__


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_tdnn_function_signature_and_endpoints(precision):
    from oracle import numpy_ref, synth
    from xvector_b200.model import variables as V
    from xvector_b200.model.tdnn import tdnn
    cfg = dict(synth.TDNN_STAT_CONFIG)
    w = synth.make_tdnn_weights(30, cfg, seed=2)
    x = synth.config1_features(3, 60, 30, seed=9)
    params = P(**cfg)
    with V.variable_store(w, precision=precision):
        out, endpoints = tdnn(x, params, is_training=False, reuse_variables=None, aux_features=None)
        ref_out, ref_ep = numpy_ref.tdnn(x, numpy_ref.OracleParams(**cfg), w)
        assert list(endpoints.keys()) == list(ref_ep.keys())
        assert params.num_nodes_pooling_layer == 1500 and params.last_layer_no_bn is False
        assert out.shape == ref_out.shape and max_rel(out, ref_out) <= TOL[precision]
        for node in ("tdnn1_conv", "tdnn2_relu", "tdnn3_bn", "tdnn3_relu", "tdnn5_bn", "pooling", "tdnn6_dense"):
            assert endpoints[node].shape == ref_ep[node].shape
            assert tensor_rel(endpoints[node], ref_ep[node]) <= TOL[precision], node
        # the reference's shapes (model/tdnn.py:38-96): conv2d endpoints keep the expanded axis, tdnn3_relu is squeezed
        assert endpoints["tdnn1_conv"].shape == (3, 1, 56, 512) and endpoints["tdnn2_relu"].shape == (3, 1, 52, 512)
        assert endpoints["tdnn3_bn"].shape == (3, 1, 46, 512) and endpoints["tdnn3_relu"].shape == (3, 46, 512)
    V.clear_engines()


def test_pooling_functions_on_arrays():
    from oracle import numpy_ref, synth
    from xvector_b200.model import variables as V
    from xvector_b200.model.pooling import general_pooling, self_attention, statistics_pooling
    rng = np.random.default_rng(4)
    x = np.maximum(rng.standard_normal((4, 120, 1500)), 0).astype(np.float32)
    params = P(pooling_type="statistics_pooling")
    ep = {}
    got = general_pooling(x, None, ep, params, False)
    assert got.shape == (4, 3000) and ep["pooling"] is got
    assert tensor_rel(got, numpy_ref.statistics_pooling(x)) < 1e-5
    assert tensor_rel(statistics_pooling(x, None, {}, params, False), numpy_ref.statistics_pooling(x)) < 1e-5
    with pytest.raises(NotImplementedError, match="Not implement lde pooling"):
        general_pooling(x, None, {}, P(pooling_type="lde"), False)
    # self_attention reads key/value from endpoints by name
    cfg = dict(synth.TDNN_ATT_CONFIG, att_num_heads=4, att_key_num_nodes=[64, 48], att_key_input="k_in", att_value_input="v_in")
    w = {}
    r2 = np.random.default_rng(8)
    for name, n_in, n_out in (("att_key0", 32, 64), ("att_key1", 64, 48)):
        base = "tdnn/attention/%s/%s" % (name, name)
        w[base + "_dense/kernel"] = (r2.standard_normal((n_in, n_out)) / 6).astype(np.float32)
        w[base + "_dense/bias"] = (0.1 * r2.standard_normal(n_out)).astype(np.float32)
    for k, v in (("gamma", r2.uniform(0.5, 1.5, 64)), ("beta", 0.1 * r2.standard_normal(64)),
                 ("moving_mean", 0.1 * r2.standard_normal(64)), ("moving_variance", r2.uniform(0.5, 1.5, 64))):
        w["tdnn/attention/att_key0/att_key0_bn/" + k] = v.astype(np.float32)
    w["tdnn/attention/query"] = (0.3 * r2.standard_normal((4, 12))).astype(np.float32)
    ep = {"k_in": r2.standard_normal((2, 50, 32)).astype(np.float32), "v_in": np.abs(r2.standard_normal((2, 50, 40))).astype(np.float32)}
    ref_ep = dict(ep)
    ref = numpy_ref.self_attention(None, None, ref_ep, numpy_ref.OracleParams(**cfg), w, scope="tdnn")
    with V.variable_store(w):
        with V.variable_scope("tdnn"):
            got = self_attention(None, None, ep, P(**cfg), False)
    assert tensor_rel(got, ref) < 1e-4
    assert tensor_rel(ep["attention_weights"], ref_ep["attention_weights"]) < 1e-4
    assert tensor_rel(ep["att_key0_relu"], ref_ep["att_key0_relu"]) < 1e-4
    V.clear_engines()


def test_common_helpers():
    from oracle import numpy_ref
    from xvector_b200.model import variables as V
    from xvector_b200.model.common import combine_last_two_dimensions, l2_scaling, prelu, shape_list, split_heads
    rng = np.random.default_rng(3)
    x = rng.standard_normal((5, 33)).astype(np.float32)
    assert tensor_rel(l2_scaling(x, 4.0), numpy_ref.l2_scaling(x, 4.0)) < 1e-6
    alpha = rng.uniform(0, 0.3, 33).astype(np.float32)
    with V.variable_store({"s/act/alpha": alpha}):
        with V.variable_scope("s"):
            got = prelu(x, name="act")
    np.testing.assert_allclose(got, np.maximum(x, 0) + alpha * (x - np.abs(x)) * 0.5, rtol=1e-6, atol=1e-7)
    y = rng.standard_normal((2, 7, 12))
    assert split_heads(y, 3).shape == (2, 3, 7, 4) and shape_list(y) == [2, 7, 12]
    np.testing.assert_array_equal(combine_last_two_dimensions(split_heads(y, 3).transpose(0, 2, 1, 3)), y)
    V.clear_engines()


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_trainer_predict_and_extractor(tmp_path, precision):
    """Trainer.predict (model/trainer.py:886-913) and the extractor's chunk/average logic (extract.py:64-93)."""
    import json
    import os
    from oracle import numpy_ref, synth
    from xvector_b200.extract import extract_many, extract_one
    from xvector_b200.misc.utils import Params
    from xvector_b200.model.trainer import Trainer
    cfg = dict(synth.TDNN_STAT_CONFIG)
    w = synth.make_tdnn_weights(30, cfg, seed=6)
    nnet = tmp_path / "nnet"
    nnet.mkdir()
    (nnet / "config.json").write_text(json.dumps(cfg))
    np.savez(str(nnet / "weights.npz"), **w)
    (nnet / "checkpoint").write_text('model_checkpoint_path: "model-1230000"\n')
    params = Params(str(nnet / "config.json"))
    trainer = Trainer(params, str(tmp_path), 30, single_cpu=True, precision=precision)
    trainer.build("predict")
    assert trainer.load() == 1230000
    x = synth.config1_features(2, 80, 34, seed=12)               # wider than dim: sliced (trainer.py:906-907)
    oparams = numpy_ref.OracleParams(**cfg)
    tol = TOL[precision]
    e3 = trainer.predict(x)
    e2 = trainer.predict(x[0])
    assert e3.shape == (2, 512) and e2.shape == (512,) and e3.dtype == np.float32
    assert max_rel(e3, numpy_ref.predict(x, oparams, w, 30)) <= tol
    assert max_rel(e2[None], e3[:1]) < 2e-3
    trainer.set_embedding("tdnn4_relu")
    f = trainer.predict(x[0])
    assert f.shape == (66, 512) and tensor_rel(f, numpy_ref.predict(x[0], oparams, w, 30, node="tdnn4_relu")) <= tol
    # conv2d endpoints keep the reference's expanded axis (model/tdnn.py:38-91): [1, T', 512] for a rank-2 input
    trainer.set_embedding("tdnn1_relu")
    f = trainer.predict(x[0])
    ref = numpy_ref.predict(x[0], oparams, w, 30, node="tdnn1_relu")
    assert f.shape == ref.shape == (1, 76, 512) and tensor_rel(f, ref) <= tol
    f = trainer.predict(x)
    assert f.shape == (2, 1, 76, 512)
    trainer.set_embedding("tdnn6_dense")
    # extractor: skip / whole / chunked, with and without normalisation
    lens = [20, 120, 530, 260]
    feats = [synth.packed_features(np.array([n]), 30, seed=40 + n) for n in lens]
    pred = lambda a: numpy_ref.predict(a, oparams, w, 30)
    for norm in (False, True):
        want = [numpy_ref.extract_embedding(ft, pred, chunk_size=200, min_chunk_size=25, normalize=norm) for ft in feats]
        many = extract_many(trainer, feats, chunk_size=200, min_chunk_size=25, normalize=norm)
        one = [extract_one(trainer, ft, chunk_size=200, min_chunk_size=25, normalize=norm) for ft in feats]
        assert want[0] is None and many[0] is None and one[0] is None
        for i in (1, 2, 3):
            assert max_rel(many[i][None], want[i][None]) <= tol
            assert max_rel(one[i][None], want[i][None]) <= tol
    trainer.close()
    t2 = Trainer(params, str(tmp_path / "missing"), 30, precision=precision)
    t2.build("predict")
    with pytest.raises(SystemExit):
        t2.predict(x)                                            # "Cannot find model in ..." (trainer.py:891-895)
    t2.close()


def test_extract_cli_ark_to_ark(tmp_path):
    """extract.py's command line (egs/voxceleb/v1/nnet/lib/extract.py:11-93) end to end: compressed-matrix ark in through
    an input pipe, native batch reader -> xv_forward -> native vector writer, embeddings ark out; against the oracle's
    restatement of the extraction loop on the features the reference reader decodes."""
    import json
    from oracle import kaldi_io_ref, numpy_ref, synth
    from xvector_b200 import kaldi_io
    from xvector_b200.extract import main
    cfg = dict(synth.TDNN_STAT_CONFIG)
    w = synth.make_tdnn_weights(30, cfg, seed=16)
    nnet = tmp_path / "nnet"
    nnet.mkdir()
    (nnet / "config.json").write_text(json.dumps(cfg))
    (nnet / "feature_dim").write_text("30\n")
    np.savez(str(nnet / "weights.npz"), **w)
    (nnet / "checkpoint").write_text('model_checkpoint_path: "model-10"\n')
    lens = [20, 120, 530, 260, 77]
    feats = [synth.packed_features(np.array([n]), 30, seed=140 + n) for n in lens]
    keys = ["spk%d-utt%d" % (i % 2, i) for i in range(len(lens))]
    ark = str(tmp_path / "feats.ark")
    with open(ark, "wb") as f:
        for i, (k, m) in enumerate(zip(keys, feats)):
            f.write(kaldi_io_ref.write_cm_bytes(m, k) if i % 2 else kaldi_io_ref.write_mat_bytes(m, k))
    decoded = dict(kaldi_io_ref.parse_ark(open(ark, "rb").read()))          # what the reference reader would hand over
    out = str(tmp_path / "xvector.ark")
    oparams = numpy_ref.OracleParams(**cfg)
    pred = lambda a: numpy_ref.predict(a, oparams, w, 30)
    for norm, spec in ((False, "ark:" + out), (True, "ark:| cat > " + out)):
        argv = ["--precision", "fp32", "-s", "200", "-m", "25"] + (["-n"] if norm else []) + [str(tmp_path), "ark:cat %s |" % ark, spec]
        main(argv)
        got = dict(kaldi_io.read_vec_flt_ark(out))
        assert list(got) == keys[1:]                                         # the 20-frame utterance is skipped
        for k in keys[1:]:
            want = numpy_ref.extract_embedding(decoded[k], pred, chunk_size=200, min_chunk_size=25, normalize=norm)
            assert got[k].dtype == np.float32 and max_rel(got[k][None], want[None]) <= TOL["fp32"], k
    # the zero-copy branch: every utterance within [min, chunk] -> one packed xv_forward per batch
    main(["--precision", "bf16", str(tmp_path), ark, out, "-m", "10"])
    got = dict(kaldi_io.read_vec_flt_ark(out))
    assert list(got) == keys
    for k in keys:
        assert max_rel(got[k][None], pred(decoded[k])[None]) <= TOL["bf16"], k


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_extract_packed_views_match_oracle(precision):
    """The chunk-and-average extractor over overlapping VIEWS of packed rows (xv_forward_views; extract.py:64-93): skip,
    whole, chunked utterances, with and without normalisation, against the oracle's restatement of the loop -- and the
    view path against the sliced-copy path (extract_many) it replaces."""
    from oracle import numpy_ref, synth
    from xvector_b200.extract import extract_many, extract_packed
    from xvector_b200.model.trainer import Trainer
    cfg = dict(synth.TDNN_STAT_CONFIG)
    w = synth.make_tdnn_weights(30, cfg, seed=6)
    trainer = Trainer(P(**cfg), "/nonexistent", 30, single_cpu=True, precision=precision, batch_frames=1024)   # several batches
    trainer.build("predict")
    trainer.set_weights(w)
    lens = np.array([20, 120, 530, 260, 77, 1203, 200, 201], np.int32)
    feats = synth.packed_features(lens, 30, seed=77)
    offs = np.concatenate([[0], np.cumsum(lens)])
    oparams = numpy_ref.OracleParams(**cfg)
    pred = lambda a: numpy_ref.predict(a, oparams, w, 30)
    for norm in (False, True):
        got, kept = extract_packed(trainer, feats, lens, chunk_size=200, min_chunk_size=25, normalize=norm)
        assert list(kept) == [1, 2, 3, 4, 5, 6, 7] and got.dtype == np.float32
        ref_many = extract_many(trainer, [feats[offs[i]:offs[i + 1]] for i in range(len(lens))], chunk_size=200, min_chunk_size=25,
                                normalize=norm)
        for row, u in zip(got, kept):
            want = numpy_ref.extract_embedding(feats[offs[u]:offs[u + 1]], pred, chunk_size=200, min_chunk_size=25, normalize=norm)
            assert max_rel(row[None], want[None]) <= TOL[precision], (u, norm)
            assert max_rel(row[None], ref_many[u][None]) <= 2e-6, (u, norm)            # same kernels, same rows
    # views straight through the C-ABI wrapper: overlapping, gapped, ascending
    row0 = np.array([0, 50, 50, 400, 1500], np.int64)
    ln = np.array([100, 100, 300, 64, 700], np.int32)
    e = trainer.predict_views(feats, row0, ln)
    for i in range(len(ln)):
        assert max_rel(e[i][None], pred(feats[row0[i]:row0[i] + ln[i]])[None]) <= TOL[precision], i
    import xvector_b200
    with pytest.raises(xvector_b200.XVectorError, match="ascending"):
        trainer.predict_views(feats, row0[::-1].copy(), ln)
    with pytest.raises(xvector_b200.XVectorError, match="outside"):
        trainer.predict_views(feats, np.array([len(feats) - 10], np.int64), np.array([50], np.int32))
    trainer.close()


def test_extract_cli_multi_gpu_is_byte_identical(tmp_path):
    """extract.py --gpus N (one reader, one host thread + context per GPU, ordered writer; replaces the run.pl JOB=1:nj
    fan-out of run_extract_embeddings.sh:68-71): same bytes as the single-GPU CLI.  With one GPU in the box the two
    contexts share it (the threading, queueing and ordering are what is under test)."""
    import json
    from oracle import synth
    from xvector_b200 import kaldi_io
    from xvector_b200.extract import extract_stream, extract_stream_mgpu, main
    from xvector_b200.misc.utils import Params
    from xvector_b200.model.trainer import Trainer
    cfg = dict(synth.TDNN_STAT_CONFIG)
    w = synth.make_tdnn_weights(30, cfg, seed=16)
    nnet = tmp_path / "nnet"
    nnet.mkdir()
    (nnet / "config.json").write_text(json.dumps(cfg))
    (nnet / "feature_dim").write_text("30\n")
    np.savez(str(nnet / "weights.npz"), **w)
    rng = np.random.default_rng(5)
    lens = rng.integers(20, 700, 60)
    ark = str(tmp_path / "feats.ark")
    with kaldi_io.open_or_fd(ark, "wb") as fd:
        for i, n in enumerate(lens):
            kaldi_io.write_mat(fd, synth.packed_features(np.array([n]), 30, seed=300 + i), "utt%03d" % i)
    one, two = str(tmp_path / "one.ark"), str(tmp_path / "two.ark")

    def make_trainer(device):
        t = Trainer(Params(str(nnet / "config.json")), str(tmp_path), 30, single_cpu=True, device=device)
        t.build("predict")
        t.load()
        return t

    # the same batch cuts on both sides (an utterance's pooling sums are grouped by the row tiles it falls on, so other cuts
    # may move the last bit)
    t1 = make_trainer(0)
    w1, s1 = extract_stream(t1, ark, one, 30, chunk_size=300, min_chunk_size=25, max_rows=4096, max_utts=16)
    t1.close()
    n_dev = min(2, max(1, len([l for l in __import__("os").popen("nvidia-smi -L 2>/dev/null").read().splitlines() if l.startswith("GPU ")])))
    devices = [0, 1] if n_dev == 2 else [0, 0]
    written, skipped = extract_stream_mgpu(make_trainer, devices, ark, two, 30, chunk_size=300, min_chunk_size=25,
                                           max_rows=4096, max_utts=16)            # small batches: many hand-offs
    assert written == int((lens >= 25).sum()) and skipped == int((lens < 25).sum()) and (w1, s1) == (written, skipped)
    assert open(one, "rb").read() == open(two, "rb").read()
    main(["-s", "300", "-m", "25", "--gpus", str(len(devices)) if n_dev == 2 else "1", str(tmp_path), ark, str(tmp_path / "cli.ark")])
    assert len(dict(kaldi_io.read_vec_flt_ark(str(tmp_path / "cli.ark")))) == written
    a = dict(kaldi_io.read_vec_flt_ark(one))
    b = dict(kaldi_io.read_vec_flt_ark(two))
    assert list(a) == list(b)
    for k in a:
        np.testing.assert_array_equal(a[k], b[k])


def test_extract_frame_and_attention_cli(tmp_path):
    """extract_frame.py (edge-replication padding, consecutive chunks) and extract_attention.py (first chunk only)."""
    import json
    from oracle import kaldi_io_ref, numpy_ref, synth
    from xvector_b200 import kaldi_io
    from xvector_b200.extract_frame import main, main_attention
    cfg = dict(synth.TDNN_ATT_CONFIG)
    w = synth.make_tdnn_weights(30, cfg, seed=26)
    nnet = tmp_path / "nnet"
    nnet.mkdir()
    (nnet / "config.json").write_text(json.dumps(cfg))
    (nnet / "feature_dim").write_text("30\n")
    np.savez(str(nnet / "weights.npz"), **w)
    (nnet / "checkpoint").write_text('model_checkpoint_path: "model-10"\n')
    lens = [20, 90, 230]
    feats = [synth.packed_features(np.array([n]), 30, seed=240 + n) for n in lens]
    keys = ["u%d" % i for i in range(len(lens))]
    ark = str(tmp_path / "feats.ark")
    with open(ark, "wb") as f:
        for k, m in zip(keys, feats):
            f.write(kaldi_io_ref.write_mat_bytes(m, k))
    oparams = numpy_ref.OracleParams(**cfg)
    out = str(tmp_path / "frames.ark")
    main(["--precision", "fp32", "--node", "tdnn4_relu", "-s", "100", str(tmp_path), ark, out])
    got = dict(kaldi_io.read_mat_ark(out))
    assert list(got) == keys[1:]

    def want_frames(x):
        pieces = [x[s:s + 100] for s in range(0, x.shape[0], 100)]
        outs = []
        for p in pieces:
            y = numpy_ref.predict(p, oparams, w, 30, node="tdnn4_relu")
            pad = (p.shape[0] - y.shape[0]) // 2
            outs.append(np.concatenate([np.repeat(y[:1], pad, 0), y, np.repeat(y[-1:], pad, 0)]))
        return np.concatenate(outs)
    for k, x in zip(keys[1:], feats[1:]):
        ref = want_frames(x)
        assert got[k].shape == (x.shape[0], 512) and tensor_rel(got[k], ref) <= TOL["fp32"], k
    main_attention(["--precision", "fp32", "-s", "100", str(tmp_path), ark, out])
    got = dict(kaldi_io.read_mat_ark(out))
    assert list(got) == keys[1:]
    for k, x in zip(keys[1:], feats[1:]):
        ref = numpy_ref.predict(x[:100], oparams, w, 30, node="attention_weights")
        assert got[k].shape == ref.shape and tensor_rel(got[k], ref) <= TOL["fp32"] * 2, k


def test_trainer_loads_a_tf_checkpoint_directly(tmp_path):
    """Trainer.load (model/trainer.py:277-294) on the reference's own storage format: a tf.train.Saver tensor bundle +
    `checkpoint` state file, no weights.npz."""
    import json
    from oracle import numpy_ref, synth, tf_bundle_ref
    from xvector_b200.misc.utils import Params
    from xvector_b200.model.trainer import Trainer
    cfg = dict(synth.TDNN_STAT_CONFIG)
    w = synth.make_tdnn_weights(30, cfg, seed=36)
    nnet = tmp_path / "nnet"
    nnet.mkdir()
    (nnet / "config.json").write_text(json.dumps(cfg))
    saved = dict(w, global_step=np.array(820000, np.int64))
    saved["tdnn/tdnn4_dense/kernel/Momentum"] = np.ones_like(w["tdnn/tdnn4_dense/kernel"])
    tf_bundle_ref.write_checkpoint(str(nnet), 820000, saved)
    trainer = Trainer(Params(str(nnet / "config.json")), str(tmp_path), 30, single_cpu=True, precision="fp32")
    trainer.build("predict")
    x = synth.config1_features(2, 70, 30, seed=3)
    e = trainer.predict(x)                                        # loads lazily like the reference's predict
    assert trainer.load() == 820000
    assert max_rel(e, numpy_ref.predict(x, numpy_ref.OracleParams(**cfg), w, 30)) <= TOL["fp32"]
    trainer.close()
