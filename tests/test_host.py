"""CPU: the C-ABI library loads and exports every declared symbol; host-side logic (Params, scheduler, extractor
planning, error behaviour).  No compute calls: there is no GPU here and no CPU fallback."""
import ctypes as C
import glob
import json
import os
import re

import numpy as np
import pytest

import xvector_b200
from xvector_b200 import _lib
from xvector_b200 import scheduler
from xvector_b200.extract import plan_many
from xvector_b200.misc.utils import Params, ParamsPlain

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    lib = xvector_b200.load()
    header = open(os.path.join(ROOT, "include", "xvector.h")).read()
    declared = set(re.findall(r"\b(xv_[a-z0-9_]+)\s*\(", header)) - {"xv_ctx"}
    assert len(declared) >= 30
    for name in sorted(declared):
        assert hasattr(lib, name), "libxvector.so does not export %s" % name
    assert declared == set(_lib.PROTOTYPES), "ctypes prototypes and include/xvector.h disagree: %s" % (
        declared ^ set(_lib.PROTOTYPES))
    assert b"sm_100a" in lib.xv_version()


NNET_CONF = sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "nnet_conf", "*.json")))


def test_all_reference_nnet_conf_load_through_params_and_the_library_parser():
    """all 42 egs/*/nnet_conf/*.json (fixtures copied by tests/golden/make_nnet_conf_fixtures.py) load
    through Params AND through the library's own JSON reader + plan builder (xv_config_check = xv_create minus the
    device).  The raw file text is used (comment-like "Note" keys, exponent floats, lists), with feature_dim spliced in."""
    from xvector_b200.engine import config_json
    assert len(NNET_CONF) == 42
    lib = xvector_b200.load()
    buf = C.create_string_buffer(512)
    seen = set()
    for path in NNET_CONF:
        params = Params(path)
        assert params.network_type in ("tdnn", "resnet_18") and "loss_func" in params.dict
        dim = 40 if params.network_type == "resnet_18" else 30
        raw = open(path).read()
        raw = raw[:raw.index("{") + 1] + '"feature_dim": %d,' % dim + raw[raw.index("{") + 1:]
        for text in (raw, config_json(params.dict, dim)):
            rc = lib.xv_config_check(text.encode("utf-8"), buf, 512)
            assert rc == 0, "%s: %s" % (os.path.basename(path), lib.xv_last_error(None))
            summary = dict(kv.split("=") for kv in buf.value.decode().split(";"))
            assert summary["embedding_node"] == params.dict.get("embedding_node", "tdnn6_dense")
            assert summary["pooling"] == params.pooling_type
            assert int(summary["pooled_dim"]) == 3000
        seen.add((params.network_type, params.loss_func, params.pooling_type))
    assert ("tdnn", "softmax", "self_attention") in seen and ("resnet_18", "softmax", "statistics_pooling") in seen
    # what the reference rejects, the parser rejects with the reference's words (model/trainer.py:109-110)
    rc = lib.xv_config_check(b'{"network_type": "ftdnn", "pooling_type": "statistics_pooling", "feature_dim": 30}', buf, 512)
    assert rc == _lib.XV_EUNSUPPORTED and b"Not implement ftdnn network" in lib.xv_last_error(None)
    assert lib.xv_config_check(b'{"network_type": "tdnn", "feature_dim": 30', buf, 512) == _lib.XV_EINVAL


def test_create_fails_loudly_without_gpu_or_with_bad_config():
    lib = xvector_b200.load()
    ctx = C.c_void_p()
    rc = lib.xv_create(0, b"{not json", C.byref(ctx))
    assert rc != 0 and not ctx.value
    assert len(lib.xv_last_error(None)) > 0
    assert lib.xv_create(0, None, C.byref(ctx)) == _lib.XV_EINVAL


def test_product_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "tf-kaldi-speaker_b200")
    for path in glob.glob(os.path.join(pkg, "**", "*.py"), recursive=True):
        src = open(path).read()
        assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), "%s imports the oracle" % path


def test_params_interface(tmp_path):
    p = tmp_path / "c.json"
    p.write_text(json.dumps({"network_type": "tdnn", "pooling_type": "statistics_pooling", "weight_l2_regularizer": 1e-2}))
    params = Params(str(p))
    assert params.network_type == "tdnn" and "pooling_type" in params.dict
    params.dict["num_nodes_pooling_layer"] = 1500
    assert params.num_nodes_pooling_layer == 1500
    params.last_layer_no_bn = False
    assert "last_layer_no_bn" in params.dict
    params.save(str(tmp_path / "d.json"))
    assert Params(str(tmp_path / "d.json")).num_nodes_pooling_layer == 1500
    pp = ParamsPlain()
    pp.x = 3
    assert pp.dict == {"x": 3}


def test_config_json_drops_non_serialisable_values():
    cfg = json.loads(xvector_b200.config_json({"network_type": "tdnn", "global_step": object()}, 30, "fp32", 4096))
    assert cfg == {"network_type": "tdnn", "feature_dim": 30, "precision": "fp32", "batch_frames": 4096}


def test_chunk_plan_matches_oracle():
    from oracle import numpy_ref
    for length, chunk in [(25000, 10000), (10001, 10000), (2501, 1001), (12345, 2000), (6000, 2000), (4001, 2000)]:
        assert scheduler.chunk_plan(length, chunk) == numpy_ref.chunk_plan(length, chunk)


def test_plan_many_skips_and_chunks():
    work = plan_many([10, 300, 4500], chunk_size=2000, min_chunk_size=25)
    assert work[0] == (1, 0, 300)
    assert [w for w in work if w[0] == 2] == [(2, s, z) for s, z in scheduler.chunk_plan(4500, 2000)]
    assert all(w[0] != 0 for w in work)


def test_lpt_sharding_balances_flops():
    from oracle import synth
    lengths = synth.config2_lengths(4096)
    for world in (1, 2, 4, 8):
        shards = scheduler.shard_lpt(lengths, world)
        all_idx = np.sort(np.concatenate(shards))
        np.testing.assert_array_equal(all_idx, np.arange(len(lengths)))
        loads = np.array([scheduler.utterance_flops(lengths[s], 30).sum() for s in shards])
        assert loads.max() / loads.mean() < 1.01
    assert abs(scheduler.utterance_flops(300, 30) / 1e9 - 2.4529) < 1e-3


def test_trainer_error_behaviour(tmp_path):
    from xvector_b200.model.trainer import Trainer
    p = ParamsPlain()
    p.network_type = "ftdnn"
    with pytest.raises(NotImplementedError, match="Not implement ftdnn network"):
        Trainer(p, str(tmp_path), 30)
    p.network_type = "tdnn"
    p.loss_func = "nope"
    with pytest.raises(NotImplementedError, match="Not implement nope loss"):
        Trainer(p, str(tmp_path), 30)
    p.loss_func = "softmax"
    t = Trainer(p, str(tmp_path), 30)
    with pytest.raises(SystemExit):
        t.load()                                           # graph not built (model/trainer.py:283-284)
    with pytest.raises(AssertionError):
        t.build("nonsense")


def test_all_reference_style_configs_load(tmp_path):
    from oracle import synth
    for i, cfg in enumerate((synth.TDNN_STAT_CONFIG, synth.TDNN_ATT_CONFIG)):
        f = tmp_path / ("c%d.json" % i)
        f.write_text(json.dumps(cfg))
        assert Params(str(f)).network_type == "tdnn"


def test_extract_packed_views_match_extract_many():
    """Host logic of the view-based extractor (extract.py:64-93 semantics): with a stand-in trainer whose "embedding" is a
    simple function of the rows, extract_packed (overlapping views of packed rows) == extract_many (sliced copies)."""
    from xvector_b200.extract import extract_many, extract_packed

    class Fake(object):
        def _emb(self, a):
            a = np.asarray(a, np.float64)
            return np.concatenate([a.mean(axis=0), a.std(axis=0), [float(len(a))]]).astype(np.float32)

        def predict_ragged(self, pieces):
            return np.stack([self._emb(p) for p in pieces])

        def predict_views(self, packed, row0, lengths):
            assert (np.diff(row0) >= 0).all()                       # the C-ABI wants ascending source rows
            return np.stack([self._emb(packed[r:r + n]) for r, n in zip(row0, lengths)])

    rng = np.random.default_rng(3)
    lengths = np.array([30, 410, 7, 200, 999, 201, 1000], np.int32)
    feats = rng.standard_normal((int(lengths.sum()), 5)).astype(np.float32)
    offs = np.concatenate([[0], np.cumsum(lengths)])
    lst = [feats[offs[i]:offs[i + 1]] for i in range(len(lengths))]
    for norm in (False, True):
        ref = extract_many(Fake(), lst, chunk_size=200, min_chunk_size=25, normalize=norm)
        got, kept = extract_packed(Fake(), feats, lengths, chunk_size=200, min_chunk_size=25, normalize=norm)
        assert list(kept) == [i for i, e in enumerate(ref) if e is not None] == [0, 1, 3, 4, 5, 6]
        for row, i in zip(got, kept):
            np.testing.assert_allclose(row, ref[i], rtol=2e-6, atol=1e-7)
    got, kept = extract_packed(Fake(), feats, lengths, chunk_size=10000, min_chunk_size=2000)
    assert got.shape[0] == 0 and len(kept) == 0


def test_plan_arrays_equals_the_per_utterance_chunk_plan():
    """scheduler.plan_arrays (the vectorised plan of extract_packed) against extract.plan_many (chunk_plan per utterance,
    egs/voxceleb/v1/nnet/lib/extract.py:65-78) over edge lengths: below the minimum, exactly chunk_size, one frame over, exact
    multiples of the half-chunk hop, odd chunk sizes (Python-2 integer division of chunk_size / 2)."""
    from xvector_b200.extract import plan_many
    from xvector_b200.scheduler import plan_arrays
    rng = np.random.default_rng(3)
    for chunk, min_len in ((2000, 25), (300, 50), (301, 25), (10000, 25), (7, 3)):
        half = chunk // 2
        lens = np.concatenate([[0, 1, min_len - 1, min_len, chunk - 1, chunk, chunk + 1, chunk + half, chunk + half + 1,
                                chunk + 2 * half, 3 * chunk, 10 * chunk + 3], rng.integers(1, 6 * chunk, 200)])
        work = plan_many(lens, chunk, min_len)
        owners, starts, sizes = plan_arrays(lens, chunk, min_len)
        assert [(int(u), int(s), int(z)) for u, s, z in zip(owners, starts, sizes)] == [(u, s, z) for u, s, z in work]
    o, s, z = plan_arrays(np.array([3, 5]), 100, 25)
    assert o.size == 0 and s.size == 0 and z.size == 0


def test_host_losses_match_the_training_oracle():
    """model/loss.py of the package (validation) against oracle/train_ref.loss_forward_backward, which is pinned on the
    reference's own numpy AM-softmax (tests/test_train_oracle.py)."""
    from oracle import numpy_ref, synth, train_ref
    from xvector_b200.model import loss as L
    rng = np.random.default_rng(8)
    f = rng.standard_normal((16, 48)) * 3
    y = rng.integers(0, 37, 16)
    w = {"softmax/output/kernel": rng.standard_normal((48, 37)).astype(np.float32),
         "softmax/output/bias": rng.standard_normal(37).astype(np.float32)}
    for cfg, step in ((dict(synth.TDNN_AMSOFTMAX_CONFIG), 0), (dict(synth.TDNN_AMSOFTMAX_CONFIG), 12345),
                      (dict(synth.TDNN_AMSOFTMAX_CONFIG, loss_func="softmax"), 7)):
        params = numpy_ref.OracleParams(**cfg)
        ref, _, _ = train_ref.loss_forward_backward(f, y, w, params, step, want_grad=False)
        got = L.valid_loss(cfg["loss_func"], f.astype(np.float32), y, w, cfg, step)
        assert abs(got - ref) <= 1e-6 * abs(ref), (cfg["loss_func"], got, ref)
    with pytest.raises(NotImplementedError):
        L.valid_loss("asoftmax", f, y, w, {}, 0)


def test_trainer_valid_with_a_stub_engine(tmp_path):
    """Trainer.valid (model/trainer.py:756-884): sequential batches of (min + max) / 2 frames, eval-mode network output from
    the device path (a stub here: the mean over frames, projected), host loss head, mean of the batch losses, embeddings
    and labels on request, at most valid_max_iterations batches."""
    from oracle import synth
    from test_data_loader import _make_data
    from xvector_b200.misc.utils import ParamsPlain
    from xvector_b200.model import loss as L
    from xvector_b200.model.trainer import Trainer
    data, spklist, mats = _make_data(tmp_path)
    cfg = dict(synth.TDNN_AMSOFTMAX_CONFIG, min_segment_len=30, max_segment_len=40, num_speakers_per_batch=2,
               num_segments_per_speaker=2, valid_max_iterations=3)
    params = ParamsPlain()
    params.dict.update(cfg)
    rng = np.random.default_rng(9)
    proj = rng.standard_normal((7, 12)).astype(np.float32)
    weights = {"softmax/output/kernel": rng.standard_normal((12, 6)).astype(np.float32)}

    class StubEngine(object):
        calls = []

        def forward(self, packed, lengths, node):
            self.calls.append(node)
            t = int(lengths[0])
            pooled = packed.reshape(len(lengths), t, -1).mean(axis=1)
            return pooled @ proj if node == "output" else pooled

        def close(self):
            pass

    tr = Trainer(params, str(tmp_path / "model"), 7, num_speakers=6, single_cpu=True)
    tr.engine, tr.is_built, tr.is_loaded, tr._weights = StubEngine(), True, True, weights
    tr._has_model = lambda: False
    loss, emb, labels = tr.valid(data, spklist, output_embeddings=True)
    keys = list(mats)
    want, want_labels = [], []
    for i in range(3):                                          # valid_max_iterations of the 5 batches
        ks = [keys[(4 * i + j) % len(keys)] for j in range(4)]
        length = min(35, min(mats[k].shape[0] for k in ks))
        feats = np.stack([mats[k][:length] for k in ks])
        lab = np.array([int(k[3]) for k in ks])
        want.append(L.additive_margin_softmax(feats.mean(axis=1) @ proj, lab, weights["softmax/output/kernel"], cfg, 0))
        want_labels.append(lab)
    assert abs(loss - np.mean(want)) <= 1e-6 * abs(np.mean(want))
    assert emb.shape == (12, 7) and np.array_equal(labels, np.concatenate(want_labels))
    loss2, emb2, labels2 = tr.valid(data, spklist)
    assert abs(loss2 - loss) < 1e-12 and emb2 is None and labels2 is None
    with pytest.raises(NotImplementedError):
        tr.valid(data, spklist, batch_type="end2end")


def test_compute_cos_pairwise_eer():
    """misc/utils.py:307-346: separable speakers -> 0, labels independent of the embeddings -> about one half, and the value
    agrees with a direct threshold sweep of the same pair scores."""
    from xvector_b200.misc.utils import compute_cos_pairwise_eer
    rng = np.random.default_rng(10)
    centers = np.eye(8, 16) * 5
    labels = np.repeat(np.arange(8), 6)
    tight = centers[labels] + 0.01 * rng.standard_normal((48, 16))
    keep = tight.copy()
    assert compute_cos_pairwise_eer(tight, labels) < 1e-6
    assert np.array_equal(tight, keep)                          # the caller's array is left alone
    noise = rng.standard_normal((200, 16))
    assert 0.4 < compute_cos_pairwise_eer(noise, rng.integers(0, 5, 200)) < 0.6
    # overlapping clusters: compare with the crossing of the empirical miss / false-alarm curves
    loose = centers[labels] + 2.5 * rng.standard_normal((48, 16))
    eer = compute_cos_pairwise_eer(loose, labels)
    e = loose / np.linalg.norm(loose, axis=1, keepdims=True)
    i, j = np.triu_indices(48, k=1)
    sc, tgt = np.sum(e[i] * e[j], axis=1), labels[i] == labels[j]
    best = min((abs(np.mean(sc[tgt] < t) - np.mean(sc[~tgt] >= t)), 0.5 * (np.mean(sc[tgt] < t) + np.mean(sc[~tgt] >= t)))
               for t in np.sort(sc))
    assert abs(eer - best[1]) < 0.02
    # more embeddings than max_num_embeddings: subsampled every n // max
    big = rng.standard_normal((2500, 8))
    assert 0.4 < compute_cos_pairwise_eer(big, rng.integers(0, 3, 2500), max_num_embeddings=500) < 0.6
