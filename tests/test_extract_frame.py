"""Host logic of extract_frame.py / extract_attention.py (edge-replication padding, consecutive chunks, first-chunk
attention) with a stand-in predictor; the GPU test of the CLIs is tests/test_gpu_dropin.py."""
import numpy as np
import pytest


class _Shrink(object):
    """predict() of a frame-level node with total context 14: drops 7 frames each side, keeps 2 channels."""

    def __init__(self):
        self.calls = []

    def predict(self, f):
        f = np.asarray(f)
        self.calls.append(f.shape)
        return f[..., 7:-7, :2] * 2.0


def test_pad_edges_and_chunks():
    from xvector_b200.extract_frame import chunk_plan_frames, extract_attention_one, extract_frames_one, pad_edges
    x = np.arange(60 * 3, dtype=np.float32).reshape(60, 3)
    y = pad_edges(x[7:-7], 60)
    assert y.shape == (60, 3) and (y[:8] == x[7]).all() and (y[-8:] == x[52]).all() and (y[7:53] == x[7:53]).all()
    with pytest.raises(ValueError):
        pad_edges(x[:45], 60)                                      # odd total context cannot be centred
    assert chunk_plan_frames(60, 25) == [(0, 25), (25, 25), (50, 10)]      # extract_frame.py:72-79
    assert chunk_plan_frames(50, 25) == [(0, 25), (25, 25)]
    t = _Shrink()
    assert extract_frames_one(t, x, chunk_size=100, min_chunk_size=61) is None and not t.calls    # too short: skipped
    whole = extract_frames_one(t, x, chunk_size=100, min_chunk_size=25)
    assert whole.shape == (60, 2) and np.array_equal(whole[7:53], x[7:53, :2] * 2)
    t = _Shrink()
    parts = extract_frames_one(t, x, chunk_size=20, min_chunk_size=25)      # three full chunks: one batched call + the last
    assert t.calls == [(2, 20, 3), (20, 3)] and parts.shape == (60, 2)
    for c in range(3):
        seg = x[20 * c:20 * c + 20]
        assert np.array_equal(parts[20 * c:20 * c + 20], pad_edges(seg[7:-7, :2] * 2, 20))
    t = _Shrink()
    tail = extract_frames_one(t, x, chunk_size=25, min_chunk_size=25)       # 10-frame tail < receptive field: merged
    assert t.calls == [(2, 25, 3), (35, 3)] and tail.shape == (60, 2)
    assert np.array_equal(tail[25:], pad_edges(x[25:][7:-7, :2] * 2, 35))
    t = _Shrink()
    att = extract_attention_one(t, x, chunk_size=40, min_chunk_size=25)     # extract_attention.py:56-58: first chunk only
    assert t.calls == [(40, 3)] and att.shape == (26, 2)
    assert extract_attention_one(t, x[:10], chunk_size=40, min_chunk_size=25) is None
