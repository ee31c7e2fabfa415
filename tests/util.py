"""Shared helpers for the parity tests: tolerances of BASELINE.json / SURVEY.md section 8(c)."""
import numpy as np

TOL = {"bf16": 1e-2, "fp32": 1e-4}       # max_i |e - e_ref| / max_i |e_ref|, per utterance
MIN_COS = 0.9999


def max_rel(out, ref):
    out = np.asarray(out, np.float64).reshape(len(ref), -1)
    ref = np.asarray(ref, np.float64).reshape(len(ref), -1)
    return (np.abs(out - ref).max(axis=1) / np.maximum(np.abs(ref).max(axis=1), 1e-30)).max()


def tensor_rel(out, ref):
    out = np.asarray(out, np.float64)
    ref = np.asarray(ref, np.float64)
    return np.abs(out - ref).max() / max(np.abs(ref).max(), 1e-30)


def min_cos(out, ref):
    out = np.asarray(out, np.float64).reshape(len(ref), -1)
    ref = np.asarray(ref, np.float64).reshape(len(ref), -1)
    return (np.sum(out * ref, axis=1) / (np.linalg.norm(out, axis=1) * np.linalg.norm(ref, axis=1))).min()


def bf16_round(x):
    """Round-to-nearest-even to bf16, returned as float32."""
    x = np.ascontiguousarray(x, np.float32)
    u = x.view(np.uint32).astype(np.uint64)
    u = (u + 0x7FFF + ((u >> 16) & 1)) & 0xFFFF0000
    return u.astype(np.uint32).view(np.float32).reshape(x.shape)


def oracle_embeddings(feats, lengths, cfg, weights, dim, node=None):
    from oracle import numpy_ref
    params = numpy_ref.OracleParams(**cfg)
    offs = np.concatenate([[0], np.cumsum(lengths)])
    return [numpy_ref.predict(feats[offs[i]:offs[i + 1]], params, weights, dim, node=node) for i in range(len(lengths))]
