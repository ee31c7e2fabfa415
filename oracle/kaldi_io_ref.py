"""Kaldi ark wire formats restated in numpy.

TEST INFRASTRUCTURE: only tests/ may import this module; the product path is lib/libxvkaldi.so and never routes
through it.  Every function cites the reference lines it restates (dataset/kaldi_io.py of the reference tree).

Pinned: tests/golden/kaldi_io_golden.npz + kaldi_io_*.ark were produced by IMPORTING the reference module
(tests/golden/make_golden.py: reference write_mat / write_vec_flt wrote the FM/DM/FV/DV arks, reference read_mat_ark
decoded the CM ark), and tests/test_kaldi_io.py checks this restatement and the native library against them bit
for bit.  The CM *compressor* below is not part of the reference (it only reads CM); it follows the published
Kaldi algorithm (src/matrix/compressed-matrix.cc, format 1 = kOneByteWithColHeaders) closely enough to produce
valid streams for the decoders and is not itself a parity claim.

Float semantics of the CM decode: the reference evaluates `min + range * 1.52590218966964e-05 * value` and
`p0 + (p25 - p0) / 64. * vec` on numpy float32 scalars / uint8 arrays.  Under NumPy >= 2 (this image: 2.3) Python
floats are weak, so the whole expression is float32, rounded after every operation in the written order -- that is
what the golden vectors hold and what both decoders reproduce bit for bit.  (NumPy 1.x promoted the same
expressions to float64 and Kaldi's C++ mixes float and double; the three agree to 1 float32 ulp.)"""
from __future__ import annotations

import struct

import numpy as np

F32 = np.float32


# ------------------------------------------------------------------------------------------------ writers
def write_mat_bytes(m, key=""):
    """write_mat (dataset/kaldi_io.py:1175-1208): [key ' '] '\\0B' 'FM '|'DM ' '\\4' rows '\\4' cols data."""
    m = np.asarray(m)
    tok = {np.dtype(np.float32): b"FM ", np.dtype(np.float64): b"DM "}[m.dtype]
    head = (key.encode("latin1") + b" ") if key != "" else b""
    return head + b"\0B" + tok + b"\4" + struct.pack("<i", m.shape[0]) + b"\4" + struct.pack("<i", m.shape[1]) + m.tobytes()


def write_vec_bytes(v, key=""):
    """write_vec_flt (dataset/kaldi_io.py:915-946)."""
    v = np.asarray(v)
    tok = {np.dtype(np.float32): b"FV ", np.dtype(np.float64): b"DV "}[v.dtype]
    head = (key.encode("latin1") + b" ") if key != "" else b""
    return head + b"\0B" + tok + b"\4" + struct.pack("<i", v.shape[0]) + v.tobytes()


# ------------------------------------------------------------------------------------------------ CM
def uint16_to_float(value, gmin, grange):
    """dataset/kaldi_io.py:1083-1084, float32 after every operation, in the written order."""
    t = F32(grange) * F32(1.52590218966964e-05)
    t = F32(t * F32(value))
    return F32(F32(gmin) + t)


def uint8_to_float(vec, p0, p25, p75, p100):
    """dataset/kaldi_io.py:1087-1100 (three linear pieces: [0,64], (64,192], (192,255])."""
    vec = np.asarray(vec, np.uint8)
    out = np.empty(vec.shape, np.float32)
    lo, hi = vec <= 64, vec > 192
    mid = ~lo & ~hi
    q0 = F32(F32(p25 - p0) / F32(64.0))
    q1 = F32(F32(p75 - p25) / F32(128.0))
    q2 = F32(F32(p100 - p75) / F32(63.0))
    out[lo] = p0 + (q0 * vec[lo].astype(np.float32)).astype(np.float32)
    out[mid] = p25 + (q1 * (vec[mid] - np.uint8(64)).astype(np.float32)).astype(np.float32)
    out[hi] = p75 + (q2 * (vec[hi] - np.uint8(192)).astype(np.float32)).astype(np.float32)
    return out


def decode_cm(body):
    """_read_compressed_mat (dataset/kaldi_io.py:1071-1115): body = bytes after the 'CM ' token."""
    gmin, grange, rows, cols = struct.unpack("<ffii", body[:16])
    heads = np.frombuffer(body, np.uint16, cols * 4, 16).reshape(cols, 4)
    data = np.frombuffer(body, np.uint8, cols * rows, 16 + cols * 8).reshape(cols, rows)
    mat = np.empty((cols, rows), np.float32)
    for c in range(cols):
        p = [uint16_to_float(h, F32(gmin), F32(grange)) for h in heads[c]]
        mat[c] = uint8_to_float(data[c], *p)
    return mat.T, 16 + cols * 8 + cols * rows


def compress_cm(m):
    """Format-1 compressed matrix of `m` (see the module docstring: test-side producer, Kaldi's published algorithm)."""
    m = np.asarray(m, np.float32)
    rows, cols = m.shape
    gmin = F32(m.min())
    grange = F32(m.max() - gmin)
    if grange <= 0:
        grange = F32(1.0e-05)

    def to_u16(v):
        f = (np.float64(v) - np.float64(gmin)) / np.float64(grange)
        return int(min(max(f, 0.0), 1.0) * 65535 + 0.499)

    heads = np.zeros((cols, 4), np.uint16)
    data = np.zeros((cols, rows), np.uint8)
    for c in range(cols):
        col = np.sort(m[:, c])
        q = rows // 4
        u0 = min(to_u16(col[0]), 65532)
        u25 = min(max(to_u16(col[q]), u0 + 1), 65533)
        u75 = min(max(to_u16(col[min(3 * q, rows - 1)]), u25 + 1), 65534)
        u100 = max(to_u16(col[-1]), u75 + 1)
        heads[c] = (u0, u25, u75, u100)
        p0, p25, p75, p100 = [np.float64(uint16_to_float(u, gmin, grange)) for u in heads[c]]
        x = m[:, c].astype(np.float64)
        lo = np.clip(np.floor((x - p0) / (p25 - p0) * 64.0 + 0.5), 0, 64)
        mid = np.clip(np.floor(64 + (x - p25) / (p75 - p25) * 128.0 + 0.5), 64, 192)
        hi = np.clip(np.floor(192 + (x - p75) / (p100 - p75) * 63.0 + 0.5), 192, 255)
        data[c] = np.where(x < p25, lo, np.where(x < p75, mid, hi)).astype(np.uint8)
    return struct.pack("<ffii", gmin, grange, rows, cols) + heads.tobytes() + data.tobytes()


def write_cm_bytes(m, key=""):
    head = (key.encode("latin1") + b" ") if key != "" else b""
    return head + b"\0BCM " + compress_cm(m)


# ------------------------------------------------------------------------------------------------ reader
def parse_ark(blob):
    """read_mat_ark / read_vec_flt_ark over a whole ark held in memory -> [(key, array)] (dataset/kaldi_io.py:693-707,
    974-1031, 838-886).  Matrices come back rank 2, vectors rank 1, with the dtype the reference returns."""
    out = []
    pos, n = 0, len(blob)
    while pos < n:
        sp = blob.find(b" ", pos)
        if sp < 0:
            break
        key = blob[pos:sp].decode("latin1").strip()
        if key == "":
            break
        pos = sp + 1
        if blob[pos:pos + 2] != b"\0B":
            raise ValueError("text objects are not restated here")
        tok = blob[pos + 2:pos + 5]
        pos += 5
        if tok in (b"FM ", b"DM "):
            rows, cols = struct.unpack("<xixi", blob[pos:pos + 10])
            dt = np.float32 if tok == b"FM " else np.float64
            pos += 10
            out.append((key, np.frombuffer(blob, dt, rows * cols, pos).reshape(rows, cols)))
            pos += rows * cols * np.dtype(dt).itemsize
        elif tok in (b"FV ", b"DV "):
            (dim,) = struct.unpack("<xi", blob[pos:pos + 5])
            dt = np.float32 if tok == b"FV " else np.float64
            pos += 5
            out.append((key, np.frombuffer(blob, dt, dim, pos)))
            pos += dim * np.dtype(dt).itemsize
        elif tok == b"CM ":
            mat, used = decode_cm(blob[pos:])
            out.append((key, mat))
            pos += used
        else:
            raise ValueError("UnknownMatrixHeader: The header contained '%s'" % tok.decode("latin1"))
    return out
