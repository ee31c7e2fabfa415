"""Kaldi ark/scp I/O with the reference's names and behaviour (dataset/kaldi_io.py), served by lib/libxvkaldi.so
(include/xvkaldi.h, csrc/xv_kaldi_io.cpp).  SURVEY.md section 8(f) N1.

Same call shapes as the reference module the extractor imports (egs/voxceleb/v1/nnet/lib/extract.py:9):

    for key, mat in read_mat_ark("ark:copy-feats scp:feats.scp ark:- |"): ...
    with open_or_fd("ark:| copy-vector ark:- ark,scp:x.ark,x.scp", "wb") as f:
        write_vec_flt(f, vec, key=key)

`open_or_fd` returns a stream object that owns a native reader or writer (the reference returns a Python file
object; every function of this module accepts either a specifier string or such a stream).  There is no pure-Python
decoding path: a missing library raises at import of the first call.

`BatchReader` is the B200-side addition: it fills pinned host buffers with whole batches of utterances (packed rows
+ lengths + keys), which is the layout xv_forward takes."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", "libxvkaldi.so")

XVK_OK = 0
XVK_ECAPACITY = -5
MAT_F32, MAT_F64, MAT_CM, MAT_TEXT, VEC_F32, VEC_F64, VEC_TEXT = 1, 2, 3, 4, 5, 6, 7


class KaldiIOError(IOError):
    pass


class UnknownMatrixHeader(KaldiIOError):
    pass


class UnknownVectorHeader(KaldiIOError):
    pass


class UnsupportedDataType(KaldiIOError):
    pass


class BadInputFormat(KaldiIOError):
    pass


class SubprocessFailed(KaldiIOError):
    pass


_p = C.c_void_p
_i32p = C.POINTER(C.c_int32)
_i64p = C.POINTER(C.c_int64)

# name -> (restype, argtypes): every symbol include/xvkaldi.h declares
PROTOTYPES = {
    "xvk_reader_open": (_p, [C.c_char_p, C.c_int, C.c_char_p, C.c_size_t]),
    "xvk_next": (C.c_int, [_p, C.POINTER(C.c_char_p), _i32p, _i32p, _i32p]),
    "xvk_read_f32": (C.c_int, [_p, _p, C.c_int64]),
    "xvk_read_f64": (C.c_int, [_p, _p, C.c_int64]),
    "xvk_read_rows_f32": (C.c_int, [_p, C.c_int32, C.c_int32, _p, C.c_int64]),
    "xvk_read_batch_f32": (C.c_int, [_p, _p, C.c_int64, C.c_int32, C.c_int32, _i32p, _p, C.c_size_t, _i32p, _i64p]),
    "xvk_reader_set_threads": (C.c_int, [_p, C.c_int32]),
    "xvk_reader_error": (C.c_char_p, [_p]),
    "xvk_reader_close": (C.c_int, [_p]),
    "xvk_scp_open": (_p, [C.c_char_p, C.c_char_p, C.c_size_t]),
    "xvk_scp_count": (C.c_int64, [_p]),
    "xvk_scp_key": (C.c_char_p, [_p, C.c_int64]),
    "xvk_scp_rxfile": (C.c_char_p, [_p, C.c_int64]),
    "xvk_scp_dims": (C.c_int, [_p, C.c_int64, _i32p, _i32p, _i32p]),
    "xvk_scp_read_f32": (C.c_int, [_p, C.c_int64, _p, C.c_int64]),
    "xvk_scp_read_many_f32": (C.c_int, [_p, _i64p, C.c_int64, _p, C.c_int64, _i64p, C.c_int64, _i32p, C.c_int32]),
    "xvk_scp_read_segments_f32": (C.c_int, [_p, _i64p, _i32p, _i32p, C.c_int64, _p, C.c_int64, _i64p, C.c_int64, C.c_int32]),
    "xvk_scp_dims_many": (C.c_int, [_p, _i64p, C.c_int64, _i32p, _i32p, C.c_int32]),
    "xvk_scp_error": (C.c_char_p, [_p]),
    "xvk_scp_close": (None, [_p]),
    "xvk_writer_open": (_p, [C.c_char_p, C.c_char_p, C.c_size_t]),
    "xvk_write_mat_f32": (C.c_int, [_p, C.c_char_p, _p, C.c_int32, C.c_int32, C.c_int64]),
    "xvk_write_mat_f64": (C.c_int, [_p, C.c_char_p, _p, C.c_int32, C.c_int32, C.c_int64]),
    "xvk_write_vec_f32": (C.c_int, [_p, C.c_char_p, _p, C.c_int32]),
    "xvk_write_vec_f64": (C.c_int, [_p, C.c_char_p, _p, C.c_int32]),
    "xvk_write_vecs_f32": (C.c_int, [_p, C.c_char_p, C.c_int64, _p, C.c_int32, C.c_int64, _p]),
    "xvk_writer_flush": (C.c_int, [_p]),
    "xvk_writer_error": (C.c_char_p, [_p]),
    "xvk_writer_close": (C.c_int, [_p]),
    "xvk_crc32c": (C.c_uint32, [C.c_char_p, C.c_size_t, C.c_uint32]),
    "xvk_version": (C.c_char_p, []),
}

_lib = None


def load():
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise KaldiIOError("%s is missing: run `make -C tf-kaldi-speaker_b200/csrc` (or __graft_entry__.build())" % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def _raise(code, msg):
    msg = msg.decode("latin1") if isinstance(msg, bytes) else str(msg)
    if "UnknownMatrixHeader" in msg:
        raise UnknownMatrixHeader(msg)
    if "BadInputFormat" in msg:
        raise BadInputFormat(msg)
    if "returned" in msg and msg.startswith("cmd "):
        raise SubprocessFailed(msg)
    raise KaldiIOError("libxvkaldi error %d: %s" % (code, msg))


def _ptr(arr):
    return arr.ctypes.data_as(C.c_void_p)


class _Reader(object):
    """Native sequential reader (an rspecifier opened for 'rb')."""
    mode = "rb"

    def __init__(self, rspecifier, keyless=False):
        err = C.create_string_buffer(512)
        self._lib = load()
        self._h = self._lib.xvk_reader_open(rspecifier.encode("latin1"), int(keyless), err, 512)
        if not self._h:
            raise KaldiIOError(err.value.decode("latin1"))
        self.name = rspecifier

    def _check(self, rc):
        if rc < 0:
            _raise(rc, self._lib.xvk_reader_error(self._h))
        return rc

    def next_header(self):
        """-> (key, kind, rows, cols) or None at the end of the stream."""
        key, kind, rows, cols = C.c_char_p(), C.c_int32(), C.c_int32(), C.c_int32()
        if self._check(self._lib.xvk_next(self._h, C.byref(key), C.byref(kind), C.byref(rows), C.byref(cols))) == 0:
            return None
        return key.value.decode("latin1"), kind.value, rows.value, cols.value

    def read_body(self, kind, rows, cols, out=None):
        """Decodes the pending object with the dtype the reference returns (float64 for DM / DV, else float32)."""
        f64 = kind in (MAT_F64, VEC_F64)
        if out is None:
            out = np.empty((rows, cols), np.float64 if f64 else np.float32)
        fn = self._lib.xvk_read_f64 if out.dtype == np.float64 else self._lib.xvk_read_f32
        self._check(fn(self._h, _ptr(out), out.strides[0] // out.itemsize if out.ndim == 2 and rows > 1 else max(cols, 1)))
        return out

    def read_rows(self, start, length, cols):
        out = np.empty((length, cols), np.float32)
        self._check(self._lib.xvk_read_rows_f32(self._h, start, length, _ptr(out), max(cols, 1)))
        return out

    def close(self):
        if self._h:
            h, self._h = self._h, None
            rc = self._lib.xvk_reader_close(h)
            if rc != XVK_OK:
                raise SubprocessFailed("cmd of %s returned a non-zero status" % self.name)

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            if self._h:
                self._lib.xvk_reader_close(self._h)
                self._h = None
        except Exception:
            pass


class _Writer(object):
    """Native writer (a wspecifier opened for 'wb')."""
    mode = "wb"

    def __init__(self, wspecifier):
        err = C.create_string_buffer(512)
        self._lib = load()
        self._h = self._lib.xvk_writer_open(wspecifier.encode("latin1"), err, 512)
        if not self._h:
            raise KaldiIOError(err.value.decode("latin1"))
        self.name = wspecifier

    def _check(self, rc):
        if rc < 0:
            _raise(rc, self._lib.xvk_writer_error(self._h))
        return rc

    def flush(self):
        self._check(self._lib.xvk_writer_flush(self._h))

    def close(self):
        if self._h:
            h, self._h = self._h, None
            rc = self._lib.xvk_writer_close(h)
            if rc != XVK_OK:
                raise KaldiIOError("closing %s failed (write error or the command returned a non-zero status)" % self.name)

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            if self._h:
                self._lib.xvk_writer_close(self._h)
                self._h = None
        except Exception:
            pass


def open_or_fd(file, mode="rb"):
    """fd = open_or_fd(file): file / gzipped file / pipe / "path:offset", with an optional ark:/scp: prefix
    (dataset/kaldi_io.py:625-654).  An already opened stream is passed through."""
    if isinstance(file, (_Reader, _Writer)):
        return file
    if not isinstance(file, str):
        raise TypeError("open_or_fd takes a specifier string or a stream returned by open_or_fd, got %r" % type(file))
    if mode in ("rb", "r"):
        return _Reader(file)
    if mode in ("wb", "w"):
        return _Writer(file)
    raise ValueError("invalid mode %s" % mode)


def read_key(fd):
    """[key] = read_key(fd): the next utterance key, None at the end of the stream (dataset/kaldi_io.py:693-707).
    The object header is parsed with the key; the matching read_mat(fd) / read_vec_flt(fd) returns the body."""
    head = fd.next_header()
    if head is None:
        fd._head = None
        return None
    fd._head = head
    return head[0]


def _body(fd, keyless_ok=True):
    head = getattr(fd, "_head", None)
    fd._head = None
    if head is None:
        head = fd.next_header()
        if head is None:
            raise BadInputFormat("BadInputFormat: end of stream where an object was expected")
    return head


# ---------------------------------------------------------------------------------------------- matrices
def read_mat(file_or_fd):
    """[mat] = read_mat(file_or_fd): one matrix, binary (FM / DM / CM) or text (dataset/kaldi_io.py:996-1011)."""
    own = not isinstance(file_or_fd, _Reader)
    fd = _Reader(file_or_fd, keyless=True) if own else file_or_fd
    try:
        _, kind, rows, cols = _body(fd)
        if kind in (VEC_F32, VEC_F64, VEC_TEXT):
            raise UnknownMatrixHeader("UnknownMatrixHeader: The header contained '%s'" % ("FV " if kind == VEC_F32 else "DV "))
        return fd.read_body(kind, rows, cols)
    finally:
        if own:
            fd.close()


def read_mat_ark(file_or_fd):
    """generator(key, mat) over an ark file / stream (dataset/kaldi_io.py:974-994)."""
    own = not isinstance(file_or_fd, _Reader)
    fd = open_or_fd(file_or_fd)
    try:
        key = read_key(fd)
        while key:
            yield key, read_mat(fd)
            key = read_key(fd)
    finally:
        if own:
            fd.close()


def read_mat_scp(file_or_fd):
    """generator(key, mat) following a Kaldi scp (dataset/kaldi_io.py:954-972)."""
    scp = ScpTable(file_or_fd)
    try:
        for i in range(len(scp)):
            yield scp.key(i), read_mat(scp.rxfile(i))
    finally:
        scp.close()


def write_mat(file_or_fd, m, key=""):
    """write_mat(f, m, key=''): binary FM / DM matrix, with its key when writing an ark (dataset/kaldi_io.py:1175-1208)."""
    m = np.asarray(m)
    if m.dtype not in (np.float32, np.float64):
        raise UnsupportedDataType("'%s', please use 'float32' or 'float64'" % m.dtype)
    if m.ndim != 2:
        raise ValueError("write_mat needs a rank-2 array, got shape %s" % (m.shape,))
    m = np.ascontiguousarray(m)
    own = not isinstance(file_or_fd, _Writer)
    fd = open_or_fd(file_or_fd, "wb")
    try:
        fn = fd._lib.xvk_write_mat_f32 if m.dtype == np.float32 else fd._lib.xvk_write_mat_f64
        fd._check(fn(fd._h, key.encode("latin1"), _ptr(m), m.shape[0], m.shape[1], max(m.shape[1], 1)))
    finally:
        if own:
            fd.close()


# ---------------------------------------------------------------------------------------------- vectors
def read_vec_flt(file_or_fd):
    """[flt-vec] = read_vec_flt(file_or_fd): binary FV / DV or text vector (dataset/kaldi_io.py:855-886)."""
    own = not isinstance(file_or_fd, _Reader)
    fd = _Reader(file_or_fd, keyless=True) if own else file_or_fd
    try:
        _, kind, rows, cols = _body(fd)
        if kind in (MAT_F32, MAT_F64, MAT_CM):
            raise UnknownVectorHeader("The header contained '%s'" % {MAT_F32: "FM ", MAT_F64: "DM ", MAT_CM: "CM "}[kind])
        if kind == MAT_TEXT:                                   # " [ 1 2 3 ]": the reference parses one line into float64
            return fd.read_body(kind, rows, cols, np.empty((rows, cols), np.float64)).reshape(-1)
        return fd.read_body(kind, rows, cols).reshape(-1)
    finally:
        if own:
            fd.close()


def read_vec_flt_ark(file_or_fd):
    """generator(key, vec) over an ark file / stream (dataset/kaldi_io.py:838-853)."""
    own = not isinstance(file_or_fd, _Reader)
    fd = open_or_fd(file_or_fd)
    try:
        key = read_key(fd)
        while key:
            yield key, read_vec_flt(fd)
            key = read_key(fd)
    finally:
        if own:
            fd.close()


def read_vec_flt_scp(file_or_fd):
    """generator(key, vec) following a Kaldi scp (dataset/kaldi_io.py:818-836)."""
    scp = ScpTable(file_or_fd)
    try:
        for i in range(len(scp)):
            yield scp.key(i), read_vec_flt(scp.rxfile(i))
    finally:
        scp.close()


def write_vec_flt(file_or_fd, v, key=""):
    """write_vec_flt(f, v, key=''): binary FV / DV vector (dataset/kaldi_io.py:915-946)."""
    v = np.asarray(v)
    if v.dtype not in (np.float32, np.float64):
        raise UnsupportedDataType("'%s', please use 'float32' or 'float64'" % v.dtype)
    v = np.ascontiguousarray(v).reshape(-1)
    own = not isinstance(file_or_fd, _Writer)
    fd = open_or_fd(file_or_fd, "wb")
    try:
        fn = fd._lib.xvk_write_vec_f32 if v.dtype == np.float32 else fd._lib.xvk_write_vec_f64
        fd._check(fn(fd._h, key.encode("latin1"), _ptr(v), v.shape[0]))
    finally:
        if own:
            fd.close()


def write_vecs_flt(file_or_fd, keys, mat, skip=None):
    """All rows of `mat` [n, dim] float32 as FV vectors under `keys` in ONE native call (the embeddings of a batch)."""
    mat = np.ascontiguousarray(mat, np.float32)
    if mat.ndim != 2 or mat.shape[0] != len(keys):
        raise ValueError("write_vecs_flt: %d keys for an array of shape %s" % (len(keys), mat.shape))
    blob = b"".join(k.encode("latin1") + b"\0" for k in keys)
    sk = None if skip is None else np.ascontiguousarray(skip, np.uint8)
    own = not isinstance(file_or_fd, _Writer)
    fd = open_or_fd(file_or_fd, "wb")
    try:
        fd._check(fd._lib.xvk_write_vecs_f32(fd._h, blob, len(keys), _ptr(mat), mat.shape[1], max(mat.shape[1], 1),
                                             None if sk is None else _ptr(sk)))
    finally:
        if own:
            fd.close()


# ---------------------------------------------------------------------------------------------- scp / batches
class ScpTable(object):
    """A parsed scp: keys, rxfiles, header-only dimension queries and multi-threaded random-access reads."""

    def __init__(self, path):
        if isinstance(path, str) and path.startswith("scp:"):
            path = path[4:]
        err = C.create_string_buffer(512)
        self._lib = load()
        self._h = self._lib.xvk_scp_open(path.encode("latin1"), err, 512)
        if not self._h:
            raise KaldiIOError(err.value.decode("latin1"))

    def _check(self, rc):
        if rc < 0:
            _raise(rc, self._lib.xvk_scp_error(self._h))

    def __len__(self):
        return int(self._lib.xvk_scp_count(self._h))

    def key(self, i):
        return self._lib.xvk_scp_key(self._h, i).decode("latin1")

    def rxfile(self, i):
        return self._lib.xvk_scp_rxfile(self._h, i).decode("latin1")

    def keys(self):
        return [self.key(i) for i in range(len(self))]

    def dims(self, idx=None, threads=8):
        """(rows, cols) int32 arrays of the entries `idx` (default: all), headers only."""
        idx = np.arange(len(self), dtype=np.int64) if idx is None else np.ascontiguousarray(idx, np.int64)
        rows, cols = np.empty(len(idx), np.int32), np.empty(len(idx), np.int32)
        self._check(self._lib.xvk_scp_dims_many(self._h, idx.ctypes.data_as(_i64p), len(idx), rows.ctypes.data_as(_i32p),
                                                cols.ctypes.data_as(_i32p), threads))
        return rows, cols

    def read_many(self, idx, out=None, rows=None, threads=8):
        """Entries `idx` packed back to back: -> (features [sum rows, dim] float32, lengths int32).  `out` may be a
        pinned array to fill in place."""
        idx = np.ascontiguousarray(idx, np.int64)
        if rows is None:
            rows, cols = self.dims(idx, threads)
            dim = int(cols[0]) if len(cols) else 0
        else:
            rows = np.ascontiguousarray(rows, np.int32)
            dim = out.shape[1] if out is not None else int(self.dims(idx[:1])[1][0])
        off = np.concatenate([[0], np.cumsum(rows, dtype=np.int64)]).astype(np.int64)
        total = int(off[-1])
        if out is None:
            out = np.empty((total, dim), np.float32)
        if out.shape[0] < total or out.dtype != np.float32 or not out.flags.c_contiguous:
            raise ValueError("read_many: destination must be C-contiguous float32 with at least %d rows" % total)
        got = np.empty(len(idx), np.int32)
        self._check(self._lib.xvk_scp_read_many_f32(self._h, idx.ctypes.data_as(_i64p), len(idx), _ptr(out), out.shape[1],
                                                    off[:-1].copy().ctypes.data_as(_i64p), out.shape[0],
                                                    got.ctypes.data_as(_i32p), threads))
        if not np.array_equal(got, rows):
            raise KaldiIOError("scp entries changed size between the header pass and the read")
        return out[:total], rows

    def read_segments(self, idx, starts, length, out=None, threads=8):
        """Rows [starts[k], starts[k] + length) of entries idx[k] -> [n, length, dim] float32 (`out`: a C-contiguous array of
        that shape to fill in place): the random segments of a training batch in one native call."""
        idx = np.ascontiguousarray(idx, np.int64)
        starts = np.ascontiguousarray(starts, np.int32)
        n, length = len(idx), int(length)
        if out is None:
            dim = int(self.dims(idx[:1])[1][0]) if n else 0
            out = np.empty((n, length, dim), np.float32)
        if out.shape[:2] != (n, length) or out.dtype != np.float32 or not out.flags.c_contiguous:
            raise ValueError("read_segments: destination must be C-contiguous float32 [%d, %d, dim]" % (n, length))
        lens = np.full(n, length, np.int32)
        off = np.arange(n, dtype=np.int64) * length
        self._check(self._lib.xvk_scp_read_segments_f32(self._h, idx.ctypes.data_as(_i64p), starts.ctypes.data_as(_i32p),
                                                        lens.ctypes.data_as(_i32p), n, _ptr(out), out.shape[2],
                                                        off.ctypes.data_as(_i64p), n * length, threads))
        return out

    def close(self):
        if self._h:
            self._lib.xvk_scp_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class BatchReader(object):
    """Streams an ark as packed batches: yields (keys, features [rows, dim] float32, lengths int32) with at most
    `max_rows` rows / `max_utts` utterances each, decoded natively into `buffer` (pass a pinned array to make the
    H2D copy of xv_forward asynchronous).  The arrays of a batch are views: consume them before the next one."""

    def __init__(self, rspecifier, dim, max_rows=1 << 20, max_utts=4096, buffer=None, threads=8, batch_rows=None):
        """max_rows: capacity of the buffer (the longest utterance that can be read); batch_rows: where a batch is cut
        (default: at the capacity).  A batch holds at most batch_rows rows; one whose first utterance alone is longer than that
        is filled up to the capacity instead."""
        self.batch_rows = None if batch_rows is None else max(int(batch_rows), 1)
        self.fd = open_or_fd(rspecifier)
        self.fd._check(self.fd._lib.xvk_reader_set_threads(self.fd._h, max(int(threads), 1)))   # regular files: parallel decode
        self.dim = int(dim)
        self.max_utts = int(max_utts)
        self.buffer = np.empty((max_rows, dim), np.float32) if buffer is None else buffer
        if self.buffer.dtype != np.float32 or self.buffer.ndim != 2 or self.buffer.shape[1] != dim or not self.buffer.flags.c_contiguous:
            raise ValueError("BatchReader buffer must be C-contiguous float32 [rows, %d]" % dim)
        self._lengths = np.empty(self.max_utts, np.int32)
        self._keys = C.create_string_buffer(64 * self.max_utts)

    def __iter__(self):
        lib, fd = self.fd._lib, self.fd
        n, rows = C.c_int32(), C.c_int64()
        more = 1
        while more:
            full = self.buffer.shape[0]
            cap = full if self.batch_rows is None else min(self.batch_rows, full)
            rc = lib.xvk_read_batch_f32(fd._h, _ptr(self.buffer), cap, self.dim, self.max_utts,
                                        self._lengths.ctypes.data_as(_i32p), self._keys, len(self._keys), C.byref(n), C.byref(rows))
            if rc == XVK_ECAPACITY and cap < full:       # one utterance longer than a batch: it gets the whole buffer
                rc = lib.xvk_read_batch_f32(fd._h, _ptr(self.buffer), full, self.dim, self.max_utts,
                                            self._lengths.ctypes.data_as(_i32p), self._keys, len(self._keys), C.byref(n), C.byref(rows))
            more = fd._check(rc)
            if n.value:
                # maxsplit: the rest of the 64 * max_utts byte buffer is NULs -- an unbounded split makes a list entry of
                # every one of them (2-3 ms per batch, a third of the reader's time on 0.5 M-row batches)
                keys = self._keys.raw.split(b"\0", n.value)[:n.value]
                yield [k.decode("latin1") for k in keys], self.buffer[:rows.value], self._lengths[:n.value].copy()
        self.close()

    def close(self):
        self.fd.close()


class FeatureReader(object):
    """dataset/kaldi_io.py:40-150: random access to the features of a Kaldi data directory by "utt path:offset" strings (the
    lines of feats.scp), whole matrices (read) or segments (read_segment), with the reference's argument rules: `length` longer
    than the utterance is cut to it, `shuffle` draws the start, an explicit `start` forbids shuffle.  -> (matrix, start).
    Reads go through the native table (one mapping per ark, no descriptor juggling); segments read only their own rows.
    FeatureReaderV2's byte-offset arithmetic (dataset/kaldi_io.py:152-338) exists for the same purpose and is covered by the
    same calls."""

    def __init__(self, data):
        import random
        self.data = data
        self._rd = random.Random()
        scp = os.path.join(data, "feats.scp")
        self._table = ScpTable(scp)
        self._index = {}
        for i in range(len(self._table)):
            self._index[self._table.key(i) + " " + self._table.rxfile(i)] = i
        self.dim = self.get_dim()
        self.utt2num_frames = {}
        n2f = os.path.join(data, "utt2num_frames")
        assert os.path.exists(n2f), "[Error] Expect utt2num_frames exists in %s " % data
        with open(n2f) as f:
            for line in f:
                utt, length = line.strip().split(" ")
                self.utt2num_frames[utt] = int(length)

    def get_dim(self):
        return int(self._table.dims(np.arange(1, dtype=np.int64))[1][0])

    def close(self):
        self._table.close()

    def _entry(self, file_or_fd):
        try:
            return self._index[file_or_fd.strip()]
        except KeyError:
            raise IOError("Cannot read features from %s" % file_or_fd)

    def read(self, file_or_fd, length=None, shuffle=False, start=None):
        i = self._entry(file_or_fd)
        try:
            mat, _ = self._table.read_many(np.array([i], np.int64), threads=1)
        except KaldiIOError:
            raise IOError("Cannot read features from %s" % file_or_fd)
        if length is not None:
            if start is None:
                num = mat.shape[0]
                length = num if length > num else length
                start = self._rd.randint(0, num - length) if shuffle else 0
            else:
                assert not shuffle, "The start point is specified, thus shuffling is invalid."
            mat = mat[start:start + length, :]
        return mat, start

    def read_segment(self, file_or_fd, length=None, shuffle=False, start=None):
        if length is None:
            return self.read(file_or_fd)
        i = self._entry(file_or_fd)
        utt = file_or_fd.strip().split(" ")[0]
        if start is None:
            num = self.utt2num_frames[utt]
            length = num if length > num else length
            start = self._rd.randint(0, num - length) if shuffle else 0
        else:
            assert not shuffle, "The start point is specified, thus shuffling is invalid."
        try:
            mat = self._table.read_segments(np.array([i], np.int64), np.array([start], np.int32), int(length), threads=1)[0]
        except KaldiIOError:
            raise IOError("Cannot read features from %s" % file_or_fd)
        return mat, start
