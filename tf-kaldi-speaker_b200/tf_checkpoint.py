"""TensorFlow checkpoints without TensorFlow.

The reference stores its models with tf.train.Saver (model/trainer.py:296-303) and restores them through
tf.train.get_checkpoint_state + saver.restore (:277-294); misc/utils.py:251-304 (get_checkpoint) picks the best step
from `valid_loss` and rewrites the `checkpoint` state file.  Neither TensorFlow nor a saved model ships with the
reference tree, so this reader follows the PUBLISHED on-disk formats:

* `checkpoint`  -- text proto: model_checkpoint_path: "..." / all_model_checkpoint_paths: "..." lines.
* `<prefix>.index` -- tensor-bundle index: a LevelDB-format table (tensorflow/core/lib/io/table*.cc): data blocks of
  prefix-compressed (shared, non_shared, value_len, key delta, value) entries + a restart array, each block followed by
  a 1-byte compression type and a masked CRC32C; an index block of block handles; a 48-byte footer ending in the magic
  0xdb4775248b80fb57.  Key "" holds BundleHeaderProto, every other key is a variable name holding BundleEntryProto
  {dtype=1, shape=2, shard_id=3, offset=4, size=5, crc32c=6} (tensorflow/core/protobuf/tensor_bundle.proto).
* `<prefix>.data-NNNNN-of-MMMMM` -- the raw little-endian tensor bytes at (offset, size).

parity unpinned: there is no golden checkpoint to test against in this environment; tests round-trip through an
independent writer of the same published format (oracle/tf_bundle_ref.py) and verify the CRCs that TensorFlow
itself would verify."""
import os
import re
import struct
import sys

import numpy as np

TABLE_MAGIC = 0xdb4775248b80fb57
_DTYPES = {1: np.float32, 2: np.float64, 3: np.int32, 4: np.uint8, 5: np.int16, 6: np.int8, 9: np.int64, 10: np.bool_,
           17: np.uint16, 19: np.float16, 22: np.uint32, 23: np.uint64}


# ------------------------------------------------------------------------------------------------ crc32c (Castagnoli)
def crc32c(data, crc=0):
    """CRC-32C of `data` (bytes): lib/libxvkaldi.so's slicing-by-8 implementation."""
    from . import kaldi_io
    data = bytes(data)
    return int(kaldi_io.load().xvk_crc32c(data, len(data), crc))


def masked_crc32c(data):
    c = crc32c(data)
    return ((((c >> 15) | (c << 17)) & 0xFFFFFFFF) + 0xa282ead8) & 0xFFFFFFFF


# ------------------------------------------------------------------------------------------------ varints / protos
def _varint(buf, pos):
    out = shift = 0
    while True:
        b = buf[pos]
        pos += 1
        out |= (b & 0x7F) << shift
        if not b & 0x80:
            return out, pos
        shift += 7


def _proto_fields(buf):
    """Minimal protobuf wire parser: yields (field number, wire type, value)."""
    pos, n = 0, len(buf)
    while pos < n:
        tag, pos = _varint(buf, pos)
        field, wire = tag >> 3, tag & 7
        if wire == 0:
            val, pos = _varint(buf, pos)
        elif wire == 1:
            val = struct.unpack_from("<Q", buf, pos)[0]
            pos += 8
        elif wire == 2:
            ln, pos = _varint(buf, pos)
            val = bytes(buf[pos:pos + ln])
            pos += ln
        elif wire == 5:
            val = struct.unpack_from("<I", buf, pos)[0]
            pos += 4
        else:
            raise ValueError("unsupported protobuf wire type %d" % wire)
        yield field, wire, val


def _signed64(v):
    return v - (1 << 64) if v >= (1 << 63) else v


def _parse_shape(buf):
    dims = []
    for field, _, val in _proto_fields(buf):
        if field == 2:                                   # repeated Dim
            size = 0
            for f2, _, v2 in _proto_fields(val):
                if f2 == 1:
                    size = _signed64(v2)
            dims.append(size)
    return tuple(dims)


def _parse_entry(buf):
    e = {"dtype": 0, "shape": (), "shard_id": 0, "offset": 0, "size": 0, "crc32c": None, "slices": 0}
    for field, _, val in _proto_fields(buf):
        if field == 1:
            e["dtype"] = val
        elif field == 2:
            e["shape"] = _parse_shape(val)
        elif field == 3:
            e["shard_id"] = val
        elif field == 4:
            e["offset"] = _signed64(val)
        elif field == 5:
            e["size"] = _signed64(val)
        elif field == 6:
            e["crc32c"] = val
        elif field == 7:
            e["slices"] += 1
    return e


# ------------------------------------------------------------------------------------------------ table
def _read_block(blob, offset, size, verify):
    contents = blob[offset:offset + size]
    ctype = blob[offset + size]
    stored = struct.unpack_from("<I", blob, offset + size + 1)[0]
    if verify and masked_crc32c(blob[offset:offset + size + 1]) != stored:
        raise IOError("tensor-bundle index: block checksum mismatch at offset %d" % offset)
    if ctype != 0:
        raise IOError("tensor-bundle index: compressed blocks (type %d) are not supported" % ctype)
    return contents


def _block_entries(block):
    n_restarts = struct.unpack_from("<I", block, len(block) - 4)[0]
    limit = len(block) - 4 - 4 * n_restarts
    pos, key = 0, b""
    while pos < limit:
        shared, pos = _varint(block, pos)
        non_shared, pos = _varint(block, pos)
        vlen, pos = _varint(block, pos)
        key = key[:shared] + bytes(block[pos:pos + non_shared])
        pos += non_shared
        yield key, bytes(block[pos:pos + vlen])
        pos += vlen


def read_table(path, verify=True):
    """All (key, value) pairs of a LevelDB-format table file, in key order."""
    blob = open(path, "rb").read()
    if len(blob) < 48 or struct.unpack_from("<Q", blob, len(blob) - 8)[0] != TABLE_MAGIC:
        raise IOError("%s is not a tensor-bundle index (bad table magic)" % path)
    footer = blob[-48:]
    pos = 0
    _, pos = _varint(footer, pos)            # metaindex handle
    _, pos = _varint(footer, pos)
    ioff, pos = _varint(footer, pos)
    isize, pos = _varint(footer, pos)
    out = []
    for _, handle in _block_entries(_read_block(blob, ioff, isize, verify)):
        boff, p = _varint(handle, 0)
        bsize, p = _varint(handle, p)
        out.extend(_block_entries(_read_block(blob, boff, bsize, verify)))
    return out


class BundleReader(object):
    """Variables of a TF checkpoint prefix (e.g. <model>/nnet/model-1230000)."""

    def __init__(self, prefix, verify=True):
        self.prefix = prefix
        self.verify = verify
        self.entries = {}
        self.num_shards = 1
        for key, value in read_table(prefix + ".index", verify):
            if key == b"":
                for field, _, val in _proto_fields(value):
                    if field == 1:
                        self.num_shards = val
                    elif field == 2 and val != 0:
                        raise IOError("big-endian tensor bundles are not supported")
            else:
                self.entries[key.decode("utf-8")] = _parse_entry(value)

    def variables(self):
        return {k: (e["shape"], _DTYPES.get(e["dtype"])) for k, e in self.entries.items()}

    def get_tensor(self, name):
        e = self.entries[name]
        if e["slices"]:
            raise IOError("%s: sliced (partitioned) variables are not supported" % name)
        dt = _DTYPES.get(e["dtype"])
        if dt is None:
            raise IOError("%s: unsupported dtype enum %d" % (name, e["dtype"]))
        path = "%s.data-%05d-of-%05d" % (self.prefix, e["shard_id"], self.num_shards)
        with open(path, "rb") as f:
            f.seek(e["offset"])
            raw = f.read(e["size"])
        if len(raw) != e["size"]:
            raise IOError("%s: short read from %s" % (name, path))
        if self.verify and e["crc32c"] is not None and masked_crc32c(raw) != e["crc32c"]:
            raise IOError("%s: data checksum mismatch" % name)
        return np.frombuffer(raw, dt).reshape(e["shape"]).copy()


# ------------------------------------------------------------------------------------------------ checkpoint state
def get_checkpoint_state(model_dir):
    """tf.train.get_checkpoint_state: -> (model_checkpoint_path, [all_model_checkpoint_paths]) or None."""
    path = os.path.join(model_dir, "checkpoint")
    if not os.path.isfile(path):
        return None
    current, every = None, []
    for line in open(path):
        m = re.match(r'\s*(model_checkpoint_path|all_model_checkpoint_paths)\s*:\s*"(.*)"\s*$', line)
        if m and m.group(1) == "model_checkpoint_path":
            current = m.group(2)
        elif m:
            every.append(m.group(2))
    return (current, every) if current else None


def checkpoint_step(name):
    """The trailing number of a checkpoint name (model/trainer.py:289)."""
    return int(next(re.finditer(r"(\d+)(?!.*\d)", os.path.basename(name))).group(0))


def get_checkpoint(model, checkpoint="-1"):
    """misc/utils.py:251-304: select a checkpoint ("last", a step, or -1 = lowest valid_loss), rewrite the state file,
    return the checkpoint path."""
    from .misc.utils import Params
    if not os.path.isfile(os.path.join(model, "checkpoint")):
        sys.exit("[ERROR] Cannot find checkpoint in %s." % model)
    state = get_checkpoint_state(model)
    if not state:
        sys.exit("[ERROR] Cannot read checkpoint %s." % os.path.join(model, "checkpoint"))
    model_checkpoint_path, every = state
    steps = sorted(int(c.rsplit("-", 1)[1]) for c in every)
    if checkpoint == "last":
        checkpoint = steps[-1]
    else:
        checkpoint = int(checkpoint)
        if checkpoint == -1:
            min_epoch, min_loss = -1, 1e10
            with open(os.path.join(model, "valid_loss")) as f:
                for line in f.readlines():
                    epoch, loss, _ = line.split(" ")
                    if float(loss) < min_loss:
                        min_loss, min_epoch = float(loss), int(epoch)
            params = Params(os.path.join(model, "config.json"))
            checkpoint = (min_epoch + 1) * params.num_steps_per_epoch      # epochs are 0-based
    assert checkpoint in steps, "The checkpoint %d not in the model directory" % checkpoint
    model_checkpoint_path = model_checkpoint_path.rsplit("-", 1)[0] + "-" + str(checkpoint)
    model_checkpoint_path = os.path.join(model, os.path.basename(model_checkpoint_path))
    with open(os.path.join(model, "checkpoint"), "w") as f:
        f.write("model_checkpoint_path: \"%s\"\n" % model_checkpoint_path)
        for c in every:
            f.write("all_model_checkpoint_paths: \"%s\"\n" % os.path.join(model, os.path.basename(c)))
    return model_checkpoint_path


def load_weights(model_dir):
    """Variables of the current checkpoint of <model_dir> as {TF variable name: array}, optimizer slots left out."""
    state = get_checkpoint_state(model_dir)
    if not state:
        raise IOError("Failed to find a checkpoint in %s" % model_dir)
    prefix = os.path.join(model_dir, os.path.basename(state[0]))
    reader = BundleReader(prefix)
    # optimizer slot variables: the reference constructs every optimizer with name="optimizer" (model/trainer.py:234-243),
    # so its slots are "<var>/optimizer" (momentum accumulator, Adam m) and "<var>/optimizer_1" (Adam v); the TF default
    # slot names are skipped as well
    skip = ("/optimizer", "/optimizer_1", "/Momentum", "/Adam", "/Adam_1", "/ExponentialMovingAverage")
    out = {}
    for name in reader.entries:
        if name.endswith(skip) or name in ("global_step", "beta1_power", "beta2_power"):
            continue
        out[name] = reader.get_tensor(name)
    return out, checkpoint_step(state[0])


# ------------------------------------------------------------------------------------------------ writer
def _put_varint(out, v):
    v &= (1 << 64) - 1
    while v >= 0x80:
        out.append((v & 0x7F) | 0x80)
        v >>= 7
    out.append(v)


def _pb_varint_field(out, field, v):
    _put_varint(out, (field << 3) | 0)
    _put_varint(out, v)


def _pb_bytes_field(out, field, payload):
    _put_varint(out, (field << 3) | 2)
    _put_varint(out, len(payload))
    out.extend(payload)


class _BlockBuilder(object):
    """One LevelDB-format block: prefix-compressed entries, a restart point every 16 keys, the restart array."""
    RESTART_INTERVAL = 16

    def __init__(self):
        self.buf = bytearray()
        self.restarts = [0]
        self.count = 0
        self.last_key = b""

    def add(self, key, value):
        shared = 0
        if self.count % self.RESTART_INTERVAL == 0:
            if self.count:
                self.restarts.append(len(self.buf))
        else:
            n = min(len(key), len(self.last_key))
            while shared < n and key[shared] == self.last_key[shared]:
                shared += 1
        _put_varint(self.buf, shared)
        _put_varint(self.buf, len(key) - shared)
        _put_varint(self.buf, len(value))
        self.buf.extend(key[shared:])
        self.buf.extend(value)
        self.last_key = key
        self.count += 1

    def size_estimate(self):
        return len(self.buf) + 4 * len(self.restarts) + 4

    def finish(self):
        out = bytearray(self.buf)
        for r in self.restarts:
            out.extend(struct.pack("<I", r))
        out.extend(struct.pack("<I", len(self.restarts)))
        return bytes(out)


def _write_block(f, contents):
    """Appends a block + its trailer (compression type 0, masked CRC32C of contents + type); returns its handle."""
    offset = f.tell()
    f.write(contents)
    f.write(b"\x00")
    f.write(struct.pack("<I", masked_crc32c(contents + b"\x00")))
    handle = bytearray()
    _put_varint(handle, offset)
    _put_varint(handle, len(contents))
    return bytes(handle)


def write_table(path, items, block_size=4096):
    """items: (key bytes, value bytes) in ascending key order -> a LevelDB-format table file (no compression, no filter)."""
    with open(path, "wb") as f:
        index = _BlockBuilder()
        block = _BlockBuilder()
        prev = None
        for key, value in items:
            if prev is not None and not key > prev:
                raise ValueError("table keys must be strictly ascending (%r after %r)" % (key, prev))
            block.add(key, value)
            prev = key
            if block.size_estimate() >= block_size:
                index.add(block.last_key, _write_block(f, block.finish()))
                block = _BlockBuilder()
        if block.count:
            index.add(block.last_key, _write_block(f, block.finish()))
        meta = _write_block(f, _BlockBuilder().finish())
        idx = _write_block(f, index.finish())
        footer = bytearray(meta + idx)
        footer.extend(b"\x00" * (40 - len(footer)))
        footer.extend(struct.pack("<Q", TABLE_MAGIC))
        f.write(footer)


_DTYPE_ENUM = {np.dtype(v): k for k, v in _DTYPES.items()}


def save_bundle(prefix, variables):
    """Writes {name: array} as a single-shard tensor bundle <prefix>.index + <prefix>.data-00000-of-00001 -- what
    tf.train.Saver.save produces (model/trainer.py:296-303): tensors laid out in name order, little endian, every tensor
    and every index block carrying its masked CRC32C."""
    names = sorted(variables, key=lambda n: n.encode("utf-8"))
    header = bytearray()
    _pb_varint_field(header, 1, 1)                       # num_shards
    version = bytearray()
    _pb_varint_field(version, 1, 1)                      # VersionDef.producer = kTensorBundleVersion
    _pb_bytes_field(header, 3, version)
    items = [(b"", bytes(header))]
    offset = 0
    with open(prefix + ".data-00000-of-00001", "wb") as f:
        for name in names:
            a = np.asarray(variables[name])
            if a.ndim:                                   # (np.ascontiguousarray would turn a scalar into shape (1,))
                a = np.ascontiguousarray(a)
            if a.dtype not in _DTYPE_ENUM:
                raise TypeError("%s: dtype %s cannot be stored in a tensor bundle" % (name, a.dtype))
            raw = a.astype(a.dtype.newbyteorder("<"), copy=False).tobytes()
            f.write(raw)
            entry = bytearray()
            _pb_varint_field(entry, 1, _DTYPE_ENUM[a.dtype])
            shape = bytearray()
            for d in a.shape:
                dim = bytearray()
                _pb_varint_field(dim, 1, int(d))
                _pb_bytes_field(shape, 2, dim)
            _pb_bytes_field(entry, 2, shape)
            if offset:
                _pb_varint_field(entry, 4, offset)
            _pb_varint_field(entry, 5, len(raw))
            _put_varint(entry, (6 << 3) | 5)
            entry.extend(struct.pack("<I", masked_crc32c(raw)))
            items.append((name.encode("utf-8"), bytes(entry)))
            offset += len(raw)
    write_table(prefix + ".index", items)


def save_checkpoint(model_dir, step, variables, max_to_keep=None, basename="model"):
    """saver.save(sess, <model_dir>/model, global_step=step) (model/trainer.py:296-303): the bundle plus the `checkpoint`
    state file (model_checkpoint_path / all_model_checkpoint_paths, as rewritten by misc/utils.py:299-303); with
    max_to_keep (= params.keep_checkpoint_max, model/trainer.py:382) the oldest bundles beyond it are deleted."""
    os.makedirs(model_dir, exist_ok=True)
    name = "%s-%d" % (basename, int(step))
    save_bundle(os.path.join(model_dir, name), variables)
    state = get_checkpoint_state(model_dir)
    every = [os.path.basename(p) for p in state[1]] if state else []
    if name in every:
        every.remove(name)
    every.append(name)
    if max_to_keep and len(every) > max_to_keep:
        for old in every[:-max_to_keep]:
            for suffix in (".index", ".data-00000-of-00001"):
                try:
                    os.remove(os.path.join(model_dir, old + suffix))
                except OSError:
                    pass
        every = every[-max_to_keep:]
    tmp = os.path.join(model_dir, "checkpoint.tmp")
    with open(tmp, "w") as f:
        f.write('model_checkpoint_path: "%s"\n' % name)
        for c in every:
            f.write('all_model_checkpoint_paths: "%s"\n' % c)
    os.replace(tmp, os.path.join(model_dir, "checkpoint"))
    return os.path.join(model_dir, name)


def main(argv=None):
    """python -m xvector_b200.tf_checkpoint <model_dir/nnet> : writes weights.npz next to the checkpoint."""
    argv = sys.argv[1:] if argv is None else argv
    if len(argv) != 1:
        sys.exit("usage: tf_checkpoint.py <model_dir>/nnet")
    weights, step = load_weights(argv[0])
    np.savez(os.path.join(argv[0], "weights.npz"), **weights)
    print("wrote %d variables of step %d to %s" % (len(weights), step, os.path.join(argv[0], "weights.npz")))


if __name__ == "__main__":
    main()
