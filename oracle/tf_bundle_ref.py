"""Writer of TensorFlow's tensor-bundle checkpoint format.  TEST INFRASTRUCTURE: produces the
files tests feed to xvector_b200/tf_checkpoint.py.

parity unpinned: neither TensorFlow nor a saved model is available in this environment; this is an independent
statement of the PUBLISHED formats (LevelDB table: tensorflow/core/lib/io/table_builder.cc, block_builder.cc,
format.cc; tensorflow/core/util/tensor_bundle/tensor_bundle.cc; tensor_bundle.proto; the `checkpoint` text proto of
tf.train.Saver).  It deliberately shares no code with the reader (own varint / protobuf / CRC-32C, RFC 3720 known
answers checked in the test).  Like tf.train.Saver it writes one shard, uncompressed blocks, restart interval 16."""
import os
import struct

import numpy as np

DT = {np.dtype(np.float32): 1, np.dtype(np.float64): 2, np.dtype(np.int32): 3, np.dtype(np.int64): 9}


def _table():
    tab = []
    for i in range(256):
        c = i
        for _ in range(8):
            c = (c >> 1) ^ 0x82F63B78 if c & 1 else c >> 1
        tab.append(c)
    return tab


_TAB = _table()


def crc32c(data):
    crc = 0xFFFFFFFF
    tab = _TAB
    for b in bytes(data):
        crc = tab[(crc ^ b) & 0xFF] ^ (crc >> 8)
    return crc ^ 0xFFFFFFFF


def mask(crc):
    return ((((crc >> 15) | (crc << 17)) & 0xFFFFFFFF) + 0xa282ead8) & 0xFFFFFFFF


def varint(v):
    out = bytearray()
    v &= (1 << 64) - 1
    while v >= 0x80:
        out.append((v & 0x7F) | 0x80)
        v >>= 7
    out.append(v)
    return bytes(out)


def _field_varint(num, v):
    return varint(num << 3) + varint(v)


def _field_bytes(num, b):
    return varint((num << 3) | 2) + varint(len(b)) + b


def entry_proto(arr, offset):
    shape = b"".join(_field_bytes(2, _field_varint(1, d)) for d in arr.shape)
    out = _field_varint(1, DT[arr.dtype])
    out += _field_bytes(2, shape)
    if offset:
        out += _field_varint(4, offset)                      # proto3: zero-valued fields are not written
    out += _field_varint(5, arr.nbytes)
    out += varint((6 << 3) | 5) + struct.pack("<I", mask(crc32c(arr.tobytes())))
    return out


def header_proto():
    version = _field_varint(1, 1)                            # VersionDef.producer = 1
    return _field_varint(1, 1) + _field_bytes(3, version)    # num_shards = 1, endianness LITTLE (0) omitted


def build_block(items, restart_interval=16):
    buf, restarts, last = bytearray(), [], b""
    for i, (k, v) in enumerate(items):
        shared = 0
        if i % restart_interval == 0:
            restarts.append(len(buf))
        else:
            while shared < min(len(last), len(k)) and last[shared] == k[shared]:
                shared += 1
        buf += varint(shared) + varint(len(k) - shared) + varint(len(v)) + k[shared:] + v
        last = k
    if not restarts:
        restarts = [0]
    for r in restarts:
        buf += struct.pack("<I", r)
    buf += struct.pack("<I", len(restarts))
    return bytes(buf)


def write_table(path, items, block_entries=40):
    """items: sorted [(key bytes, value bytes)].  Several data blocks once there are more than block_entries keys."""
    out = bytearray()
    index = []

    def emit(block):
        off = len(out)
        out.extend(block)
        out.append(0)                                        # kNoCompression
        out.extend(struct.pack("<I", mask(crc32c(block + b"\0"))))
        return varint(off) + varint(len(block))
    for i in range(0, len(items), block_entries):
        chunk = items[i:i + block_entries]
        handle = emit(build_block(chunk))
        index.append((chunk[-1][0] + (b"" if i + block_entries >= len(items) else b"\x00"), handle))   # separator >= last key
    meta = emit(build_block([]))
    idx = emit(build_block(index, restart_interval=1))
    footer = meta + idx
    footer += b"\0" * (40 - len(footer))
    footer += struct.pack("<Q", 0xdb4775248b80fb57)
    out.extend(footer)
    with open(path, "wb") as f:
        f.write(bytes(out))


def write_checkpoint(model_dir, step, variables, keep=()):
    """<model_dir>/model-<step>.{index,data-00000-of-00001} + the `checkpoint` state file (tf.train.Saver layout)."""
    prefix = os.path.join(model_dir, "model-%d" % step)
    items, data, offset = [(b"", header_proto())], bytearray(), 0
    for name in sorted(variables):
        arr = np.asarray(variables[name])
        arr = arr if arr.flags.c_contiguous else arr.copy()       # (np.ascontiguousarray would turn a scalar into rank 1)
        items.append((name.encode("utf-8"), entry_proto(arr, offset)))
        data += arr.tobytes()
        offset += arr.nbytes
    write_table(prefix + ".index", items)
    with open(prefix + ".data-00000-of-00001", "wb") as f:
        f.write(bytes(data))
    with open(os.path.join(model_dir, "checkpoint"), "w") as f:
        f.write('model_checkpoint_path: "%s"\n' % prefix)
        for s in list(keep) + [step]:
            f.write('all_model_checkpoint_paths: "%s"\n' % os.path.join(model_dir, "model-%d" % s))
    return prefix
