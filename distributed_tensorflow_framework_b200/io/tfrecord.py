"""TFRecord files and tf.train.Example records on the native reader (no TensorFlow).

Parity: TFRecordReader + parse_single_example used at /root/reference/distribute_input.py:92-99. Feature specs mirror
tf.FixedLenFeature(shape, dtype) as used by the reference entry script (/root/reference/distribute.py:30-35).
"""
import ctypes
import struct
from collections import namedtuple

import numpy as np

from ._iolib import lib

FixedLenFeature = namedtuple("FixedLenFeature", ["shape", "dtype"])
STRING, INT64, FLOAT32 = "string", "int64", "float32"


class TFRecordWriter(object):
    def __init__(self, path):
        self._h = lib().dtf_tfrecord_writer_open(path.encode())
        if not self._h:
            raise IOError("cannot open %s for writing" % path)

    def write(self, record: bytes):
        if lib().dtf_tfrecord_write(self._h, record, len(record)) != 0:
            raise IOError("TFRecord write failed")

    def close(self):
        if self._h:
            lib().dtf_tfrecord_writer_close(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def tf_record_iterator(path, check_crc=True):
    """Yield the raw records of one TFRecord file."""
    h = lib().dtf_tfrecord_open(path.encode(), 1 if check_crc else 0)
    if not h:
        raise IOError("cannot open %s" % path)
    data = ctypes.c_void_p()
    n = ctypes.c_ulonglong()
    try:
        while True:
            rc = lib().dtf_tfrecord_next(h, ctypes.byref(data), ctypes.byref(n))
            if rc == 0:
                return
            if rc < 0:
                raise IOError("corrupt TFRecord %s (code %d)" % (path, rc))
            yield ctypes.string_at(data.value, n.value) if n.value else b""
    finally:
        lib().dtf_tfrecord_close(h)


# ---- tf.train.Example encoding (pure Python; writers are not on a hot path)
def _varint(v):
    out = bytearray()
    v &= (1 << 64) - 1
    while v >= 0x80:
        out.append((v & 0x7f) | 0x80)
        v >>= 7
    out.append(v)
    return bytes(out)


def _len_field(field, payload):
    return _varint((field << 3) | 2) + _varint(len(payload)) + payload


def make_example(features: dict) -> bytes:
    """features: name -> bytes | list[bytes] | list[int] | list[float] | np.ndarray (int64 / float32)."""
    entries = b""
    for name in sorted(features):
        v = features[name]
        if isinstance(v, (bytes, bytearray)):
            v = [bytes(v)]
        if isinstance(v, np.ndarray):
            v = v.reshape(-1).tolist()
        if not isinstance(v, (list, tuple)):
            v = [v]
        if len(v) and isinstance(v[0], (bytes, bytearray)):
            feat = _len_field(1, b"".join(_len_field(1, bytes(b)) for b in v))
        elif len(v) and isinstance(v[0], float):
            feat = _len_field(2, _len_field(1, struct.pack("<%df" % len(v), *v)))
        else:
            feat = _len_field(3, _len_field(1, b"".join(_varint(int(i)) for i in v)))
        entry = _len_field(1, name.encode()) + _len_field(2, feat)
        entries += _len_field(1, entry)
    return _len_field(1, entries)


def parse_single_example(serialized: bytes, features: dict) -> dict:
    """name -> FixedLenFeature(shape, dtype) ==> name -> bytes (scalar string) | np.ndarray."""
    L = lib()
    out = {}
    n = len(serialized)
    for name, spec in features.items():
        shape, dtype = tuple(spec[0]), spec[1]
        dtype = getattr(dtype, "name", dtype)
        count = int(np.prod(shape)) if len(shape) else 1
        key = name.encode()
        if dtype in (STRING, "bytes", bytes, str):
            vals = []
            off, ln = ctypes.c_ulonglong(), ctypes.c_ulonglong()
            for i in range(count):
                if L.dtf_example_bytes(serialized, n, key, i, ctypes.byref(off), ctypes.byref(ln)) != 0:
                    raise ValueError("feature %r: missing bytes value %d" % (name, i))
                vals.append(serialized[off.value:off.value + ln.value])
            out[name] = vals[0] if not shape else np.array(vals, dtype=object).reshape(shape)
        elif dtype in (INT64, "int"):
            buf = np.empty(count, dtype=np.int64)
            got = L.dtf_example_int64s(serialized, n, key, buf.ctypes.data, count)
            if got != count:
                raise ValueError("feature %r: expected %d int64 values, found %d" % (name, count, got))
            out[name] = buf.reshape(shape) if shape else buf[0]
        elif dtype in (FLOAT32, "float"):
            buf = np.empty(count, dtype=np.float32)
            got = L.dtf_example_floats(serialized, n, key, buf.ctypes.data, count)
            if got != count:
                raise ValueError("feature %r: expected %d float values, found %d" % (name, count, got))
            out[name] = buf.reshape(shape) if shape else buf[0]
        else:
            raise TypeError("unsupported feature dtype %r" % (dtype,))
    return out


class ShufflingRecordQueue(object):
    """Background-thread record reader with a shuffle pool (tf.train.shuffle_batch's queue,
    /root/reference/distribute_input.py:54-69): `capacity` records are buffered and dequeued at random
    (min_after_dequeue = 0 like the reference)."""

    def __init__(self, paths, capacity, shuffle=True, loop=True, seed=0):
        if isinstance(paths, str):
            paths = [paths]
        self._h = lib().dtf_prefetch_open("\n".join(paths).encode(), 1 if shuffle else 0, 1 if shuffle else 0,
                                          1 if loop else 0, int(max(capacity, 1)), int(seed))
        self._data = ctypes.c_void_p()
        self._n = ctypes.c_ulonglong()

    def get(self):
        rc = lib().dtf_prefetch_next(self._h, ctypes.byref(self._data), ctypes.byref(self._n))
        if rc == 0:
            raise StopIteration
        if rc < 0:
            raise IOError("record prefetch failed (code %d)" % rc)
        return ctypes.string_at(self._data.value, self._n.value)

    def close(self):
        if self._h:
            lib().dtf_prefetch_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
