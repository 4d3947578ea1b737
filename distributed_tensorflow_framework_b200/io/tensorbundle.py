"""TF1 Saver V2 / TensorBundle checkpoints on the native reader/writer (no TensorFlow).

Layout: `<prefix>.index` (LevelDB-format SSTable: key "" -> BundleHeaderProto, tensor name ->
BundleEntryProto) + `<prefix>.data-XXXXX-of-YYYYY` raw little-endian tensor bytes, plus the `checkpoint` state file.
The reference gets this from MonitoredTrainingSession(checkpoint_dir=model_dir) (/root/reference/distribute_train.py:169-175).
"""
import ctypes
import os
import re

import numpy as np

from ._iolib import lib

# tensorflow DataType enum values
DT_FLOAT, DT_INT32, DT_INT64, DT_BFLOAT16, DT_HALF, DT_UINT8, DT_DOUBLE = 1, 3, 9, 14, 19, 4, 2
_NP2DT = {np.dtype("float32"): DT_FLOAT, np.dtype("int32"): DT_INT32, np.dtype("int64"): DT_INT64,
          np.dtype("float16"): DT_HALF, np.dtype("uint8"): DT_UINT8, np.dtype("float64"): DT_DOUBLE}
_DT2NP = {v: k for k, v in _NP2DT.items()}
_DT2NP[DT_BFLOAT16] = np.dtype("uint16")  # raw bits; callers reinterpret


def write_bundle(prefix, tensors, shard_of=None, num_shards=1):
    """tensors: name -> np.ndarray (or (np.uint16 array, 'bfloat16')). shard_of: name -> shard id."""
    L = lib()
    os.makedirs(os.path.dirname(os.path.abspath(prefix)), exist_ok=True)
    h = L.dtf_bundle_writer_open(prefix.encode(), num_shards)
    for name in sorted(tensors):
        arr = tensors[name]
        dt = None
        if isinstance(arr, tuple):
            arr, tag = arr
            dt = DT_BFLOAT16 if tag == "bfloat16" else None
        arr = np.require(arr, requirements="C")  # (ascontiguousarray would turn a scalar into shape [1])
        if dt is None:
            dt = _NP2DT[arr.dtype]
        shape = (ctypes.c_longlong * max(arr.ndim, 1))(*arr.shape)
        shard = 0 if shard_of is None else int(shard_of.get(name, 0))
        rc = L.dtf_bundle_writer_add(h, name.encode(), dt, shape, arr.ndim, arr.ctypes.data, arr.nbytes, shard)
        if rc != 0:
            raise IOError("bundle write failed for %s" % name)
    if L.dtf_bundle_writer_finish(h) != 0:
        raise IOError("bundle finish failed for %s" % prefix)


class BundleReader(object):
    def __init__(self, prefix):
        self._h = lib().dtf_bundle_reader_open(prefix.encode())
        if not self._h:
            raise IOError("cannot open tensor bundle %s" % prefix)
        self.entries = {}
        L = lib()
        name = ctypes.create_string_buffer(4096)
        dtype, rank, shard = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        nbytes = ctypes.c_longlong()
        shape = (ctypes.c_longlong * 8)()
        for i in range(L.dtf_bundle_reader_num_tensors(self._h)):
            rc = L.dtf_bundle_reader_entry(self._h, i, name, 4096, ctypes.byref(dtype), shape, ctypes.byref(rank),
                                           ctypes.byref(nbytes), ctypes.byref(shard))
            if rc != 0:
                raise IOError("bad bundle entry %d" % i)
            self.entries[name.value.decode()] = (dtype.value, tuple(shape[d] for d in range(rank.value)),
                                                 nbytes.value, shard.value)
        self.num_shards = L.dtf_bundle_reader_num_shards(self._h)

    def keys(self):
        return sorted(self.entries)

    def has_tensor(self, name):
        return name in self.entries

    def get_tensor(self, name):
        dt, shape, nbytes, _ = self.entries[name]
        arr = np.empty(shape, dtype=_DT2NP[dt])
        rc = lib().dtf_bundle_reader_read(self._h, name.encode(), arr.ctypes.data, nbytes)
        if rc != 0:
            raise IOError("reading %s failed (code %d: -4 = crc mismatch)" % (name, rc))
        return arr

    def dtype_of(self, name):
        return self.entries[name][0]

    def close(self):
        if self._h:
            lib().dtf_bundle_reader_close(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


# ---- `checkpoint` state file (text-format CheckpointState)
def read_checkpoint_state(model_dir):
    path = os.path.join(model_dir, "checkpoint")
    if not os.path.exists(path):
        return None, []
    latest, all_paths = None, []
    for line in open(path):
        m = re.match(r'\s*(model_checkpoint_path|all_model_checkpoint_paths)\s*:\s*"(.*)"', line)
        if not m:
            continue
        if m.group(1) == "model_checkpoint_path":
            latest = m.group(2)
        else:
            all_paths.append(m.group(2))
    return latest, all_paths


def write_checkpoint_state(model_dir, latest, all_paths):
    tmp = os.path.join(model_dir, "checkpoint.tmp")
    with open(tmp, "w") as f:
        f.write('model_checkpoint_path: "%s"\n' % latest)
        for p in all_paths:
            f.write('all_model_checkpoint_paths: "%s"\n' % p)
    os.replace(tmp, os.path.join(model_dir, "checkpoint"))


def latest_checkpoint(model_dir):
    latest, _ = read_checkpoint_state(model_dir)
    if latest is None:
        return None
    prefix = latest if os.path.isabs(latest) else os.path.join(model_dir, latest)
    return prefix if os.path.exists(prefix + ".index") else None
