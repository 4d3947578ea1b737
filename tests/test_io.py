"""Format tests for the native I/O library: crc32c vectors, TFRecord framing, Example decode, TensorBundle round-trip
checked by an independent pure-Python SSTable parser, checkpoint state file, event files."""
import os
import struct

import numpy as np
import pytest

from distributed_tensorflow_framework_b200.io import _iolib, events, tensorbundle as tb, tfrecord as tfr


def test_crc32c_known_vectors():
    L = _iolib.lib()
    assert L.dtf_crc32c(b"123456789", 9) == 0xE3069283
    assert L.dtf_crc32c(bytes(32), 32) == 0x8A9136AA          # RFC 3720 B.4
    assert L.dtf_crc32c(bytes([0xff] * 32), 32) == 0x62A8AB43
    assert L.dtf_crc32c(bytes(range(32)), 32) == 0x46DD794E
    crc = L.dtf_crc32c(b"foo", 3)
    masked = L.dtf_crc32c_masked(b"foo", 3)
    assert masked == (((crc >> 15) | (crc << 17)) + 0xa282ead8) & 0xffffffff


def _py_masked_crc(data):
    # independent bitwise implementation
    crc = 0xffffffff
    for b in data:
        crc ^= b
        for _ in range(8):
            crc = (crc >> 1) ^ (0x82f63b78 if crc & 1 else 0)
    crc ^= 0xffffffff
    return (((crc >> 15) | (crc << 17)) + 0xa282ead8) & 0xffffffff


def test_tfrecord_framing_and_example(tmp_path):
    path = str(tmp_path / "data.tfrecord")
    img = np.arange(4 * 4 * 3, dtype=np.uint8).tobytes()
    with tfr.TFRecordWriter(path) as w:
        for i in range(5):
            w.write(tfr.make_example({"image_raw": img, "label": [i], "h": [4], "w": [4], "score": [0.5, float(i)]}))
    # byte-level check of the first record with a struct-based reader
    raw = open(path, "rb").read()
    (length,) = struct.unpack("<Q", raw[:8])
    assert struct.unpack("<I", raw[8:12])[0] == _py_masked_crc(raw[:8])
    payload = raw[12:12 + length]
    assert struct.unpack("<I", raw[12 + length:16 + length])[0] == _py_masked_crc(payload)
    recs = list(tfr.tf_record_iterator(path))
    assert len(recs) == 5 and recs[0] == payload
    feats = {"image_raw": tfr.FixedLenFeature([], "string"), "label": tfr.FixedLenFeature([], "int64"),
             "score": tfr.FixedLenFeature([2], "float32")}
    for i, r in enumerate(recs):
        ex = tfr.parse_single_example(r, feats)
        assert ex["image_raw"] == img and ex["label"] == i
        np.testing.assert_allclose(ex["score"], [0.5, i])
    with pytest.raises(ValueError):
        tfr.parse_single_example(recs[0], {"missing": tfr.FixedLenFeature([], "int64")})
    # corruption is detected
    bad = bytearray(raw)
    bad[20] ^= 0xff
    open(path, "wb").write(bytes(bad))
    with pytest.raises(IOError):
        list(tfr.tf_record_iterator(path))


def test_shuffling_queue(tmp_path):
    paths = []
    for f in range(2):
        p = str(tmp_path / ("f%d.tfrecord" % f))
        with tfr.TFRecordWriter(p) as w:
            for i in range(50):
                w.write(struct.pack("<i", f * 50 + i))
        paths.append(p)
    q = tfr.ShufflingRecordQueue(paths, capacity=32, shuffle=True, loop=False, seed=3)
    got = []
    while True:
        try:
            got.append(struct.unpack("<i", q.get())[0])
        except StopIteration:
            break
    q.close()
    assert sorted(got) == list(range(100)) and got != list(range(100))
    q = tfr.ShufflingRecordQueue(paths[:1], capacity=8, shuffle=False, loop=True)
    got = [struct.unpack("<i", q.get())[0] for _ in range(120)]
    q.close()
    assert got[:50] == list(range(50)) and got[50:100] == list(range(50))


# ---- independent pure-Python SSTable / bundle parser (never touches the C++ reader)
def _varint(b, i):
    v = s = 0
    while True:
        c = b[i]
        i += 1
        v |= (c & 0x7f) << s
        s += 7
        if not c & 0x80:
            return v, i


def _parse_block(b):
    (nrestarts,) = struct.unpack("<I", b[-4:])
    end = len(b) - 4 - 4 * nrestarts
    i, key, out = 0, b"", []
    while i < end:
        shared, i = _varint(b, i)
        non_shared, i = _varint(b, i)
        vlen, i = _varint(b, i)
        key = key[:shared] + b[i:i + non_shared]
        i += non_shared
        out.append((key, b[i:i + vlen]))
        i += vlen
    return out


def _py_read_index(path):
    data = open(path, "rb").read()
    assert struct.unpack("<Q", data[-8:])[0] == 0xdb4775248b80fb57
    footer = data[-48:-8]
    i = 0
    _, i = _varint(footer, i)
    _, i = _varint(footer, i)
    ioff, i = _varint(footer, i)
    isize, i = _varint(footer, i)
    out = []
    for _, handle in _parse_block(data[ioff:ioff + isize]):
        off, j = _varint(handle, 0)
        size, j = _varint(handle, j)
        assert data[off + size] == 0  # uncompressed
        assert struct.unpack("<I", data[off + size + 1:off + size + 5])[0] == _py_masked_crc(data[off:off + size + 1])
        out += _parse_block(data[off:off + size])
    return out


def test_tensorbundle_roundtrip_and_independent_parse(tmp_path):
    rng = np.random.RandomState(0)
    tensors = {
        "conv1/weights": rng.randn(5, 5, 3, 64).astype(np.float32),
        "conv1/biases": np.zeros(64, np.float32),
        "global_step": np.array(1234, dtype=np.int64),
        "local3/weights/Adam": rng.randn(17, 9).astype(np.float32),
        "bf16w": (rng.randint(0, 65535, (4, 8)).astype(np.uint16), "bfloat16"),
    }
    for i in range(40):  # enough keys to exercise prefix compression and restart points
        tensors["layer_%02d/weights" % i] = rng.randn(3, i + 1).astype(np.float32)
    prefix = str(tmp_path / "model.ckpt-1234")
    shard_of = {k: (hash(k) % 2) for k in tensors}
    tb.write_bundle(prefix, tensors, shard_of=shard_of, num_shards=2)
    assert os.path.exists(prefix + ".index")
    assert os.path.exists(prefix + ".data-00000-of-00002") and os.path.exists(prefix + ".data-00001-of-00002")
    with tb.BundleReader(prefix) as r:
        assert r.num_shards == 2
        assert set(r.keys()) == set(tensors)
        for k, v in tensors.items():
            ref = v[0] if isinstance(v, tuple) else v
            np.testing.assert_array_equal(r.get_tensor(k), ref)
        assert r.dtype_of("bf16w") == tb.DT_BFLOAT16 and r.dtype_of("global_step") == tb.DT_INT64
        assert r.get_tensor("global_step").shape == ()  # a scalar, as tf.train.Saver writes global_step
    # independent parse: keys sorted bytewise, header first, entries decode to the right offsets
    kv = _py_read_index(prefix + ".index")
    keys = [k for k, _ in kv]
    assert keys == sorted(keys) and keys[0] == b""
    assert kv[0][1][:2] == bytes([0x08, 0x02])  # BundleHeaderProto.num_shards = 2
    entry = dict(kv)[b"conv1/biases"]
    # field 1 (dtype) = DT_FLOAT, field 5 (size) = 256 bytes, field 6 fixed32 crc
    assert entry[:2] == bytes([0x08, 0x01])
    assert bytes([0x28]) + bytes([0x80, 0x02]) in entry
    crc_pos = entry.index(bytes([0x35]))
    shard = shard_of["conv1/biases"]
    blob = open(prefix + ".data-%05d-of-00002" % shard, "rb").read()
    # locate via offset field (4) if present
    off = 0
    if bytes([0x20]) in entry[:crc_pos]:
        j = entry.index(bytes([0x20]))
        off, _ = _varint(entry, j + 1)
    assert struct.unpack("<I", entry[crc_pos + 1:crc_pos + 5])[0] == _py_masked_crc(blob[off:off + 256])
    # corruption of a data shard is detected by the crc
    with open(prefix + ".data-%05d-of-00002" % shard, "r+b") as f:
        f.seek(off)
        f.write(b"\x01")
    with tb.BundleReader(prefix) as r:
        with pytest.raises(IOError):
            r.get_tensor("conv1/biases")


def test_checkpoint_state_file(tmp_path):
    d = str(tmp_path)
    assert tb.latest_checkpoint(d) is None
    tb.write_bundle(os.path.join(d, "model.ckpt-10"), {"a": np.ones(3, np.float32)})
    tb.write_checkpoint_state(d, "model.ckpt-10", ["model.ckpt-5", "model.ckpt-10"])
    text = open(os.path.join(d, "checkpoint")).read()
    assert text.splitlines()[0] == 'model_checkpoint_path: "model.ckpt-10"'
    latest, allp = tb.read_checkpoint_state(d)
    assert latest == "model.ckpt-10" and allp == ["model.ckpt-5", "model.ckpt-10"]
    assert tb.latest_checkpoint(d) == os.path.join(d, "model.ckpt-10")


def test_event_writer(tmp_path):
    w = events.EventFileWriter(str(tmp_path))
    for s in range(3):
        w.add_scalar("total_loss", 1.0 / (s + 1), s)
    w.add_scalar("cross_entropy", 0.25, 2)
    w.close()
    got = events.read_scalars(w.path)
    assert [g[0] for g in got] == [0, 1, 2, 2]
    assert got[3][1] == "cross_entropy" and abs(got[1][2] - 0.5) < 1e-7
    try:
        from tensorboard.backend.event_processing import event_file_loader
    except Exception:
        return
    evs = list(event_file_loader.LegacyEventFileLoader(w.path).Load())
    assert evs[0].file_version == "brain.Event:2"
    vals = [(e.step, e.summary.value[0].tag, e.summary.value[0].simple_value) for e in evs[1:]]
    assert vals[0][1] == "total_loss" and abs(vals[2][2] - 1 / 3) < 1e-6
