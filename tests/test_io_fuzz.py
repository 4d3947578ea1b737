"""Corrupted inputs never crash the native readers: truncated / bit-flipped / padded checkpoint shards, index files,
TFRecord files and serialized Examples either parse or raise a Python exception (crc / framing / bounds checks of
io/csrc/dtf_io.cc)."""
import os
import random

import numpy as np

from distributed_tensorflow_framework_b200.io import tensorbundle as tb, tfrecord as tfr


def _mutate(rng, blob):
    data = bytearray(blob)
    mode = rng.randrange(3)
    if mode == 0 and len(data) > 1:
        data = data[:rng.randrange(len(data))]
    elif mode == 1:
        for _ in range(rng.randrange(1, 6)):
            data[rng.randrange(len(data))] = rng.randrange(256)
    else:
        pos = rng.randrange(len(data))
        data[pos:pos] = bytes(rng.randrange(256) for _ in range(rng.randrange(1, 40)))
    return bytes(data)


def test_corrupted_checkpoints_raise_instead_of_crashing(tmp_path):
    prefix = str(tmp_path / "model.ckpt-1")
    tensors = {"a/weights": np.random.rand(5, 5, 3, 8).astype(np.float32), "a/biases": np.zeros(8, np.float32),
               "global_step": np.array(1, np.int64), "b/w": np.random.rand(100, 7).astype(np.float32)}
    tb.write_bundle(prefix, tensors, shard_of={k: i % 2 for i, k in enumerate(tensors)}, num_shards=2)
    files = sorted(os.listdir(tmp_path))
    assert len(files) == 3
    orig = {f: (tmp_path / f).read_bytes() for f in files}
    rng = random.Random(0)
    ok = bad = 0
    for _ in range(200):
        f = rng.choice(files)
        (tmp_path / f).write_bytes(_mutate(rng, orig[f]))
        try:
            with tb.BundleReader(prefix) as r:
                for k in r.keys():
                    r.get_tensor(k)
            ok += 1
        except Exception:
            bad += 1
        (tmp_path / f).write_bytes(orig[f])
    assert bad > 150 and ok + bad == 200  # nearly every corruption is detected (crc32c), none kills the process
    with tb.BundleReader(prefix) as r:  # and the intact files still read back exactly
        assert np.array_equal(r.get_tensor("b/w"), tensors["b/w"])


def test_corrupted_tfrecords_and_examples_raise_instead_of_crashing(tmp_path):
    p = tmp_path / "x.tfrecord"
    with tfr.TFRecordWriter(str(p)) as w:
        for i in range(20):
            w.write(tfr.make_example({"img": bytes(range(50)), "label": [i], "f": [0.5, 1.5]}))
    orig = p.read_bytes()
    feats = {"img": tfr.FixedLenFeature([], "string"), "label": tfr.FixedLenFeature([], "int64"),
             "f": tfr.FixedLenFeature([2], "float32")}
    rng = random.Random(1)
    bad = 0
    for _ in range(200):
        p.write_bytes(_mutate(rng, orig))
        try:
            for rec in tfr.tf_record_iterator(str(p)):
                tfr.parse_single_example(rec, feats)
        except Exception:
            bad += 1
    assert bad > 150
    example = tfr.make_example({"img": bytes(range(50)), "label": [3], "f": [0.5, 1.5]})
    for _ in range(1000):
        try:
            tfr.parse_single_example(_mutate(rng, example), feats)
        except Exception:
            pass
    assert int(np.asarray(tfr.parse_single_example(example, feats)["label"]).reshape(-1)[0]) == 3


from hypothesis import given, settings, strategies as st  # noqa: E402

_names = st.text(alphabet="abcdefghij/_0123456789ABC", min_size=1, max_size=60).filter(
    lambda s: not s.startswith("/") and not s.endswith("/"))
_dtypes = st.sampled_from(["float32", "int32", "int64", "float16", "uint8", "float64"])
_shapes = st.lists(st.integers(0, 7), min_size=0, max_size=4)


@settings(max_examples=40, deadline=None)
@given(items=st.dictionaries(_names, st.tuples(_dtypes, _shapes, st.integers(0, 2 ** 31 - 1)), min_size=1, max_size=30),
       num_shards=st.integers(1, 4))
def test_checkpoint_round_trip_for_random_names_shapes_dtypes_and_shards(tmp_path_factory, items, num_shards):
    d = tmp_path_factory.mktemp("bundle")
    prefix = str(d / "model.ckpt-7")
    tensors, shard_of = {}, {}
    for i, (name, (dt, shape, seed)) in enumerate(sorted(items.items())):
        rng = np.random.RandomState(seed)
        tensors[name] = (rng.rand(*shape) * 100).astype(dt) if shape else np.array(rng.rand() * 100).astype(dt)
        shard_of[name] = i % num_shards
    tb.write_bundle(prefix, tensors, shard_of=shard_of, num_shards=num_shards)
    with tb.BundleReader(prefix) as r:
        assert r.num_shards == num_shards and set(r.keys()) == set(tensors)
        for name, ref in tensors.items():
            got = r.get_tensor(name)
            assert got.dtype == ref.dtype and tuple(got.shape) == tuple(ref.shape), name
            assert np.array_equal(got, ref), name


def _py_crc32c(data):
    crc = 0xffffffff
    for b in data:
        crc ^= b
        for _ in range(8):
            crc = (crc >> 1) ^ (0x82f63b78 if crc & 1 else 0)
    return crc ^ 0xffffffff


@settings(max_examples=80, deadline=None)
@given(data=st.binary(min_size=0, max_size=700))
def test_crc32c_matches_a_bitwise_reference_for_random_buffers(data):
    from distributed_tensorflow_framework_b200.io import _iolib
    assert _iolib.lib().dtf_crc32c(data, len(data)) == _py_crc32c(data)


_feat_name = st.text(alphabet="abcdefgh_/XYZ019", min_size=1, max_size=24)
_feat_value = st.one_of(
    st.lists(st.binary(min_size=0, max_size=40), min_size=1, max_size=4),
    st.lists(st.integers(-2 ** 63, 2 ** 63 - 1), min_size=1, max_size=9),
    st.lists(st.floats(width=32, allow_nan=False, allow_infinity=False), min_size=1, max_size=9))


@settings(max_examples=80, deadline=None)
@given(features=st.dictionaries(_feat_name, _feat_value, min_size=1, max_size=8))
def test_example_round_trip_for_random_features(tmp_path_factory, features):
    """make_example -> TFRecord file -> C++ reader -> C++ Example parser gives back every value (bytes lists,
    negative / 64-bit integers as varints, packed floats), whatever the feature names and their order."""
    path = str(tmp_path_factory.mktemp("rec") / "x.tfrecord")
    with tfr.TFRecordWriter(path) as w:
        w.write(tfr.make_example(features))
        w.write(tfr.make_example(features))
    spec = {}
    for name, v in features.items():
        kind = "string" if isinstance(v[0], bytes) else ("float32" if isinstance(v[0], float) else "int64")
        spec[name] = tfr.FixedLenFeature([len(v)], kind)
    n = 0
    for rec in tfr.tf_record_iterator(path):
        got = tfr.parse_single_example(rec, spec)
        n += 1
        for name, v in features.items():
            if isinstance(v[0], bytes):
                assert list(got[name]) == v, name
            elif isinstance(v[0], float):
                assert np.array_equal(got[name], np.array(v, dtype=np.float32)), name
            else:
                assert got[name].tolist() == v, name
    assert n == 2
