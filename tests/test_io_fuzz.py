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
