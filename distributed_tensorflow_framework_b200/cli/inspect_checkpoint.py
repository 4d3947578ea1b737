"""List or print the tensors of a TF1-layout checkpoint (what `inspect_checkpoint` is to tf.train.Saver files).

  python -m distributed_tensorflow_framework_b200.cli.inspect_checkpoint <model_dir | prefix> [--tensor_name N]
                                                                            [--all_tensors] [--events]

`<model_dir>`: the latest checkpoint named by its `checkpoint` state file. `--events` also prints the scalars of the
TensorBoard event files found next to it.
"""
import argparse
import os
import sys

import numpy as np

from ..io import events, tensorbundle as tb

_DT_NAMES = {tb.DT_FLOAT: "float32", tb.DT_INT32: "int32", tb.DT_INT64: "int64", tb.DT_BFLOAT16: "bfloat16",
             tb.DT_HALF: "float16", tb.DT_UINT8: "uint8", tb.DT_DOUBLE: "float64"}


def _bf16_to_f32(bits):
    return (bits.astype(np.uint32) << 16).view(np.float32)


def describe(prefix, tensor_name=None, all_tensors=False, out=sys.stdout):
    total = 0
    with tb.BundleReader(prefix) as r:
        print("checkpoint %s (%d data shard%s)" % (prefix, r.num_shards, "" if r.num_shards == 1 else "s"), file=out)
        for name in r.keys():
            dt, shape, nbytes, shard = r.entries[name]
            if tensor_name is not None and name != tensor_name:
                continue
            total += nbytes
            print("%-60s %-9s %-20s shard %d  %d bytes" % (name, _DT_NAMES.get(dt, "dtype %d" % dt),
                                                           list(shape), shard, nbytes), file=out)
            if all_tensors or tensor_name is not None:
                v = r.get_tensor(name)
                if dt == tb.DT_BFLOAT16:
                    v = _bf16_to_f32(v)
                print(np.array2string(v, threshold=40, precision=6), file=out)
        if tensor_name is not None and total == 0 and tensor_name not in r.entries:
            raise KeyError("tensor %r is not in the checkpoint" % tensor_name)
    print("total: %.2f MB" % (total / 1e6), file=out)
    return total


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("path", help="model directory or checkpoint prefix (model.ckpt-N)")
    ap.add_argument("--tensor_name", default=None)
    ap.add_argument("--all_tensors", action="store_true")
    ap.add_argument("--events", action="store_true", help="also dump the scalar summaries of the event files")
    a = ap.parse_args(argv)
    prefix, model_dir = a.path, os.path.dirname(os.path.abspath(a.path))
    if os.path.isdir(a.path):
        model_dir = a.path
        prefix = tb.latest_checkpoint(a.path)
        if prefix is None:
            print("no checkpoint state file in %s" % a.path, file=sys.stderr)
            return 1
    describe(prefix, a.tensor_name, a.all_tensors)
    if a.events:
        for root, _, files in os.walk(model_dir):
            for f in sorted(files):
                if f.startswith("events.out.tfevents"):
                    print("event file %s" % os.path.join(root, f))
                    for step, tag, value in events.read_scalars(os.path.join(root, f)):
                        print("  step %-8d %-30s %.6g" % (step, tag, value))
    return 0


if __name__ == "__main__":
    sys.exit(main())
