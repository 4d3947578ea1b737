"""Command-line flags with the reference's names and defaults.

Parity: /root/reference/distribute_flags.py:9-62 (tf.app.flags). Implemented on argparse so that it has no TF / absl
global state; `FLAGS.<name>` attribute access, `parse(argv)` and `app_run(main)` (tf.app.run equivalent,
/root/reference/distribute.py:135-136) are provided. Quirk fixes: the two thread flags had a trailing space in their
name; `data_load_option` used by DistributeExperiment was never defined (Q13). Extensions:
`sync_replicas` / `sync_mode` select sync, backup-worker or async parameter-server mode.
"""
from __future__ import annotations

import argparse
import sys
from typing import Any, Callable, Dict, List, Optional


def _str2bool(v):
    if isinstance(v, bool):
        return v
    return str(v).lower() in ("1", "true", "t", "yes", "y")


_DEFS = [
    # name, type, default, help                                         (reference line)
    ("use_fp16", _str2bool, False, "Train the model using fp16 variables (accepted for compatibility: the CUDA path "
     "always computes in bf16 over fp32 master weights, the CPU oracle path stays fp32)."),  # :10
    ("project_name", str, "distribute", "Name of the project."),  # :14
    ("job_name", str, "", "One of 'ps', 'worker'."),  # :16
    ("ps_hosts", str, "", "Comma-separated list of hostname:port pairs."),  # :18
    ("worker_hosts", str, "", "Comma-separated list of hostname:port pairs."),  # :20
    ("task_index", int, None, "Index of task within the job."),  # :22
    ("replicas_to_aggregate", int, None,
     "Number of replicas to aggregate before the parameter update (default: num_workers)."),  # :26
    ("intra_op_parallelism_threads", int, 0, "Threads for intra-op parallelism (host side)."),  # :32
    ("inter_op_parallelism_threads", int, 0, "Threads for inter-op parallelism (host side)."),  # :38
    ("log_device_placement", _str2bool, False, "Whether to log device placement."),  # :45
    ("input_image_height", int, 224, "Input image height."),  # :51
    ("input_image_width", int, 224, "Input image width."),  # :53
    ("sample_number", int, 100000, "Total sample number to train."),  # :55
    ("train_learning_rate", float, 0.001, "Initial learning rate."),  # :57
    ("learning_rate_json", str, "./learning_rate.json", "Path of the persisted learning rate."),  # :61
    # ---- defined here because the reference reads it without defining it (distribute_experiment.py:30)
    ("data_load_option", str, "tfrecord", "One of tfrecord | placeholder | datapath."),
    # ---- extensions
    ("sync_replicas", _str2bool, True, "Synchronous replicas (False = asynchronous parameter-server mode)."),
    ("sync_mode", str, "", "sync | backup | async; overrides sync_replicas when set."),
    ("comm_algo", str, "auto", "auto | oneshot | twoshot | nccl (gradient collective)."),
    ("placement", str, "bytes", "Owner-shard policy: bytes | round_robin (replica_device_setter order)."),
    ("save_checkpoint_secs", int, 600, "Chief checkpoint cadence in seconds (distribute_train.py:175)."),
    ("metrics_jsonl", str, "", "Append machine-readable step metrics (JSON lines) to this file; empty = off."),
    ("save_checkpoint_steps", int, 0, "Checkpoint every N global steps instead of every save_checkpoint_secs."),
    ("max_staleness", int, -1, "Async mode: drop a pushed gradient computed more than this many owner updates ago "
     "(-1 = no bound)."),
    ("use_cuda_graph", _str2bool, False, "Capture the training step into a CUDA graph and replay it."),
    ("lrn_compat_noop", _str2bool, False, "Reproduce the reference's `norm` no-op (distribute_tools.py:183-187)."),
]


class _Flags(object):
    def __init__(self):
        object.__setattr__(self, "_values", {})
        object.__setattr__(self, "_parsed", False)
        self.reset()

    def reset(self):
        vals: Dict[str, Any] = {}
        for name, _, default, _ in _DEFS:
            vals[name] = default
        object.__setattr__(self, "_values", vals)
        object.__setattr__(self, "_parsed", False)

    def __getattr__(self, name):
        vals = object.__getattribute__(self, "_values")
        name = name.strip()
        if name in vals:
            return vals[name]
        raise AttributeError("unknown flag %r" % name)

    def __setattr__(self, name, value):
        self._values[name.strip()] = value

    def define(self, name, typ, default, help_=""):
        _DEFS.append((name, typ, default, help_))
        self._values.setdefault(name, default)

    def parse(self, argv: Optional[List[str]] = None) -> List[str]:
        parser = argparse.ArgumentParser(add_help=True, allow_abbrev=False)
        for name, typ, default, help_ in _DEFS:
            parser.add_argument("--" + name, type=typ, default=default, help=help_)
        ns, rest = parser.parse_known_args(sys.argv[1:] if argv is None else argv)
        for k, v in vars(ns).items():
            self._values[k] = v
        object.__setattr__(self, "_parsed", True)
        return rest

    def as_dict(self):
        return dict(self._values)


FLAGS = _Flags()


def parse(argv=None):
    return FLAGS.parse(argv)


def app_run(main: Callable, argv: Optional[List[str]] = None):
    """tf.app.run(): parse flags, call main(argv_rest), exit with its return value."""
    rest = FLAGS.parse(argv)
    rc = main([sys.argv[0]] + rest)
    if rc:
        sys.exit(rc)
