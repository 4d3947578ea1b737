"""NVTX ranges around steps / buckets / phases (the reference has no tracing; its only timing is the
wall clock around sess.run, /root/reference/distribute_train.py:184,194). Ranges are no-ops without CUDA."""
import contextlib

import torch


@contextlib.contextmanager
def range(name):  # noqa: A001  (mirrors torch.cuda.nvtx.range)
    on = torch.cuda.is_available()
    if on:
        torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        if on:
            torch.cuda.nvtx.range_pop()


def mark(name):
    if torch.cuda.is_available():
        torch.cuda.nvtx.mark(name)
