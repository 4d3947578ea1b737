"""TensorBoard event-file writer (scalar summaries) on the native TFRecord framing.

Parity: tf.summary.scalar per loss term (/root/reference/distribute_tower.py:137-143) written by the default
SummarySaverHook of MonitoredTrainingSession into model_dir.
"""
import os
import socket
import time

from ._iolib import lib


class EventFileWriter(object):
    def __init__(self, logdir, suffix=""):
        os.makedirs(logdir, exist_ok=True)
        name = "events.out.tfevents.%010d.%s%s" % (int(time.time()), socket.gethostname(), suffix)
        self.path = os.path.join(logdir, name)
        self._h = lib().dtf_events_open(self.path.encode(), time.time())
        if not self._h:
            raise IOError("cannot create %s" % self.path)

    def add_scalar(self, tag, value, step, wall_time=None):
        lib().dtf_events_scalar(self._h, tag.encode(), float(value), int(step),
                                time.time() if wall_time is None else wall_time)

    def flush(self):
        lib().dtf_events_flush(self._h)

    def close(self):
        if self._h:
            lib().dtf_events_close(self._h)
            self._h = None


def read_scalars(path):
    """Independent pure-Python reader (used by tests): [(step, tag, value)]."""
    import struct
    from .tfrecord import tf_record_iterator

    def varint(b, i):
        v = s = 0
        while True:
            c = b[i]
            i += 1
            v |= (c & 0x7f) << s
            s += 7
            if not c & 0x80:
                return v, i

    def fields(b):
        i = 0
        while i < len(b):
            tag, i = varint(b, i)
            f, w = tag >> 3, tag & 7
            if w == 0:
                v, i = varint(b, i)
            elif w == 1:
                v = b[i:i + 8]
                i += 8
            elif w == 5:
                v = b[i:i + 4]
                i += 4
            else:
                n, i = varint(b, i)
                v = b[i:i + n]
                i += n
            yield f, w, v

    out = []
    for rec in tf_record_iterator(path):
        step, summary = 0, None
        for f, w, v in fields(rec):
            if f == 2:
                step = v
            elif f == 5:
                summary = v
        if summary is None:
            continue
        for f, w, v in fields(summary):
            if f != 1:
                continue
            tag, val = None, None
            for f2, w2, v2 in fields(v):
                if f2 == 1:
                    tag = v2.decode()
                elif f2 == 2:
                    val = struct.unpack("<f", v2)[0]
            out.append((step, tag, val))
    return out
