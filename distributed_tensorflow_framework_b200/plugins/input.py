"""Input subsystem: loader kinds and the shuffle-batch pipeline.

Parity: /root/reference/distribute_input.py:19-208 -- `InputOptions`, `Dataloader` ABC (`load_train_batch`,
`load_eval_batch`, `_generate_image_batch`), `TFRecordDataLoader`, `DataPathDataLoader`, `PlaceholderDataLoader`.
TF's graph-mode queues become an eager pull pipeline: every call of `load_train_batch` returns the NEXT batch as
torch tensors; records come from the native threaded shuffling reader (io/csrc/dtf_io.cc) and batches are staged in
pinned host memory so the training loop can issue an asynchronous H2D copy.

Fixes: hooks use single-underscore names and the name-mangled legacy spellings are still looked
up (Q11); the non-shuffle branch no longer dereferences an int (Q12); DataPath readers receive the name queue (Q12);
the placeholder path builds its queue once and feeds every tower (Q10).
"""
import abc
import multiprocessing
import queue
import random
import threading
from enum import Enum

import numpy as np
import torch

from ..api import annotations as _ann
from ..api import constants
from ..api import flags as _flags
from ..io import tfrecord as tfr


class InputOptions(Enum):
    TF_RECORD = 0
    PLACEHOLDER = 1
    DATAPATHLOADER = 2


def _find_hook(obj, base_cls_names, name):
    """Look up a user hook by its plain name, then by the reference's name-mangled spellings."""
    fn = getattr(obj, name, None)
    if fn is not None:
        return fn
    for cls in [type(obj).__name__] + list(base_cls_names):
        fn = getattr(obj, "_%s_%s" % (cls.lstrip("_"), name), None)
        if fn is not None:
            return fn
    return None


def _to_tensor(x):
    if isinstance(x, torch.Tensor):
        return x
    return torch.as_tensor(np.asarray(x))


class FilenameQueue(object):
    """tf.train.string_input_producer([...], shuffle=True): endless, reshuffled every epoch."""

    def __init__(self, names, shuffle=True, seed=0):
        self.names = list(names) if not isinstance(names, str) else [names]
        self.shuffle = shuffle
        self._rng = random.Random(seed)
        self._cur = []

    def get(self):
        if not self._cur:
            self._cur = list(self.names)
            if self.shuffle:
                self._rng.shuffle(self._cur)
        return self._cur.pop()


class ShuffleBatcher(object):
    """tf.train.shuffle_batch / tf.train.batch: `num_threads` producer threads pull examples from `example_fn`,
    a pool of `capacity` examples is kept, batches are drawn at random (min_after_dequeue = 0) and assembled in
    pinned memory."""

    def __init__(self, example_fn, batch_size, capacity, num_threads=2, shuffle=True, seed=0, pin=True):
        self.example_fn = example_fn
        self.batch_size = batch_size
        self.capacity = max(int(capacity), batch_size)
        self.shuffle = shuffle
        self._rng = random.Random(seed)
        self._pool = []
        self._lock = threading.Condition()
        self._stop = False
        self._error = None
        self._exhausted = 0
        self._pin = pin and torch.cuda.is_available()
        self._nthreads = max(1, int(num_threads))
        self._threads = [threading.Thread(target=self._produce, daemon=True) for _ in range(self._nthreads)]
        for t in self._threads:
            t.start()

    def _produce(self):
        try:
            while not self._stop:
                try:
                    ex = self.example_fn()
                except StopIteration:
                    break
                ex = [_to_tensor(e) for e in ex]
                with self._lock:
                    while len(self._pool) >= self.capacity and not self._stop:
                        self._lock.wait(0.1)
                    self._pool.append(ex)
                    self._lock.notify_all()
        except Exception as e:  # surfaced to the consumer
            self._error = e
        finally:
            with self._lock:
                self._exhausted += 1
                self._lock.notify_all()

    def next_batch(self):
        picked = []
        with self._lock:
            while len(picked) < self.batch_size:
                while not self._pool:
                    if self._error is not None:
                        raise self._error
                    if self._exhausted >= self._nthreads:
                        raise StopIteration
                    self._lock.wait(0.1)
                idx = self._rng.randrange(len(self._pool)) if self.shuffle else 0
                if self.shuffle:
                    self._pool[idx], self._pool[-1] = self._pool[-1], self._pool[idx]
                    picked.append(self._pool.pop())
                else:
                    picked.append(self._pool.pop(0))
                self._lock.notify_all()
        out = []
        for j in range(len(picked[0])):
            first = picked[0][j]
            buf = torch.empty((self.batch_size,) + tuple(first.shape), dtype=first.dtype, pin_memory=self._pin)
            for i, ex in enumerate(picked):
                buf[i].copy_(ex[j])
            out.append(buf)
        return out

    def close(self):
        self._stop = True
        with self._lock:
            self._lock.notify_all()


class Dataloader(metaclass=abc.ABCMeta):
    """Base loader. Attributes injected by the entry script (/root/reference/distribute.py:91,98-101):
    batch_size, sample_number, data_dir, gpu_num, features."""
    batch_size = 1
    sample_number = 1
    data_dir = ""
    gpu_num = 1

    def __init_subclass__(cls, **kwargs):
        super().__init_subclass__(**kwargs)
        _ann.register_plugin(cls)

    @abc.abstractmethod
    def load_train_batch(self, name_queue=None, *args, **kwargs):
        """Return (raw_data, ground_truth) for the next training batch."""

    @abc.abstractmethod
    def load_eval_batch(self, *args, **kwargs):
        """Return (raw_data, ground_truth) for the next evaluation batch, or None when exhausted."""

    def _generate_image_batch(self, example_fn, min_queue_examples, num_thread, shuffle=True):
        """distribute_input.py:54-69: capacity = min_queue_examples + 3 * batch_size, min_after_dequeue = 0."""
        return ShuffleBatcher(example_fn, self.batch_size, capacity=min_queue_examples + 3 * self.batch_size,
                              num_threads=num_thread, shuffle=shuffle)


class TFRecordDataLoader(Dataloader, metaclass=abc.ABCMeta):
    type = "TFRecordDataLoader"
    features = None

    def load_train_batch(self, name_queue=None, *args, **kwargs):
        if name_queue is None:
            raise RuntimeError("Cannot find get the queue from tf-record.")
        return self.load_batch_from_tfrecord(name_queue, *args, **kwargs)

    def load_eval_batch(self, *args, **kwargs):
        return None

    def _decode_raw_data(self, raw_features, height, width, *args, **kwargs):
        """User hook: parsed feature dict -> [raw_data, ground_truth] (arrays / tensors) for ONE example."""
        raise NotImplementedError("implement _decode_raw_data(raw_features, height, width)")

    def load_batch_from_tfrecord(self, filename_queue, *args, **kwargs):
        """distribute_input.py:92-106: read -> parse_single_example(self.features) -> user decode -> shuffle batch."""
        if getattr(self, "_batcher", None) is None:
            height = _flags.FLAGS.input_image_height
            width = _flags.FLAGS.input_image_width
            names = filename_queue.names if isinstance(filename_queue, FilenameQueue) else filename_queue
            min_queue_examples = int(self.sample_number * constants.MIN_FRACTION_OF_EXAMPLE_IN_QUEUE)
            records = tfr.ShufflingRecordQueue(names, capacity=max(min_queue_examples, 4 * self.batch_size),
                                               shuffle=True, loop=True, seed=getattr(self, "seed", 0))
            decode = _find_hook(self, ["TFRecordDataLoader"], "_decode_raw_data")
            if self.features is None:
                raise ValueError("Please use @current_feature to create your features for data_loader")
            lock = threading.Lock()

            def example_fn():
                with lock:  # the record handle returned by the native queue is valid until the next get()
                    rec = records.get()
                feats = tfr.parse_single_example(rec, self.features)
                return decode(feats, height, width, *args, **kwargs)

            self._records = records
            self._batcher = self._generate_image_batch(example_fn, min_queue_examples,
                                                       min(multiprocessing.cpu_count() * 2, 8), shuffle=True)
        raw, gt = self._batcher.next_batch()[:2]
        return raw, gt


class DataPathDataLoader(Dataloader, metaclass=abc.ABCMeta):
    type = "DataPathDataLoader"

    def load_train_batch(self, name_queue=None, *args, **kwargs):
        assert name_queue is not None, "name queue cannot be None for DataPathDataLoader!"
        if getattr(self, "_batcher", None) is None:
            create_reader = _find_hook(self, ["DataPathDataLoader"], "_create_reader")
            parse = _find_hook(self, ["DataPathDataLoader"], "_parse_raw_data")
            reader = create_reader()
            lock = threading.Lock()

            def example_fn():
                with lock:
                    name = name_queue.get()
                value = reader(name) if callable(reader) else reader.read(name)
                return parse(value)

            min_queue_examples = int(self.sample_number * constants.MIN_FRACTION_OF_EXAMPLE_IN_QUEUE)
            self._batcher = self._generate_image_batch(example_fn, min_queue_examples,
                                                       min(multiprocessing.cpu_count() * 2, 8), shuffle=True)
        raw, gt = self._batcher.next_batch()[:2]
        return raw, gt

    def load_eval_batch(self, *args, **kwargs):
        return None

    def _parse_raw_data(self, raw_reader_data):
        """User hook: reader output -> [raw_data, ground_truth] for one example."""
        raise NotImplementedError

    def _create_reader(self):
        """User hook: return a callable `reader(path) -> value` (or an object with .read(path))."""
        raise NotImplementedError

    @abc.abstractmethod
    def create_name_queue(self, data_dir):
        """Return a FilenameQueue (or any object with .get()) over the sample paths."""


class PlaceholderDataLoader(Dataloader, metaclass=abc.ABCMeta):
    type = "PlaceholderDataLoader"

    def load_train_batch(self, name_queue=None, *args, **kwargs):
        create = _find_hook(self, ["PlaceholderDataLoader"], "_create_placeholder")
        return create(*args, **kwargs)

    def load_eval_batch(self, *args, **kwargs):
        return None

    def load_queue_for_placeholder(self, *args, **kwargs):
        batch_queue = queue.Queue()
        put = _find_hook(self, ["PlaceholderDataLoader"], "_put_names_dict_into_queue")
        put(batch_queue, *args, **kwargs)
        return batch_queue

    def _create_placeholder(self, *args, **kwargs):
        """User hook: return (raw_spec, gt_spec) placeholders (shape/dtype descriptions; informational here)."""
        return None, None

    @abc.abstractmethod
    def _put_names_dict_into_queue(self, queue, *args, **kwargs):
        """User hook: put one path-dict per sample into `queue`."""

    @abc.abstractmethod
    def decode_data_from_path_name(self, paths, *args, **kwargs):
        """User hook: path dict -> {'raw_data': array, 'ground_truth': array}."""

    def load_placeholder_data(self, sample_path_queue, *args, **kwargs):
        """distribute_input.py:198-208: pop batch_size path dicts (re-queued at the tail), decode on the host."""
        raw_batch, ground_truth_batch = [], []
        for _ in range(self.batch_size):
            paths = sample_path_queue.get()
            data = self.decode_data_from_path_name(paths, *args, **kwargs)
            raw_batch.append(_to_tensor(data["raw_data"]))
            ground_truth_batch.append(_to_tensor(data["ground_truth"]))
            sample_path_queue.put(paths)
        return torch.stack(raw_batch), torch.stack(ground_truth_batch)


class SyntheticDataLoader(PlaceholderDataLoader):
    """Synthetic data of a named shape (BASELINE.json: "synthetic data of the named shape"). A small pool of
    pinned host batches is cycled so the host->device copy of every step is real. Class attributes / injected
    attributes: image_shape (H, W, C), num_classes (0 => image-to-image target of `target_shape`)."""
    type = "PlaceholderDataLoader"
    image_shape = (32, 32, 3)
    num_classes = 10
    target_shape = None
    pool_batches = 4
    seed = 0

    def _make_pool(self):
        g = torch.Generator().manual_seed(self.seed)
        pin = torch.cuda.is_available()
        self._pool = []
        for _ in range(self.pool_batches):
            x = torch.rand((self.batch_size,) + tuple(self.image_shape), generator=g)
            if self.num_classes:
                y = torch.randint(0, self.num_classes, (self.batch_size,), generator=g)
            else:
                y = torch.rand((self.batch_size,) + tuple(self.target_shape or self.image_shape), generator=g)
            if pin:
                x, y = x.pin_memory(), y.pin_memory()
            self._pool.append((x, y))
        self._i = 0

    def next_host_batch(self):
        if getattr(self, "_pool", None) is None or self._pool[0][0].shape[0] != self.batch_size:
            self._make_pool()
        b = self._pool[self._i % len(self._pool)]
        self._i += 1
        return b

    def load_train_batch(self, name_queue=None, *args, **kwargs):
        return self.next_host_batch()

    def load_eval_batch(self, *args, **kwargs):
        n = getattr(self, "_eval_served", 0)
        if n >= getattr(self, "eval_batches", 4):
            self._eval_served = 0
            return None
        self._eval_served = n + 1
        return self.next_host_batch()

    def _put_names_dict_into_queue(self, q, *args, **kwargs):
        for i in range(self.batch_size * self.pool_batches):
            q.put({"index": i})

    def decode_data_from_path_name(self, paths, *args, **kwargs):
        if getattr(self, "_pool", None) is None:
            self._make_pool()
        i = paths["index"]
        x, y = self._pool[(i // self.batch_size) % len(self._pool)]
        return {"raw_data": x[i % self.batch_size], "ground_truth": y[i % self.batch_size]}

    def load_placeholder_data(self, sample_path_queue, *args, **kwargs):
        return self.next_host_batch()
