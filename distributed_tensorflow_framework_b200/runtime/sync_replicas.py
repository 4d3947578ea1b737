"""Synchronous replicas with optional backup workers, and the asynchronous parameter-server mode.

Parity: tf.train.SyncReplicasOptimizer as used at /root/reference/distribute_train.py:146-160:
  * every worker pushes its gradients, stamped with the global step it started from (`local_step`), into one
    ConditionalAccumulator per variable; a push whose stamp is older than the accumulator's global step is DROPPED
    (stale gradient of a straggler);
  * the chief's queue runner takes the mean once `replicas_to_aggregate` gradients have arrived, applies it with the
    wrapped optimizer, increments `global_step` and enqueues `tokens_per_step = max(total_num_replicas,
    replicas_to_aggregate)` tokens carrying the new global step;
  * every worker dequeues one token (the barrier) and adopts it as its new local_step.
`replicas_to_aggregate < total_num_replicas` gives backup workers. The reference has no async mode
(it always wraps the optimizer, :151); `AsyncReplicasOptimizer` is the extension required by BASELINE.json configs[4].

Two backends share the protocol:
  * `InProcessParameterServer` -- the deterministic fake backend of SURVEY section 4: N worker THREADS share host
    buffers; used for protocol tests (straggler / staleness injection) and for CPU multi-worker simulation.
  * collective backend -- real ranks (torch.distributed gloo on CPU, fused peer-memory kernels on CUDA); see
    `SyncReplicasOptimizer.apply_gradients`.
"""
import queue
import threading
from typing import Dict

import torch

from ..api import log as logger
from ..engine import variables as V


class ConditionalAccumulator(object):
    """tf.ConditionalAccumulator: sums gradients whose local_step >= the accumulator's global step."""

    def __init__(self, shape, dtype=torch.float32):
        self._sum = torch.zeros(shape, dtype=dtype)
        self._count = 0
        self._global_step = 0
        self._cv = threading.Condition()
        self.dropped = 0

    def set_global_step(self, step):
        with self._cv:
            self._global_step = max(self._global_step, int(step))

    def apply_grad(self, grad, local_step):
        with self._cv:
            if local_step < self._global_step:
                self.dropped += 1
                return False
            self._sum += grad
            self._count += 1
            self._cv.notify_all()
            return True

    def num_accumulated(self):
        with self._cv:
            return self._count

    def take_grad(self, num_required, timeout=None):
        """Block until >= num_required gradients were accumulated; return their mean, reset, bump the step."""
        with self._cv:
            if not self._cv.wait_for(lambda: self._count >= num_required, timeout=timeout):
                raise TimeoutError("accumulator: %d of %d gradients after %.1f s" % (self._count, num_required,
                                                                                      timeout))
            mean = self._sum / float(self._count)
            self._sum = torch.zeros_like(self._sum)
            self._count = 0
            self._global_step += 1
            return mean


class _TokenQueue(object):
    """The shared token FIFO of SyncReplicasOptimizer with one guard the TF queue does not have: a worker is never
    handed a second token of the step it already adopted. With a plain FIFO a fast worker can take two tokens of
    step s, push twice for step s+1 and starve a slow worker; in TF the run continues and the stale push is simply
    dropped, but with a fixed number of iterations per worker (tests, epoch-bounded jobs) the missing contribution
    stalls the last aggregation. Seeded tokens (replicas_to_aggregate > total_num_replicas: a worker must contribute
    more than once per step) are exempt."""

    def __init__(self):
        self._cv = threading.Condition()
        self._items = []  # (step, seeded)

    def put(self, step, seeded=False):
        with self._cv:
            self._items.append((int(step), bool(seeded)))
            self._cv.notify_all()

    def qsize(self):
        with self._cv:
            return len(self._items)

    def get(self, timeout=None, after_step=None):
        def pick():
            for i, (step, seeded) in enumerate(self._items):
                if seeded or after_step is None or step > after_step:
                    return i
            return None

        with self._cv:
            if not self._cv.wait_for(lambda: pick() is not None, timeout=timeout):
                raise queue.Empty()
            return self._items.pop(pick())[0]


class InProcessParameterServer(object):
    """Shared "ps" state for worker threads: parameters, accumulators, token queue, done queue, global step."""

    def __init__(self, store: V.VariableStore, optimizer, replicas_to_aggregate, total_num_replicas):
        self.store, self.optimizer = store, optimizer
        self.replicas_to_aggregate = replicas_to_aggregate
        self.total_num_replicas = total_num_replicas
        self.tokens_per_step = max(total_num_replicas, replicas_to_aggregate)
        self.vars = store.trainable_variables()
        self.accumulators: Dict[str, ConditionalAccumulator] = {
            v.name: ConditionalAccumulator(v.master.shape) for v in self.vars}
        self.token_queue = _TokenQueue()
        self.done_queue: "queue.Queue[int]" = queue.Queue()
        self.global_step = 0
        self._lock = threading.Lock()
        self._runner = None
        self._stop = threading.Event()
        self.applied_steps = 0

    # ---- chief side
    def start_chief_queue_runner(self, num_tokens=-1):
        if num_tokens == -1:
            # TF seeds the queue with `replicas_to_aggregate` tokens, which lets every worker run one step ahead of
            # the aggregation and makes its second push stale. Here only the tokens that are NEEDED for progress are
            # seeded (replicas_to_aggregate > total_num_replicas), so workers move in lock-step with the global step.
            num_tokens = max(0, self.replicas_to_aggregate - self.total_num_replicas)
        for _ in range(num_tokens):
            self.token_queue.put(self.global_step, seeded=True)
        if self._runner is None:
            self._runner = threading.Thread(target=self._chief_loop, daemon=True)
            self._runner.start()

    def _chief_loop(self):
        while not self._stop.is_set():
            try:
                self.aggregate_and_apply(timeout=0.2)
            except TimeoutError:
                continue

    def aggregate_and_apply(self, timeout=None):
        grads = []
        first = True
        for v in self.vars:
            # only the first accumulator may time out cleanly; once one mean is taken the step is committed
            grads.append((self.accumulators[v.name].take_grad(self.replicas_to_aggregate,
                                                              timeout if first else None), v))
            first = False
        with self._lock:
            self.optimizer.apply_gradients(grads, global_step=self.global_step, store=self.store)
            self.global_step += 1
            self.applied_steps += 1
            step = self.global_step
        for acc in self.accumulators.values():
            acc.set_global_step(step)
        for _ in range(self.tokens_per_step):
            self.token_queue.put(step)
        return step

    def stop(self):
        self._stop.set()
        if self._runner is not None:
            self._runner.join(timeout=2.0)

    # ---- worker side
    def push(self, grads_and_vars, local_step):
        accepted = 0
        for g, v in grads_and_vars:
            if self.accumulators[v.name].apply_grad(g.detach().to(torch.float32).cpu(), local_step):
                accepted += 1
        return accepted

    def take_token(self, timeout=None, after_step=None):
        """One token = permission to start the next step. `after_step`: the step of the token this worker took last;
        it will not be handed a second token of that same step (see _TokenQueue)."""
        return self.token_queue.get(timeout=timeout, after_step=after_step)


class SyncReplicasOptimizer(object):
    def __init__(self, opt, replicas_to_aggregate, total_num_replicas=None, use_locking=False, name="sync_replicas",
                 backend=None, engine=None, group=None, server=None):
        if total_num_replicas is None:
            total_num_replicas = replicas_to_aggregate
        self._opt = opt
        self.replicas_to_aggregate = int(replicas_to_aggregate)
        self.total_num_replicas = int(total_num_replicas)
        self.tokens_per_step = max(self.total_num_replicas, self.replicas_to_aggregate)
        self.name = name
        self._backend = backend  # InProcessParameterServer or None (collective)
        self._engine = engine  # parallel.comm.GradientEngine on CUDA
        self._group = group
        self._server = server
        self._local_step = 0
        self._last_token = None  # step of the last token taken from the fake backend
        self._gradients_applied = False
        if use_locking:
            logger.warn("use_locking=True is ignored: owner-side updates are lock-free by design")

    # delegate
    def compute_gradients(self, *args, **kwargs):
        return self._opt.compute_gradients(*args, **kwargs)

    @property
    def wrapped(self):
        return self._opt

    def apply_gradients(self, grads_and_vars, global_step=None, store=None):
        """Push gradients; block on the token barrier; return the new global step."""
        self._gradients_applied = True
        store = store or V.default_store()
        if self._backend is not None:  # fake backend: worker thread
            self._backend.push(grads_and_vars, self._local_step)
            self._local_step = self._backend.take_token(after_step=self._last_token)
            self._last_token = self._local_step
            return self._local_step
        world = self.total_num_replicas
        if self._engine is not None:  # CUDA: fused peer-memory collective (+ optimizer)
            self._engine.finish_step()
            self._local_step = store.global_step
            return self._local_step
        if world > 1:
            import torch.distributed as dist
            included = self._arrival_gate(store.global_step)
            flat = torch.cat([g.reshape(-1).float() for g, _ in grads_and_vars])
            if not included:
                flat.zero_()  # a dropped (stale) straggler contributes nothing
            dist.all_reduce(flat, group=self._group)
            flat /= float(min(self.replicas_to_aggregate, world))
            out, off = [], 0
            for g, v in grads_and_vars:
                n = g.numel()
                out.append((flat[off:off + n].view(g.shape), v))
                off += n
            grads_and_vars = out
        step = self._opt.apply_gradients(grads_and_vars, global_step=store.global_step, store=store)
        self._local_step = step
        return step

    def _arrival_gate(self, step):
        """Backup workers on real ranks: the first `replicas_to_aggregate` arrivals of a step are aggregated."""
        if self.replicas_to_aggregate >= self.total_num_replicas or self._server is None or self._server.store is None:
            return True
        order = int(self._server.store.add("sync_replicas/arrive/%d" % step, 1))
        return order <= self.replicas_to_aggregate

    def make_session_run_hook(self, is_chief, num_tokens=-1):
        return _SyncReplicasHook(self, is_chief, num_tokens)


class _SyncReplicasHook(object):
    """Chief: start the queue runner and seed the token queue; everybody: set local_step = global_step."""

    def __init__(self, sync_opt, is_chief, num_tokens):
        self._sync, self._is_chief, self._num_tokens = sync_opt, is_chief, num_tokens

    def begin(self):
        pass

    def after_create_session(self, session):
        b = self._sync._backend
        self._sync._local_step = session.global_step
        if b is not None and self._is_chief:
            b.global_step = session.global_step
            for acc in b.accumulators.values():
                acc.set_global_step(b.global_step)
            b.start_chief_queue_runner(self._num_tokens)

    def before_run(self, run_context):
        return None

    def after_run(self, run_context, run_values):
        pass

    def end(self, session):
        b = self._sync._backend
        if b is not None and self._is_chief:
            b.stop()


class AsyncReplicasOptimizer(object):
    """Asynchronous parameter-server mode (extension): every push is applied by the owner as soon as it arrives,
    possibly computed from stale weights; no token barrier. With the fake backend a lock-free shared store is
    updated directly (use_locking=False semantics); on CUDA the K-ASYNC push/apply kernels are used."""

    def __init__(self, opt, total_num_replicas=1, backend=None, engine=None, max_staleness=None):
        self._opt, self.total_num_replicas = opt, total_num_replicas
        self._backend, self._engine = backend, engine
        self.max_staleness = max_staleness
        self._local_step = 0
        self.dropped = 0

    def compute_gradients(self, *args, **kwargs):
        return self._opt.compute_gradients(*args, **kwargs)

    def apply_gradients(self, grads_and_vars, global_step=None, store=None, computed_at_step=None):
        store = store or V.default_store()
        if self._engine is not None:
            self._engine.finish_step()
            return store.global_step
        if self._backend is not None:
            b = self._backend
            with b._lock:  # serialises the step counter only; parameter reads by other workers stay lock-free
                if self.max_staleness is not None and computed_at_step is not None and \
                        b.global_step - computed_at_step > self.max_staleness:
                    self.dropped += 1
                    return b.global_step
                b.optimizer.apply_gradients(grads_and_vars, global_step=b.global_step, store=b.store)
                b.global_step += 1
                b.applied_steps += 1
                return b.global_step
        return self._opt.apply_gradients(grads_and_vars, global_step=store.global_step, store=store)

    def make_session_run_hook(self, is_chief, num_tokens=-1):
        return None  # no token barrier in asynchronous mode
