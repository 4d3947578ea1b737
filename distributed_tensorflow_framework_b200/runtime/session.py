"""MonitoredTrainingSession equivalent: init-or-restore, hooks, stop-at-step, chief-only periodic checkpoints,
recovery from lost workers.

Parity: /root/reference/distribute_train.py:169-180 -- MonitoredTrainingSession(master, is_chief, checkpoint_dir,
scaffold=Scaffold(init_op, init_fn), hooks=[StopAtStepHook, sync hook], save_checkpoint_secs=600, config,
stop_grace_period_secs=60, log_step_count_steps=100) and the loop `while not sess.should_stop(): sess.run(...)`.
In eager form `sess.run(step_fn)` executes one training step (a callable returning a dict with at least 'loss').
Checkpoints use the TF1 Saver V2 layout (io/tensorbundle.py): model.ckpt-<step>.index / .data-0000k-of-0000n +
`checkpoint` state file, keys = variable names, `global_step` int64, optimizer slots under TF names.
"""
import os

import numpy as np
import torch

from ..api import log as logger
from ..io import tensorbundle as tb
from . import hooks as H


class WorkerLostError(RuntimeError):
    """Recoverable: a peer stopped responding (the analogue of tf.errors.AbortedError / UnavailableError)."""


class Scaffold(object):
    def __init__(self, init_op=None, init_fn=None, saver=None):
        self.init_op, self.init_fn, self.saver = init_op, init_fn, saver


class Saver(object):
    """tf.train.Saver(sharded=True, max_to_keep=5) over a VariableStore (+ optimizer slots)."""

    def __init__(self, store, optimizer=None, max_to_keep=5, num_shards=1, gather_fn=None, lr_handler=None):
        """`gather_fn`: collective that completes the fp32 masters / slots on every rank in the owner-sharded modes
        (GradientEngine.gather_masters) -- `save` must then be called by EVERY rank at the same step, and only the
        rank with write=True touches the disk. `num_shards` > 1 writes one data shard per owner rank, the layout the
        reference's ps sharding produces (one shard per ps device). `lr_handler`: a LearningRate whose JSON file is
        written next to every checkpoint (/root/reference/distribute_learningrate.py:21-32)."""
        self.store, self.optimizer, self.max_to_keep = store, optimizer, max_to_keep
        self.num_shards = max(int(num_shards), 1)
        self.gather_fn = gather_fn
        self.lr_handler = lr_handler

    def save(self, model_dir, step, basename="model.ckpt", write=True):
        if self.gather_fn is not None:
            self.gather_fn()
        if not write:
            return None
        tensors, shard_of = {}, {}
        for v in self.store.all_variables():
            tensors[v.name] = v.logical(v.master.detach()).float().cpu().contiguous().numpy()
            shard_of[v.name] = v.owner % self.num_shards
        tensors["global_step"] = np.array(int(step), dtype=np.int64)
        if self.optimizer is not None:
            for k, t in self.optimizer.slot_tensors(self.store).items():
                tensors[k] = t.numpy()
                base = k.rsplit("/", 1)[0]
                shard_of[k] = shard_of.get(base, 0)
        prefix = os.path.join(model_dir, "%s-%d" % (basename, int(step)))
        tb.write_bundle(prefix, tensors, shard_of=shard_of, num_shards=self.num_shards)
        _, all_paths = tb.read_checkpoint_state(model_dir)
        name = os.path.basename(prefix)
        all_paths = [p for p in all_paths if p != name] + [name]
        while len(all_paths) > self.max_to_keep:
            old = all_paths.pop(0)
            for fn in os.listdir(model_dir):
                if fn.startswith(old + ".index") or fn.startswith(old + ".data-"):
                    try:
                        os.remove(os.path.join(model_dir, fn))
                    except OSError:
                        pass
        tb.write_checkpoint_state(model_dir, name, all_paths)
        if self.lr_handler is not None and getattr(self.lr_handler, "save_path", None):
            self.lr_handler.save()
        logger.info("Saving checkpoints for %d into %s.", int(step), prefix)
        return prefix

    def restore(self, prefix):
        with tb.BundleReader(prefix) as r:
            state = {k: torch.from_numpy(r.get_tensor(k)) for k in r.keys() if r.dtype_of(k) != tb.DT_BFLOAT16}
        missing = self.store.load_state_dict(state)
        step = int(state["global_step"]) if "global_step" in state else 0
        self.store.global_step = step
        if self.optimizer is not None:
            self.optimizer.load_slot_tensors(self.store, state)
        if missing:
            logger.warn("variables missing from checkpoint %s: %s", prefix, missing)
        logger.info("Restoring parameters from %s", prefix)
        return step


class MonitoredTrainingSession(object):
    def __init__(self, master="", is_chief=True, checkpoint_dir=None, scaffold=None, hooks=None, chief_only_hooks=None,
                 save_checkpoint_secs=600, save_checkpoint_steps=None, save_summaries_steps=100, config=None,
                 stop_grace_period_secs=60, log_step_count_steps=100, store=None, optimizer=None, saver=None,
                 summary_writer=None, barrier_fn=None, max_recoveries=3, collective_save=False, agree_fn=None,
                 on_recover=None):
        """`collective_save`: the checkpoint needs a collective gather (owner-sharded masters), so the saver hook runs
        on EVERY rank and the ranks agree on the saving step (step cadence, or the chief's wall-clock decision passed
        through `agree_fn(step, chief_decision) -> bool` on the control plane); only the chief writes.
        `on_recover`: called after a recoverable failure before the checkpoint is reloaded (re-arm the communicator)."""
        self.master, self.is_chief, self.checkpoint_dir = master, is_chief, checkpoint_dir
        self.scaffold = scaffold or Scaffold()
        self.store = store
        self.saver = saver or self.scaffold.saver or (Saver(store, optimizer) if store is not None else None)
        self.config = config
        self.stop_grace_period_secs = stop_grace_period_secs
        self._hooks = list(hooks or [])
        self._barrier_fn = barrier_fn
        self._max_recoveries = max_recoveries
        self._recoveries = 0
        self._stop = False
        self._writer = summary_writer
        self._on_recover = on_recover
        want_ckpt = checkpoint_dir and self.saver is not None and (save_checkpoint_secs or save_checkpoint_steps)
        if want_ckpt and collective_save and not is_chief:
            self._hooks.append(H.CheckpointSaverHook(self.saver, checkpoint_dir, save_checkpoint_secs,
                                                     save_checkpoint_steps, write=False, collective=True,
                                                     agree_fn=agree_fn))
        if is_chief:
            self._hooks += list(chief_only_hooks or [])
            if want_ckpt:
                self._hooks.append(H.CheckpointSaverHook(self.saver, checkpoint_dir, save_checkpoint_secs,
                                                         save_checkpoint_steps, write=True,
                                                         collective=collective_save, agree_fn=agree_fn))
            if checkpoint_dir and log_step_count_steps:
                self._hooks.append(H.StepCounterHook(every_n_steps=log_step_count_steps, summary_writer=self._writer))
            if self._writer is not None and save_summaries_steps:
                self._hooks.append(H.SummarySaverHook(self._writer, save_summaries_steps))

    # ------------------------------------------------------------------ life cycle
    @property
    def global_step(self):
        return self.store.global_step if self.store is not None else 0

    def __enter__(self):
        for h in self._hooks:
            h.begin()
        self._init_or_restore()
        for h in self._hooks:
            h.after_create_session(self)
        return self

    def _init_or_restore(self):
        """Restore from the latest checkpoint if there is one; otherwise run init_op, then init_fn(scaffold, sess) --
        as in tf.train.SessionManager.prepare_session, a model recovered from a checkpoint is NOT re-initialised (an
        init_fn that warm-starts from pretrained weights must not overwrite a resumed run).
        Non-chief workers restore the same checkpoint (every replica holds the full model) and then wait for the
        chief at the barrier -- the analogue of "Waiting for session to be initialized" (distribute_train.py:164-167)."""
        restored = False
        if self.checkpoint_dir and self.saver is not None:
            prefix = tb.latest_checkpoint(self.checkpoint_dir)
            if prefix is not None:
                self.saver.restore(prefix)
                restored = True
        if not restored:
            if self.scaffold.init_op is not None:
                self.scaffold.init_op()
            if self.scaffold.init_fn is not None:
                self.scaffold.init_fn(self.scaffold, self)
        if self._barrier_fn is not None:
            self._barrier_fn()

    def should_stop(self):
        return self._stop

    def request_stop(self):
        self._stop = True

    def run(self, step_fn, *args, **kwargs):
        """One step: before_run hooks -> step_fn -> after_run hooks. A WorkerLostError triggers recovery: every
        surviving rank reloads the last checkpoint and the step is retried."""
        ctx = H.RunContext(self)
        for h in self._hooks:
            h.before_run(ctx)
        while True:
            try:
                values = step_fn(*args, **kwargs)
                for h in self._hooks:
                    h.after_run(ctx, values)
                break
            except WorkerLostError as e:
                self._recoveries += 1
                if self._recoveries > self._max_recoveries:
                    raise
                logger.warn("recoverable failure (%s): restoring from the last checkpoint", e)
                if self._on_recover is not None:
                    self._on_recover(e)
                self._init_or_restore()
        if ctx.stop_requested:
            self._stop = True
        return values

    def close(self):
        self._stop = True

    def __exit__(self, exc_type, exc, tb_):
        if exc_type is None:
            for h in self._hooks:
                h.end(self)
        if self._writer is not None:
            self._writer.flush()
        return False
