"""Session hooks (tf.train.SessionRunHook protocol in eager form).

Parity: StopAtStepHook / CheckpointSaverHook / StepCounterHook / SummarySaverHook are the hooks that
MonitoredTrainingSession installs for the reference (/root/reference/distribute_train.py:169-180);
`ExamplesPerSecondHook` is /root/reference/distribute_utils.py:57-116 (average and current examples/sec every N
steps or seconds; exactly one of the two intervals must be given).
"""
import json
import time

from ..api import log as logger


class SessionRunHook(object):
    def begin(self):
        pass

    def after_create_session(self, session):
        pass

    def before_run(self, run_context):
        return None

    def after_run(self, run_context, run_values):
        pass

    def end(self, session):
        pass


class RunContext(object):
    def __init__(self, session):
        self.session = session
        self._stop = False

    def request_stop(self):
        self._stop = True

    @property
    def stop_requested(self):
        return self._stop


class SecondOrStepTimer(object):
    """basic_session_run_hooks.SecondOrStepTimer."""

    def __init__(self, every_secs=None, every_steps=None):
        if (every_secs is None) == (every_steps is None):
            raise ValueError("Either every_secs or every_steps should be provided.")
        self._every_secs, self._every_steps = every_secs, every_steps
        self.reset()

    def reset(self):
        self._last_time = None
        self._last_step = None

    def should_trigger_for_step(self, step):
        if self._last_step is None:
            return True
        if self._last_step == step:
            return False
        if self._every_secs is not None and time.time() >= self._last_time + self._every_secs:
            return True
        if self._every_steps is not None and step >= self._last_step + self._every_steps:
            return True
        return False

    def update_last_triggered_step(self, step):
        now = time.time()
        if self._last_time is None:
            elapsed_secs, elapsed_steps = None, None
        else:
            elapsed_secs, elapsed_steps = now - self._last_time, step - self._last_step
        self._last_time, self._last_step = now, step
        return elapsed_secs, elapsed_steps

    def last_triggered_step(self):
        return self._last_step


class StopAtStepHook(SessionRunHook):
    def __init__(self, num_steps=None, last_step=None):
        if (num_steps is None) == (last_step is None):
            raise ValueError("exactly one of num_steps and last_step must be given")
        self._num_steps, self._last_step = num_steps, last_step

    def after_create_session(self, session):
        if self._last_step is None:
            self._last_step = session.global_step + self._num_steps

    def after_run(self, run_context, run_values):
        if run_context.session.global_step >= self._last_step:
            run_context.request_stop()


class StepCounterHook(SessionRunHook):
    """Logs global_step/sec every N steps (log_step_count_steps=100, distribute_train.py:180)."""

    def __init__(self, every_n_steps=100, every_n_secs=None, summary_writer=None):
        self._timer = SecondOrStepTimer(every_steps=every_n_steps if every_n_secs is None else None,
                                        every_secs=every_n_secs)
        self._writer = summary_writer
        self.last_rate = None

    def after_run(self, run_context, run_values):
        step = run_context.session.global_step
        if self._timer.should_trigger_for_step(step):
            secs, steps = self._timer.update_last_triggered_step(step)
            if secs:
                self.last_rate = steps / secs
                logger.info("global_step/sec: %g", self.last_rate)
                if self._writer is not None:
                    self._writer.add_scalar("global_step/sec", self.last_rate, step)


class ExamplesPerSecondHook(SessionRunHook):
    def __init__(self, batch_size, every_n_steps=100, every_n_secs=None):
        if (every_n_steps is None) == (every_n_secs is None):
            raise ValueError("exactly one of every_n_steps and every_n_secs should be provided.")
        self._timer = SecondOrStepTimer(every_steps=every_n_steps, every_secs=every_n_secs)
        self._step_train_time = 0
        self._total_steps = 0
        self._batch_size = batch_size
        self.average_examples_per_sec = None
        self.current_examples_per_sec = None

    def after_run(self, run_context, run_values):
        global_step = run_context.session.global_step
        if self._timer.should_trigger_for_step(global_step):
            elapsed_time, elapsed_steps = self._timer.update_last_triggered_step(global_step)
            if elapsed_time is not None:
                steps_per_sec = elapsed_steps / elapsed_time
                self._step_train_time += elapsed_time
                self._total_steps += elapsed_steps
                self.average_examples_per_sec = self._batch_size * (self._total_steps / self._step_train_time)
                self.current_examples_per_sec = steps_per_sec * self._batch_size
                logger.info("%s: %g (%g), step = %g", "Average examples/sec", self.average_examples_per_sec,
                            self.current_examples_per_sec, self._total_steps)


class SummarySaverHook(SessionRunHook):
    """Writes the scalar summaries returned by the step function into a TensorBoard event file."""

    def __init__(self, writer, save_steps=100):
        self._writer, self._save_steps = writer, save_steps

    def after_run(self, run_context, run_values):
        step = run_context.session.global_step
        if not isinstance(run_values, dict) or step % self._save_steps != 0:
            return
        for tag, val in (run_values.get("summaries") or {}).items():
            self._writer.add_scalar(tag, float(val), step)
        self._writer.flush()

    def end(self, session):
        self._writer.flush()


class CheckpointSaverHook(SessionRunHook):
    """Periodic save (save_checkpoint_secs=600, distribute_train.py:175) + final save at end.

    Plain mode (replicated masters): installed on the chief only, local timer. Collective mode (owner-sharded
    masters: saving needs a gather that every rank must enter): installed on EVERY rank; the ranks save at the same
    global step -- a step cadence is deterministic, a wall-clock cadence is decided by the chief and agreed through
    `agree_fn(step, chief_decision)` every `probe_every` steps; `write` is True on the chief only."""

    def __init__(self, saver, checkpoint_dir, save_secs=600, save_steps=None, write=True, collective=False,
                 agree_fn=None, probe_every=10):
        self._saver, self._dir = saver, checkpoint_dir
        self._timer = SecondOrStepTimer(every_secs=save_secs if save_steps is None else None, every_steps=save_steps)
        self._by_steps = save_steps is not None
        self._write, self._collective, self._agree, self._probe = write, collective, agree_fn, max(int(probe_every), 1)
        self._last_saved = None

    def after_create_session(self, session):
        self._timer.update_last_triggered_step(session.global_step)

    def after_run(self, run_context, run_values):
        step = run_context.session.global_step
        if not self._collective or self._by_steps:
            trigger = self._timer.should_trigger_for_step(step)
        elif step % self._probe != 0:
            trigger = False
        else:
            mine = self._timer.should_trigger_for_step(step) if self._write else False
            trigger = self._agree(step, mine) if self._agree is not None else mine
        if trigger:
            self._timer.update_last_triggered_step(step)
            self._save(run_context.session, step)

    def _save(self, session, step):
        if self._last_saved == step:
            return
        self._saver.save(self._dir, step, write=self._write)
        self._last_saved = step

    def end(self, session):
        self._save(session, session.global_step)


class HeartbeatHook(SessionRunHook):
    """Failure detection (extension; the reference has none). Two signals:
    * every step, the communicator's host-mapped error record is read (no CUDA call): a device-side wait that ran
      into its time bound (a peer that never reached a barrier) raises WorkerLostError at once;
    * every N steps each rank bumps a counter in every peer's flag page; a peer whose counter has not moved for
      `timeout_s` raises WorkerLostError too.
    The session treats WorkerLostError as recoverable (reload the last checkpoint)."""

    def __init__(self, comm, every_n_steps=50, timeout_s=120.0):
        self._comm, self._every, self._timeout = comm, every_n_steps, timeout_s
        self._beat = 0
        self._last_seen = {}
        self._last_change = {}

    def after_run(self, run_context, run_values):
        if self._comm is None:
            return
        self._comm.check()
        step = run_context.session.global_step
        if step % self._every != 0:
            return
        self._beat += 1
        seen = self._comm.heartbeat(self._beat).tolist()
        now = time.time()
        for r in range(self._comm.world):
            if self._last_seen.get(r) != seen[r]:
                self._last_seen[r] = seen[r]
                self._last_change[r] = now
            elif now - self._last_change.get(r, now) > self._timeout:
                from .session import WorkerLostError
                raise WorkerLostError("rank %d silent for %.0f s" % (r, now - self._last_change[r]))


class MetricsJsonlHook(SessionRunHook):
    """Machine-readable training metrics (extension): one JSON object per line every N steps with the
    step, the loss (if the step function returned one), wall-clock examples/sec over the interval and -- when a
    `device_ms_fn` is given (e.g. a CUDA-event timer of the trainer) -- the device-timed milliseconds per step.
    The file is opened in append mode, so a resumed run continues the same log."""

    def __init__(self, path, batch_size, every_n_steps=10, device_ms_fn=None, extra=None):
        self._path, self._batch, self._every = path, int(batch_size), int(every_n_steps)
        self._device_ms_fn, self._extra = device_ms_fn, dict(extra or {})
        self._t0 = self._step0 = None
        self._f = None

    def begin(self):
        self._f = open(self._path, "a", buffering=1)

    def after_create_session(self, session):
        self._t0, self._step0 = time.time(), session.global_step

    def after_run(self, run_context, run_values):
        step = run_context.session.global_step
        if self._f is None or self._t0 is None or step % self._every != 0 or step == self._step0:
            return
        now = time.time()
        rec = {"step": int(step), "time": now,
               "examples_per_sec": self._batch * (step - self._step0) / max(now - self._t0, 1e-9)}
        if isinstance(run_values, dict):
            loss = run_values.get("loss")
        else:
            loss = run_values[0] if isinstance(run_values, (tuple, list)) and run_values else run_values
        try:
            if loss is not None:
                rec["loss"] = float(loss)
        except (TypeError, ValueError):
            pass
        if self._device_ms_fn is not None:
            rec["device_ms_per_step"] = float(self._device_ms_fn())
        rec.update(self._extra)
        self._f.write(json.dumps(rec) + "\n")
        self._t0, self._step0 = now, step

    def end(self, session):
        if self._f is not None:
            self._f.close()
            self._f = None
