"""Estimator wrapper. Parity: /root/reference/distribute_estimator.py:25-35 -- `DistributeEstimator(model_fn,
eval_model_fn, config, params, input_class)`: a tf.estimator.Estimator that additionally carries the loader class
(by argument or by `@current_input`) and a separate eval model function. Fixes: the broken import is
gone and the check is on the instance. The base class `Estimator` drives the framework's Trainer, Saver and
MonitoredTrainingSession (training with TF1-layout checkpoints in model_dir, evaluation from the latest checkpoint).

`RunConfig.uid` is /root/reference/distribute_utils.py:23-54.
"""
import collections

from ..plugins import input as Input


class RunConfig(object):
    _DEFAULT_UID_WHITE_LIST = ["tf_random_seed", "save_summary_steps", "save_checkpoints_steps",
                               "save_checkpoints_secs", "session_config", "keep_checkpoint_max",
                               "keep_checkpoint_every_n_hours", "log_step_count_steps"]

    def __init__(self, model_dir=None, cluster_spec=None, task_type="worker", task_id=0, **kwargs):
        self._model_dir, self._cluster_spec = model_dir, cluster_spec
        self._task_type, self._task_id = task_type, task_id
        for k, v in kwargs.items():
            setattr(self, "_" + k, v)

    model_dir = property(lambda self: self._model_dir)
    cluster_spec = property(lambda self: self._cluster_spec)
    task_type = property(lambda self: self._task_type)
    task_id = property(lambda self: self._task_id)
    is_chief = property(lambda self: self._task_id == 0)

    def uid(self, whitelist=None):
        if whitelist is None:
            whitelist = self._DEFAULT_UID_WHITE_LIST
        state = {k: v for k, v in self.__dict__.items() if not k.startswith("__")}
        for k in whitelist:
            state.pop("_" + k, None)
        ordered_state = collections.OrderedDict(sorted(state.items(), key=lambda t: t[0]))
        if ordered_state.get("_cluster_spec") is not None:
            ordered_state["_cluster_spec"] = collections.OrderedDict(
                sorted(ordered_state["_cluster_spec"].as_dict().items(), key=lambda t: t[0]))
        return ", ".join("%s=%r" % (k, v) for (k, v) in ordered_state.items())


class EstimatorSpec(object):
    """What a `model_fn(features, labels, mode, params)` returns so that the Estimator can drive the framework:
    `net` (a plugins.net.Net or a Model), `loss` (a plugins.loss.Loss) and, for training, `optimizer`.
    `predictions_fn(logits)` optionally maps the network output to predictions."""

    def __init__(self, net=None, loss=None, optimizer=None, predictions_fn=None):
        self.net, self.loss, self.optimizer, self.predictions_fn = net, loss, optimizer, predictions_fn


class Estimator(object):
    """tf.estimator.Estimator over this framework's Trainer / Saver / MonitoredTrainingSession.

    `model_fn(features, labels, mode, params)` is called once per mode with the first batch and returns an
    EstimatorSpec. `train` then builds a Trainer (fused kernels + collectives on CUDA, oracle path on CPU), restores
    the latest checkpoint of `model_dir` if there is one, runs the steps inside a MonitoredTrainingSession (periodic
    TF1-layout checkpoints, final save) and returns self; `evaluate` restores the latest checkpoint and returns
    {"loss", "accuracy" (classifiers), "global_step"}; `predict` yields per-batch predictions.
    For backward compatibility a model_fn that returns anything else (e.g. a dict) is simply mapped over the input."""
    TRAIN, EVAL, PREDICT = "train", "eval", "infer"

    def __init__(self, model_fn, model_dir=None, config=None, params=None, device=None):
        self._model_fn, self._model_dir = model_fn, model_dir or (config.model_dir if config else None)
        self.config, self.params = config, params or {}
        self._device = device
        self._trainer = None

    # ------------------------------------------------------------------ helpers
    def _pick_device(self):
        import torch
        if self._device is not None:
            return torch.device(self._device)
        return torch.device("cuda", 0) if torch.cuda.is_available() else torch.device("cpu")

    @staticmethod
    def _as_net(net):
        from ..plugins import net as net_mod
        return net if isinstance(net, net_mod.Net) else net_mod.Net(model=net)

    def _save_secs(self):
        return getattr(self.config, "_save_checkpoints_secs", None) if self.config is not None else None

    def _save_steps(self):
        return getattr(self.config, "_save_checkpoints_steps", None) if self.config is not None else None

    # ------------------------------------------------------------------ train
    def train(self, input_fn, steps=None, hooks=None, max_steps=None):
        from ..engine.trainer import Trainer
        from . import hooks as H
        from .session import MonitoredTrainingSession, Saver
        it = iter(input_fn())
        try:
            first = next(it)
        except StopIteration:
            return self
        spec = self._model_fn(first[0], first[1], self.TRAIN, self.params)
        if not isinstance(spec, EstimatorSpec):  # legacy: plain function mapped over the batches
            out, n = spec, 1
            for features, labels in it:
                if steps is not None and n >= steps:
                    break
                out = self._model_fn(features, labels, self.TRAIN, self.params)
                n += 1
            return out
        if self._trainer is None:
            from ..engine.optimizers import AdamOptimizer
            self._trainer = Trainer(self._as_net(spec.net), spec.loss, spec.optimizer or AdamOptimizer(0.001),
                                    device=self._pick_device())
        tr = self._trainer
        tr.build(first[0], first[1])
        saver = Saver(tr.store, tr.optimizer)
        sess_hooks = list(hooks or [])
        with MonitoredTrainingSession(is_chief=True, checkpoint_dir=self._model_dir, hooks=sess_hooks,
                                      save_checkpoint_secs=self._save_secs() or (600 if not self._save_steps() else None),
                                      save_checkpoint_steps=self._save_steps(), store=tr.store,
                                      optimizer=tr.optimizer, saver=saver) as sess:
            start = sess.global_step
            last = max_steps if max_steps is not None else (start + steps if steps is not None else None)
            batch = first
            while not sess.should_stop():
                if last is not None and sess.global_step >= last:
                    break
                values = sess.run(lambda b=batch: {"loss": tr.step(b), "summaries": tr.last_summaries})
                self.last_loss = values["loss"]
                try:
                    batch = next(it)
                except StopIteration:
                    break
        self.last_loss = float(self.last_loss) if getattr(self, "last_loss", None) is not None else None
        return self

    # ------------------------------------------------------------------ evaluate / predict
    def _eval_setup(self, first, mode, model_fn=None):
        import torch
        from ..engine import variables as V
        from ..io import tensorbundle as tb
        from .session import Saver
        spec = (model_fn or self._model_fn)(first[0], first[1], mode, self.params)
        if not isinstance(spec, EstimatorSpec):
            return None, None, None, spec
        device = self._pick_device()
        store = V.VariableStore(device)
        net = self._as_net(spec.net)
        from ..engine import layers
        layers.set_training(False)
        try:
            with V.using_store(store), torch.no_grad():
                net.process(first[0].to(device))  # creates the variables
        finally:
            layers.set_training(True)
        step = 0
        prefix = tb.latest_checkpoint(self._model_dir) if self._model_dir else None
        if prefix is not None:
            step = Saver(store).restore(prefix)
        return spec, store, step, None

    def evaluate(self, input_fn, steps=None, model_fn=None):
        import torch
        from ..engine import layers, variables as V
        it = iter(input_fn())
        try:
            first = next(it)
        except StopIteration:
            return {}
        spec, store, step, legacy = self._eval_setup(first, self.EVAL, model_fn)
        if spec is None:
            results = [legacy]
            for i, (features, labels) in enumerate(it, 1):
                if steps is not None and i >= steps:
                    break
                results.append((model_fn or self._model_fn)(features, labels, self.EVAL, self.params))
            return results
        net, dev = self._as_net(spec.net), store.device
        total, n, hits, seen = 0.0, 0, 0, 0
        layers.set_training(False)
        try:
            with V.using_store(store), torch.no_grad():
                batch = first
                while batch is not None and (steps is None or n < steps):
                    x, y = batch[0].to(dev), batch[1].to(dev)
                    logits = net.process(x)
                    if spec.loss is not None:
                        total += float(spec.loss.loss(logits, y))
                    if y.dtype in (torch.int64, torch.int32) and logits.dim() == 2:
                        hits += int((logits.float().argmax(1) == y.reshape(-1)).sum())
                        seen += logits.shape[0]
                    n += 1
                    batch = next(it, None)
        finally:
            layers.set_training(True)
        out = {"loss": total / max(n, 1), "global_step": step}
        if seen:
            out["accuracy"] = hits / float(seen)
        return out

    def predict(self, input_fn):
        import torch
        from ..engine import layers, variables as V
        it = iter(input_fn())
        first = next(it, None)
        if first is None:
            return
        spec, store, _, legacy = self._eval_setup(first, self.PREDICT)
        if spec is None:
            yield legacy
            for features, _ in it:
                yield self._model_fn(features, None, self.PREDICT, self.params)
            return
        net, dev = self._as_net(spec.net), store.device
        layers.set_training(False)
        try:
            with V.using_store(store), torch.no_grad():
                batch = first
                while batch is not None:
                    logits = net.process(batch[0].to(dev))
                    yield spec.predictions_fn(logits) if spec.predictions_fn is not None else logits
                    batch = next(it, None)
        finally:
            layers.set_training(True)


class DistributeEstimator(Estimator):
    def __init__(self, model_fn, eval_model_fn=None, config=None, params=None, input_class=None):
        super().__init__(model_fn, config=config, params=params)
        if input_class is None:
            input_class = getattr(self, "input", None) or getattr(type(self), "input", None)
        if isinstance(input_class, str):
            from ..api import annotations as _ann
            input_class = getattr(Input, input_class, None) or _ann.registered(input_class)
        if isinstance(input_class, type):
            input_class = input_class.__new__(input_class)
        assert input_class is None or isinstance(input_class, Input.Dataloader), \
            "input_class must be a Dataloader (use @current_input or the input_class argument)"
        self.input_class = input_class
        self.eval_model_fn = eval_model_fn

    def evaluate(self, input_fn, steps=None, model_fn=None):
        return super().evaluate(input_fn, steps, model_fn or self.eval_model_fn)
