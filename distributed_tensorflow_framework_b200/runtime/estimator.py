"""Estimator wrapper. Parity: /root/reference/distribute_estimator.py:25-35 -- `DistributeEstimator(model_fn,
eval_model_fn, config, params, input_class)`: a tf.estimator.Estimator that additionally carries the loader class
(by argument or by `@current_input`) and a separate eval model function. Fixes: the broken import is
gone and the check is on the instance. The base class is a small train/evaluate/predict driver over the Trainer.

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


class Estimator(object):
    """Minimal tf.estimator.Estimator: model_fn(features, labels, mode, params) -> dict(loss=..., predictions=...)."""
    TRAIN, EVAL, PREDICT = "train", "eval", "infer"

    def __init__(self, model_fn, model_dir=None, config=None, params=None):
        self._model_fn, self._model_dir = model_fn, model_dir or (config.model_dir if config else None)
        self.config, self.params = config, params or {}

    def train(self, input_fn, steps=None, hooks=None):
        out = None
        for i, (features, labels) in enumerate(input_fn()):
            if steps is not None and i >= steps:
                break
            out = self._model_fn(features, labels, self.TRAIN, self.params)
        return out

    def evaluate(self, input_fn, steps=None, model_fn=None):
        fn = model_fn or self._model_fn
        results = []
        for i, (features, labels) in enumerate(input_fn()):
            if steps is not None and i >= steps:
                break
            results.append(fn(features, labels, self.EVAL, self.params))
        return results

    def predict(self, input_fn):
        for features, _ in input_fn():
            yield self._model_fn(features, None, self.PREDICT, self.params)


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
