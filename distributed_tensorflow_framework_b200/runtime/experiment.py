"""Experiment wrapper. Parity: /root/reference/distribute_experiment.py:11-83 -- `DistributeExperiment(mode, train_fn,
train_dataloader, eval_fn, eval_dataloader, features)` with `train(pre_train_fn, post_train_fn)`,
`evaluation(pre_eval_fn, post_evaluation_fn)` and `run()`; local decorator `current_input(train_input=, eval_input=)`.
Fixes: the `data_load_option` flag exists; the loader kind is read from the loader class attribute."""
from ..api import annotations as _ann
from ..api import flags
from ..plugins import input as Input

_OPTION_BY_FLAG = {"tfrecord": Input.InputOptions.TF_RECORD, "placeholder": Input.InputOptions.PLACEHOLDER,
                   "datapath": Input.InputOptions.DATAPATHLOADER}
_OPTION_BY_TYPE = {"TFRecordDataLoader": Input.InputOptions.TF_RECORD,
                   "PlaceholderDataLoader": Input.InputOptions.PLACEHOLDER,
                   "DataPathDataLoader": Input.InputOptions.DATAPATHLOADER}


def current_input(**kwargs):
    """@current_input(train_input='Cls', eval_input='Cls') -- distribute_experiment.py:11-20."""

    def deco(obj):
        for key in ("train_input", "eval_input"):
            if key in kwargs:
                setattr(obj, key, kwargs[key])
        return obj

    return deco


class DistributeExperiment(object):
    def __init__(self, mode, train_fn=None, train_dataloader=None, eval_fn=None, eval_dataloader=None, features=None):
        if mode not in ("Train", "Eval"):
            raise ValueError("mode must be 'Train' or 'Eval'")
        self.mode = mode
        self.train_fn, self.eval_fn = train_fn, eval_fn
        self.features = features
        self.train_dataloader = self._make_loader(train_dataloader, "train_input")
        self.eval_dataloader = self._make_loader(eval_dataloader, "eval_input")
        if mode == "Train" and (train_fn is None or self.train_dataloader is None):
            raise ValueError("Train mode needs train_fn and a train dataloader")
        if mode == "Eval" and (eval_fn is None or self.eval_dataloader is None):
            raise ValueError("Eval mode needs eval_fn and an eval dataloader")
        loader = self.train_dataloader if mode == "Train" else self.eval_dataloader
        kind = getattr(type(loader), "type", None)
        self.input_mode = _OPTION_BY_TYPE.get(kind) or _OPTION_BY_FLAG.get(flags.FLAGS.data_load_option)
        if self.input_mode is None:
            raise ValueError("unknown data_load_option %r" % flags.FLAGS.data_load_option)
        if self.input_mode == Input.InputOptions.TF_RECORD:
            if features is None:
                raise ValueError("TFRecord loaders need `features`")
            loader.features = features

    def _make_loader(self, spec, attr):
        if spec is None:
            spec = getattr(self, attr, None) or getattr(type(self), attr, None)
        if spec is None:
            return None
        if isinstance(spec, str):
            cls = getattr(Input, spec, None) or _ann.registered(spec)
            if cls is None:
                raise ValueError("loader class %r not found" % spec)
            return cls.__new__(cls)
        if isinstance(spec, type):
            return spec.__new__(spec)
        return spec

    def train(self, pre_train_fn=None, post_train_fn=None):
        return self.train_fn(self.train_dataloader, self.input_mode, pre_train_fn, post_train_fn)

    def evaluation(self, pre_eval_fn=None, post_evaluation_fn=None):
        return self.eval_fn(self.eval_dataloader, pre_eval_fn, post_evaluation_fn)

    def run(self):
        if self.mode == "Train":
            return self.train()
        return self.evaluation()
