"""Annotation / reflection registry (IoC shell of the framework).

Parity: /root/reference/distribute_annotations.py:8-339 -- 18 decorator factories, each copying exactly one
whitelisted keyword argument onto the decorated object, plus the two getters. Differences from the reference are
the quirk fixes listed in SURVEY.md section 9: `ps_host` is accepted as an alias of `ps_hosts` (Q1), the advice key
`post_processs_fn` (sic) is accepted as well as `post_process_fn` (Q5), advice handlers must be callables (Q6),
and class lookup also consults the plugin registry so user classes do not have to be pasted into framework modules.
"""
from __future__ import annotations

import types
from typing import Any, Callable, Dict, Iterable, Optional

# name -> class; filled by @register_plugin and by the plugin base-class modules
_REGISTRY: Dict[str, type] = {}


def register_plugin(cls=None, *, name: Optional[str] = None):
    """Register a user class so that `@current_model(model='Name')` style lookups find it."""

    def deco(c):
        _REGISTRY[name or c.__name__] = c
        return c

    return deco(cls) if cls is not None else deco


def registered(name: str) -> Optional[type]:
    return _REGISTRY.get(name)


def _single_kwarg_decorator(attr: str, accepted: Iterable[str]) -> Callable[..., Callable]:
    accepted = tuple(accepted)

    def factory(**kwargs):
        def deco(obj):
            for key in accepted:
                if key in kwargs:
                    setattr(obj, attr, kwargs[key])
                    break
            return obj

        return deco

    factory.__name__ = attr
    return factory


def current_model(**kwargs):
    """@current_model(model='ClassName') -- distribute_annotations.py:8-23."""
    return _single_kwarg_decorator("model", ("model",))(**kwargs)


def current_input(**kwargs):
    """@current_input(input='ClassName') -- distribute_annotations.py:26-41."""
    return _single_kwarg_decorator("input", ("input",))(**kwargs)


def current_mode(**kwargs):
    """@current_mode(mode='Train'|'Eval') -- distribute_annotations.py:44-58."""
    return _single_kwarg_decorator("mode", ("mode",))(**kwargs)


def current_feature(**kwargs):
    """@current_feature(features={...}) -- distribute_annotations.py:61-78."""
    return _single_kwarg_decorator("features", ("features",))(**kwargs)


_ADVICE_KEYS = {
    "pre_fn": "pre_fn",
    "post_fn": "post_fn",
    "pre_process_fn": "pre_process_fn",
    "post_process_fn": "post_process_fn",
    "post_processs_fn": "post_process_fn",  # reference spelling (distribute_annotations.py:101)
    "init_fn": "init_fn",
    "pre_eval_fn": "pre_eval_fn",
    "post_eval_fn": "post_eval_fn",
}


def get_advice(**kwargs):
    """AOP advices woven at fixed cut-points of Train/Eval -- distribute_annotations.py:81-104.

    Unlike the reference (where `pre_fn=print(...)` silently became None) non-callable handlers are
    rejected, except None which means "no advice".
    """

    def deco(obj):
        for key, value in kwargs.items():
            if key not in _ADVICE_KEYS:
                continue
            if value is not None and not callable(value):
                raise TypeError("advice %s must be callable or None, got %r" % (key, type(value)))
            if value is not None and isinstance(obj, type):
                value = staticmethod(value)  # keep plain-function call signature on instances
            setattr(obj, _ADVICE_KEYS[key], value)
        return obj

    return deco


def parse_data_dir(**kwargs):
    """@parse_data_dir(parse_data_dir_fn=f): data_dir = f(data_dir) -- distribute_annotations.py:107-124."""

    def deco(obj):
        fn = kwargs.get("parse_data_dir_fn")
        if fn is not None:
            if not callable(fn):
                raise TypeError("parse_data_dir_fn must be callable")
            setattr(obj, "parse_data_dir_fn", staticmethod(fn) if isinstance(obj, type) else fn)
        return obj

    return deco


def gpu_num(**kwargs):
    """distribute_annotations.py:127-139."""
    return _single_kwarg_decorator("gpu_num", ("gpu_num",))(**kwargs)


def ps_hosts(**kwargs):
    """distribute_annotations.py:142-155; `ps_host` tolerated (distribute.py:37 typo)."""
    return _single_kwarg_decorator("ps_hosts", ("ps_hosts", "ps_host"))(**kwargs)


def worker_hosts(**kwargs):
    """distribute_annotations.py:158-171."""
    return _single_kwarg_decorator("worker_hosts", ("worker_hosts", "worker_host"))(**kwargs)


def job_name(**kwargs):
    """distribute_annotations.py:174-186."""
    return _single_kwarg_decorator("job_name", ("job_name",))(**kwargs)


def task_index(**kwargs):
    """distribute_annotations.py:189-201."""
    return _single_kwarg_decorator("task_index", ("task_index",))(**kwargs)


def batch_size(**kwargs):
    """distribute_annotations.py:204-215."""
    return _single_kwarg_decorator("batch_size", ("batch_size",))(**kwargs)


def sample_number(**kwargs):
    """distribute_annotations.py:218-229."""
    return _single_kwarg_decorator("sample_number", ("sample_number",))(**kwargs)


def epoch_num(**kwargs):
    """distribute_annotations.py:232-243."""
    return _single_kwarg_decorator("epoch_num", ("epoch_num",))(**kwargs)


def model_dir(**kwargs):
    """distribute_annotations.py:246-257."""
    return _single_kwarg_decorator("model_dir", ("model_dir",))(**kwargs)


def data_dir(**kwargs):
    """distribute_annotations.py:260-279."""
    return _single_kwarg_decorator("data_dir", ("data_dir",))(**kwargs)


def optimizer(**kwargs):
    """@optimizer(optimizer=<instance or None>) -- distribute_annotations.py:282-297."""
    return _single_kwarg_decorator("optimizer", ("optimizer",))(**kwargs)


def loss(**kwargs):
    """@loss(loss='ClassName') -- distribute_annotations.py:300-315."""
    return _single_kwarg_decorator("loss", ("loss",))(**kwargs)


# extensions (not in the reference): replica mode and eval handler selection
def sync_mode(**kwargs):
    """@sync_mode(sync_mode='sync'|'async'|'backup') -- flag-selectable sync vs async (BASELINE.json north star)."""
    return _single_kwarg_decorator("sync_mode", ("sync_mode",))(**kwargs)


def learning_rate(**kwargs):
    """@learning_rate(learning_rate=<LearningRate instance or class name>)."""
    return _single_kwarg_decorator("learning_rate", ("learning_rate",))(**kwargs)


def get_instance_from_annotation(obj: Any, attr: str, module: Optional[types.ModuleType] = None):
    """Class name stored under `attr` -> class (looked up in `module`, then in the plugin registry) -> instance
    created with `__new__` only (no `__init__`), exactly like distribute_annotations.py:318-333."""
    name = get_value_from_annotation(obj, attr)
    if isinstance(name, type):
        cls = name
    else:
        cls = getattr(module, name, None) if module is not None else None
        if cls is None:
            cls = _REGISTRY.get(name)
        if cls is None:
            raise ValueError("class %r (annotation %r) not found in module %s nor in the plugin registry"
                             % (name, attr, getattr(module, "__name__", None)))
    return cls.__new__(cls)


def get_value_from_annotation(obj: Any, attr: str):
    """distribute_annotations.py:336-339: raises ValueError when the annotation is missing."""
    if not hasattr(obj, attr):
        raise ValueError("annotation %r is missing on %r" % (attr, getattr(obj, "__name__", obj)))
    return getattr(obj, attr)
