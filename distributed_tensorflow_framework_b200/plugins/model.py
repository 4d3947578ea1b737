"""Model plugin base class. Parity: /root/reference/distribute_model.py:13-22.

The reference expects user model classes to be pasted into this module so that `@current_model(model='Name')`
can find them (distribute.py:123). Here user classes can live anywhere: subclassing `Model` registers them in the
plugin registry automatically (and explicit `@register_plugin` works for classes that do not subclass).
"""
import abc

from ..api import annotations as _ann


class Model(metaclass=abc.ABCMeta):
    def __init_subclass__(cls, **kwargs):
        super().__init_subclass__(**kwargs)
        _ann.register_plugin(cls)

    @abc.abstractmethod
    def inference(self, input_data, *args, **kwargs):
        """Forward pass: NHWC input batch -> network output (logits / image)."""
        raise NotImplementedError
