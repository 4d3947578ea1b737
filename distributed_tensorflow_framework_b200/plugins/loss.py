"""Loss plugin base class + built-in losses. Parity: /root/reference/distribute_loss.py:11-19.

Built-ins cover the reference's implied workloads: softmax cross-entropy (CIFAR / ResNet) and MSE
(dehazing FCN). On CUDA they run the fused forward+backward kernels; on CPU they are plain torch ops (the oracle).
"""
import abc

import torch.nn.functional as F

from ..api import annotations as _ann


class Loss(metaclass=abc.ABCMeta):
    def __init_subclass__(cls, **kwargs):
        super().__init_subclass__(**kwargs)
        _ann.register_plugin(cls)

    @abc.abstractmethod
    def loss(self, predict, ground_truth):
        """Return a scalar tensor."""
        raise NotImplementedError


class SoftmaxCrossEntropyLoss(Loss):
    """Mean sparse softmax cross-entropy; `ground_truth` holds int64 class ids."""

    def loss(self, predict, ground_truth):
        from ..engine import functional as Fn
        labels = ground_truth.reshape(-1).long()
        logits = predict.reshape(labels.shape[0], -1)
        if logits.is_cuda:
            return Fn.softmax_xent(logits, labels)
        return F.cross_entropy(logits.float(), labels)


class MSELoss(Loss):
    """Mean squared error between the prediction and a same-shaped target (image-to-image nets)."""

    def loss(self, predict, ground_truth):
        from ..engine import functional as Fn
        if predict.is_cuda:
            return Fn.mse(predict, ground_truth.to(predict.dtype))
        return F.mse_loss(predict.float(), ground_truth.float())
