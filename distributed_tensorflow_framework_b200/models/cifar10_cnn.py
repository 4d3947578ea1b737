"""CIFAR-10 tutorial CNN (the lineage of the reference's layer library: /root/reference/distribute_tools.py:1-13):
conv5x5(64) - pool3x3/2 - lrn - conv5x5(64) - lrn - pool3x3/2 - FC384 - FC192 - logits(10), NHWC, 32x32x3 inputs
(BASELINE.json configs[0] and [1])."""
from ..engine import layers as L
from ..plugins.input import SyntheticDataLoader
from ..plugins.model import Model


class Cifar10CNN(Model):
    num_classes = 10

    def inference(self, images, *args, **kwargs):
        x = L.conv("conv1", images, 3, 64, kernel_size=[5, 5])
        x = L.pool("pool1", x, kernel=[1, 3, 3, 1], stride=[1, 2, 2, 1])
        x = L.norm("norm1", x, lsize=4)
        x = L.conv("conv2", x, 64, 64, kernel_size=[5, 5])
        x = L.norm("norm2", x, lsize=4)
        x = L.pool("pool2", x, kernel=[1, 3, 3, 1], stride=[1, 2, 2, 1])
        x = L.FC_layer("local3", x, 384)
        x = L.FC_layer("local4", x, 192)
        return L.fc_nonacti("softmax_linear", x, self.num_classes)


class Cifar10SyntheticLoader(SyntheticDataLoader):
    image_shape = (32, 32, 3)
    num_classes = 10
