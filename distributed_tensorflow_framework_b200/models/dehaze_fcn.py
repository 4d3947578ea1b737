"""Fully-convolutional image-to-image net in the style of the reference's implied GMAN dehazing model
("GMAN tower" /root/reference/distribute_tower.py:53; features hazed_image_raw / clear_image_raw
/root/reference/distribute.py:30-35; `deconv` default output [32,224,224,64] /root/reference/distribute_tools.py:131):
conv encoder with two stride-2 stages, residual trunk, two transposed-conv decoder stages, 3-channel output, MSE loss.
BASELINE.json configs[3] (256x256)."""
from ..engine import layers as L
from ..plugins.input import SyntheticDataLoader
from ..plugins.model import Model


class DehazeFCN(Model):
    width = 64
    trunk_blocks = 4
    out_channels = 3

    def inference(self, images, *args, **kwargs):
        n, h, w = images.shape[0], images.shape[1], images.shape[2]
        c = self.width
        x = L.conv("conv1", images, 3, c, kernel_size=[3, 3])
        e1 = L.conv("conv2", x, c, c, kernel_size=[3, 3])
        x = L.conv("down1", e1, c, 2 * c, kernel_size=[3, 3], stride=[1, 2, 2, 1])
        x = L.conv("down2", x, 2 * c, 4 * c, kernel_size=[3, 3], stride=[1, 2, 2, 1])
        for i in range(self.trunk_blocks):
            y = L.conv("res%d_a" % i, x, 4 * c, 4 * c, kernel_size=[3, 3])
            y = L.conv_nonacti("res%d_b" % i, y, 4 * c, 4 * c, kernel_size=[3, 3])
            x = L.add_relu(x, y, relu=True)
        h2, w2 = -(-h // 2), -(-w // 2)
        x = L.deconv("up1", x, 2 * c, 4 * c, output_shape=[n, h2, w2, 2 * c], stride=[1, 2, 2, 1])
        x = L.acti_layer(x)
        x = L.deconv("up2", x, c, 2 * c, output_shape=[n, h, w, c], stride=[1, 2, 2, 1])
        x = L.acti_layer(x)
        x = L.add_relu(x, e1, relu=False)
        x = L.conv("conv3", x, c, c, kernel_size=[3, 3])
        return L.conv_nonacti("out", x, c, self.out_channels, kernel_size=[3, 3])


class DehazeSyntheticLoader(SyntheticDataLoader):
    image_shape = (256, 256, 3)
    num_classes = 0
    target_shape = (256, 256, 3)
    pool_batches = 2
