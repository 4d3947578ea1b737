"""ResNet-18 / ResNet-50 (v1.5: stride on the 3x3 of the bottleneck) on the framework's layer library, NHWC,
224x224x3 inputs -- BASELINE.json configs[2] (ResNet-50 bf16 sync data-parallel) and [4] (ResNet-18 async ps).
The reference ships no ResNet; these are example plugins in its Model API (/root/reference/distribute_model.py:13-22)."""
from ..engine import layers as L
from ..engine.variables import variable_scope
from ..plugins.input import SyntheticDataLoader
from ..plugins.model import Model


# Batch-norm statistics can be produced by the epilogue of the kernel that writes the tensor instead of a separate
# pass over it. 1 (default): the 1x1 convolutions (dense persistent GEMM: per-warp column sums kept in shared memory
# across the CTA's tiles, accumulated while the staged tile is written out). 2: also the KxK halo convolutions, whose
# epilogue statistics were measured NOT to pay on ResNet-50 (25.8 ms/step with, 25.3 without). 0: never.
import os as _os

EPILOGUE_BN_STATS = int(_os.environ.get("DTF_EPILOGUE_STATS", "1"))


_RELU = True


def set_relu(flag):
    """Test hook: build the networks without activations (a linear conv / batch-norm / residual stack), so that a
    gradient comparison against the CPU oracle is not blurred by ReLU decisions. Returns the previous setting."""
    global _RELU
    old, _RELU = _RELU, bool(flag)
    return old


def _conv_only(name, x, cin, cout, k, stride):
    """A convolution whose batch norm is applied by its consumer (projection shortcut): (raw output, statistics)."""
    emit = EPILOGUE_BN_STATS >= 2 or (EPILOGUE_BN_STATS == 1 and k == 1)
    if emit:
        y, _, part = L.conv_nonacti(name, x, cin, cout, kernel_size=[k, k], stride=[1, stride, stride, 1],
                                    use_bias=False, emit_stats=True)
        return y, part
    return L.conv_nonacti(name, x, cin, cout, kernel_size=[k, k], stride=[1, stride, stride, 1], use_bias=False), None


def _conv_bn(name, x, cin, cout, k, stride, relu=True, residual=None, fork_input=False, residual_bn=None,
             residual_part=None):
    relu = relu and _RELU
    emit = EPILOGUE_BN_STATS >= 2 or (EPILOGUE_BN_STATS == 1 and k == 1)
    if fork_input or emit:
        y, x_fwd, part = L.conv_nonacti(name, x, cin, cout, kernel_size=[k, k], stride=[1, stride, stride, 1],
                                        use_bias=False, fork_input=fork_input, emit_stats=emit)
    else:
        y, x_fwd, part = L.conv_nonacti(name, x, cin, cout, kernel_size=[k, k], stride=[1, stride, stride, 1],
                                        use_bias=False), None, None
    out = L.batch_norm_channel(name + "/bn", y, relu=relu, residual=residual, stats_part=part,
                               residual_bn=residual_bn, residual_part=residual_part)
    return (out, x_fwd) if fork_input else out  # x_fwd: x forwarded for the shortcut branch


def _basic_block(name, x, cin, cout, stride):
    with variable_scope(name):
        # conv1 forwards x to the shortcut branch so that both gradients of x meet inside conv1's backward
        y, shortcut = _conv_bn("conv1", x, cin, cout, 3, stride, fork_input=True)
        if stride != 1 or cin != cout:
            # projection shortcut: its batch norm is applied inside conv2's closing batch norm (one pass over both)
            raw_s, part_s = _conv_only("shortcut", shortcut, cin, cout, 1, stride)
            return _conv_bn("conv2", y, cout, cout, 3, 1, relu=True, residual=raw_s, residual_bn="shortcut/bn",
                            residual_part=part_s)
        return _conv_bn("conv2", y, cout, cout, 3, 1, relu=True, residual=shortcut)


def _bottleneck(name, x, cin, mid, stride):
    cout = mid * 4
    with variable_scope(name):
        y, shortcut = _conv_bn("conv1", x, cin, mid, 1, 1, fork_input=True)
        y = _conv_bn("conv2", y, mid, mid, 3, stride)
        if stride != 1 or cin != cout:
            raw_s, part_s = _conv_only("shortcut", shortcut, cin, cout, 1, stride)
            return _conv_bn("conv3", y, mid, cout, 1, 1, relu=True, residual=raw_s, residual_bn="shortcut/bn",
                            residual_part=part_s)
        return _conv_bn("conv3", y, mid, cout, 1, 1, relu=True, residual=shortcut)


class _ResNet(Model):
    num_classes = 1000
    blocks = (2, 2, 2, 2)
    bottleneck = False

    def inference(self, images, *args, **kwargs):
        if _RELU:  # stem: batch norm + ReLU are applied inside the max-pool kernel
            y, part = _conv_only("conv1", images, 3, 64, 7, 2)
            x = L.batch_norm_relu_pool("conv1/bn", "pool1", y, kernel=[1, 3, 3, 1], stride=[1, 2, 2, 1],
                                       stats_part=part)
        else:
            x = _conv_bn("conv1", images, 3, 64, 7, 2)
            x = L.pool("pool1", x, kernel=[1, 3, 3, 1], stride=[1, 2, 2, 1])
        cin = 64
        for stage, n in enumerate(self.blocks):
            width = 64 * (2 ** stage)
            for b in range(n):
                stride = 2 if (b == 0 and stage > 0) else 1
                name = "stage%d/block%d" % (stage + 1, b)
                if self.bottleneck:
                    x = _bottleneck(name, x, cin, width, stride)
                    cin = width * 4
                else:
                    x = _basic_block(name, x, cin, width, stride)
                    cin = width
        x = L.global_avg_pool(x)
        return L.fc_nonacti("logits", x, self.num_classes)


class ResNet18(_ResNet):
    blocks = (2, 2, 2, 2)
    bottleneck = False


class ResNet50(_ResNet):
    blocks = (3, 4, 6, 3)
    bottleneck = True


class ImageNetSyntheticLoader(SyntheticDataLoader):
    image_shape = (224, 224, 3)
    num_classes = 1000
    pool_batches = 2
