"""Convolution / deconvolution variants a user of the reference's layer library may write: rectangular and even
kernels, kernels larger than the image, strides 1-4, batch sizes 1 / 5 / 7, rectangular images. Each case is a small
model (1x1 'pre' convolution -> the layer) compared with the CPU path for one training step
(tests/scenarios._one_step_parity). One line per case; failures are reported, not raised."""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import distributed_tensorflow_framework_b200 as dtf  # noqa: E402
import scenarios  # noqa: E402

T = dtf.tools
CASES = []
for k in ([3, 5], [1, 7], [7, 1], [2, 2], [4, 4], [5, 5], [9, 9], [11, 11]):
    for s in (1, 2):
        CASES.append(("conv k%dx%d s%d" % (k[0], k[1], s),
                      lambda x, k=k, s=s: T.conv("c", x, x.shape[-1], 16, kernel_size=k, stride=[1, s, s, 1])))
for s in (3, 4):
    CASES.append(("conv k3x3 s%d" % s, lambda x, s=s: T.conv("c", x, x.shape[-1], 16, stride=[1, s, s, 1])))
    CASES.append(("conv k1x1 s%d" % s, lambda x, s=s: T.conv("c", x, x.shape[-1], 16, kernel_size=[1, 1],
                                                            stride=[1, s, s, 1])))
for k in ([2, 2], [3, 3], [4, 4], [5, 5]):
    for s in (1, 2, 3):
        CASES.append(("deconv k%dx%d s%d" % (k[0], k[1], s),
                      lambda x, k=k, s=s: T.deconv("d", x, 8, x.shape[-1], kernel_size=k, stride=[1, s, s, 1],
                                                   output_shape=[x.shape[0], s * x.shape[1], s * x.shape[2], 8])))
CASES.append(("deconv k3x3 s2 odd out", lambda x: T.deconv("d", x, 8, x.shape[-1], stride=[1, 2, 2, 1],
              output_shape=[x.shape[0], 2 * x.shape[1] - 1, 2 * x.shape[2] - 1, 8])))
for k, s in (([2, 2], 2), ([3, 3], 2), ([3, 3], 1), ([2, 2], 1), ([5, 5], 3), ([2, 3], 2)):
    for mx in (True, False):
        CASES.append(("pool %s k%dx%d s%d" % ("max" if mx else "avg", k[0], k[1], s),
                      lambda x, k=k, s=s, mx=mx: T.pool("p", x, kernel=[1, k[0], k[1], 1], stride=[1, s, s, 1],
                                                        is_max_pool=mx)))

bad = 0
for (n, h, w, C) in ((1, 6, 10, 16), (5, 3, 3, 8), (7, 17, 9, 24)):
    for name, fn in CASES:
        class M(dtf.Model):
            def inference(self, x, *a, **k):
                return fn(T.conv_nonacti("pre", x, 3, C, kernel_size=[1, 1]))
        try:
            torch.manual_seed(0)
            x = torch.rand(n, h, w, 3)
            probe = dtf.Trainer(dtf.Net(M()), dtf.MSELoss(), dtf.GradientDescentOptimizer(0.05), device="cpu")
            probe.build(x, None)
            shape = tuple(probe.forward(x).shape)
            if not torch.cuda.is_available():
                probe.step((x, torch.rand(*shape)))
                print("%-24s b%d %dx%dx%d cpu ok, out %s" % (name, n, h, w, C, shape), flush=True)
                continue
            k = scenarios._one_step_parity(M, dtf.MSELoss(), x, torch.rand(*shape))
            print("%-24s b%d %dx%dx%d ok   (%d vars, out %s)" % (name, n, h, w, C, k, shape), flush=True)
        except Exception as e:  # noqa
            bad += 1
            print("%-24s b%d %dx%dx%d FAIL %s: %s" % (name, n, h, w, C, type(e).__name__,
                                                      str(e)[:160].replace("\n", " ")), flush=True)
print("bad cases:", bad)
