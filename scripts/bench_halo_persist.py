"""Halo convolution: persistent kernel vs one tile per CTA on the 3x3 layers of ResNet-50 / the dehazing net.
Numerics (persistent == one-tile, bit for bit) and CUDA-event timing with an L2 flush between iterations."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_framework_b200.ops import raw
lib = raw._lib.lib()
shapes = [(256, 56, 64, 64), (256, 28, 128, 128), (256, 14, 256, 256), (256, 7, 512, 512), (8, 256, 64, 64),
          (8, 64, 256, 256), (3, 28, 128, 64), (2, 17, 64, 192), (5, 7, 64, 72)]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def timeit(fn, iters=5):
    fn(); fn(); torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        flush.fill_(1)
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / iters
for (B, H, Ci, Co) in shapes:
    x = torch.randn(B, H, H, Ci, device="cuda").to(torch.bfloat16)
    w = (torch.randn(3, 3, Ci, Co, device="cuda") * 0.05).to(torch.bfloat16).reshape(9 * Ci, Co)
    bias = torch.randn(Co, device="cuda")
    res = {}
    for mode in (2, 1, 0):
        lib.dtf_set_persistent_halo(mode)
        y = raw.conv_fprop(x, w, bias, 3, 3, 1, relu=True)
        t = timeit(lambda: raw.conv_fprop(x, w, bias, 3, 3, 1, relu=True))
        res[mode] = (y, t)
    lib.dtf_set_persistent_halo(2)
    same = torch.equal(res[0][0], res[1][0])
    err = (res[2][0].float() - res[0][0].float()).abs().max().item() / (res[0][0].float().abs().max().item() + 1e-6)
    fl = 2.0 * B * H * H * Co * 9 * Ci / 1e9
    print("b%-3d %3dx%-3d %4d->%-4d pixels-as-N %.3f ms (%5.0f TFLOP/s) relerr %.1e | persistent %.3f ms (%5.0f)  one-tile %.3f ms (%5.0f) equal=%s" % (
        B, H, H, Ci, Co, res[2][1], fl / res[2][1], err, res[1][1], fl / res[1][1], res[0][1], fl / res[0][1], same), flush=True)
