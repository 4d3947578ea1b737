"""One convolution shape, forward / data gradient / filter gradient, one warm-up and one measured launch of each: the
harness for `ncu --set full --profile-from-start off -k regex:'halo|igemm'` captures (the measured launches only).
usage: one_conv_shape.py H Cin Cout k stride [batch]   (several shapes: separate them with '/')"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_framework_b200.ops import raw


def run(H, Ci, Co, k, s, B):
    x = torch.randn(B, H, H, Ci, device="cuda").to(torch.bfloat16)
    w = (torch.randn(k, k, Ci, Co, device="cuda") * 0.05).to(torch.bfloat16)
    w2 = w.reshape(k * k * Ci, Co)
    wt = w.permute(0, 1, 3, 2).contiguous().reshape(k * k * Co, Ci)
    y = raw.conv_fprop(x, w2, None, k, k, s)
    dy = torch.randn_like(y)
    dw = torch.zeros(k * k * Ci, Co, device="cuda")
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    for it in range(2):
        torch.cuda.synchronize()
        if it == 1:
            torch.cuda.profiler.start()  # ncu --profile-from-start off: only the warm iteration is captured
        ev[0].record()
        raw.conv_fprop(x, w2, None, k, k, s)
        ev[1].record()
        raw.conv_dgrad(dy, wt, (H, H), k, k, s)
        ev[2].record()
        raw.conv_wgrad(x, dy, k, k, s, out=dw)
        ev[3].record()
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    fl = 2.0 * B * y.shape[1] * y.shape[2] * Co * k * k * Ci / 1e9
    t = [ev[i].elapsed_time(ev[i + 1]) for i in range(3)]
    print("%dx%d %d->%d k%d s%d b%d: fprop %.3f dgrad %.3f wgrad %.3f ms (%.0f / %.0f / %.0f TFLOP/s, warm L2)" % (
        H, H, Ci, Co, k, s, B, t[0], t[1], t[2], fl / t[0], fl / t[1], fl / t[2]))


if __name__ == "__main__":
    specs = " ".join(sys.argv[1:]).split("/")
    for spec in specs:
        v = [int(a) for a in spec.split()]
        run(v[0], v[1], v[2], v[3], v[4], v[5] if len(v) > 5 else 256)
