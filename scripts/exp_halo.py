import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_framework_b200.ops import raw, _lib
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
for (B, H, Ci, Co) in [(256, 56, 64, 64), (256, 14, 256, 256), (256, 28, 128, 128)]:
    x = torch.randn(B, H, H, Ci, device="cuda").to(torch.bfloat16)
    w = (torch.randn(3, 3, Ci, Co, device="cuda") * 0.05).to(torch.bfloat16).reshape(9 * Ci, Co)
    out = torch.empty(B, H, H, Co, device="cuda", dtype=torch.bfloat16)
    for dbg, name in ((0, "normal"), (2, "no tap shift"), (4, "no stores"), (8, "no A loads"), (14, "all three")):
        t = timeit(lambda: _lib.call("dtf_conv_halo_bf16", x, w, out, None, B, H, H, Ci, 3, 3, Co, 1, 1, 0, dbg << 8, raw._sp()))
        print("b%d %dx%d %d->%d  %-14s %.3f ms" % (B, H, H, Ci, Co, name, t), flush=True)
