"""Per-layer conv timing (ResNet-50 shapes, batch 64): ours vs cuDNN through torch. CUDA events, L2-cold-ish."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from distributed_tensorflow_framework_b200.ops import raw
torch.backends.cudnn.benchmark = True
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
shapes = [  # H, Cin, Cout, k, stride, count
    (224, 8, 64, 7, 2, 1), (56, 64, 64, 1, 1, 1), (56, 64, 64, 3, 1, 3), (56, 64, 256, 1, 1, 4), (56, 256, 64, 1, 1, 2),
    (56, 256, 128, 1, 1, 1), (56, 128, 128, 3, 2, 1), (28, 128, 512, 1, 1, 4), (56, 256, 512, 1, 2, 1),
    (28, 512, 128, 1, 1, 3), (28, 128, 128, 3, 1, 3), (28, 512, 256, 1, 1, 1), (28, 256, 256, 3, 2, 1),
    (14, 256, 1024, 1, 1, 6), (28, 512, 1024, 1, 2, 1), (14, 1024, 256, 1, 1, 5), (14, 256, 256, 3, 1, 5),
    (14, 1024, 512, 1, 1, 1), (14, 512, 512, 3, 2, 1), (7, 512, 2048, 1, 1, 3), (14, 1024, 2048, 1, 2, 1),
    (7, 2048, 512, 1, 1, 2), (7, 512, 512, 3, 1, 2),
]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def timeit(fn, iters=5):
    fn(); fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        flush.fill_(1)
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / iters
tot = {"of": 0, "od": 0, "ow": 0, "cf": 0, "cd": 0, "cw": 0}
print("%-28s %8s %8s %8s | %8s %8s %8s | TFLOP/s ours f/d/w" % ("layer", "our_f", "our_d", "our_w", "cud_f", "cud_d", "cud_w"))
for (H, Ci, Co, k, s, cnt) in shapes:
    x = torch.randn(B, H, H, Ci, device="cuda").to(torch.bfloat16)
    w = (torch.randn(k, k, Ci, Co, device="cuda") * 0.05).to(torch.bfloat16)
    w2 = w.reshape(k * k * Ci, Co); wt = w.permute(0, 1, 3, 2).contiguous().reshape(k * k * Co, Ci)
    y = raw.conv_fprop(x, w2, None, k, k, s)
    dy = torch.randn_like(y)
    dw = torch.zeros(k * k * Ci, Co, device="cuda")
    tf = timeit(lambda: raw.conv_fprop(x, w2, None, k, k, s))
    td = timeit(lambda: raw.conv_dgrad(dy, wt, (H, H), k, k, s))
    tw = timeit(lambda: raw.conv_wgrad(x, dy, k, k, s, out=dw))
    xc = x.permute(0, 3, 1, 2); wc = w.permute(3, 2, 0, 1).contiguous(memory_format=torch.channels_last)
    dyc = dy.permute(0, 3, 1, 2)
    pad = k // 2
    cf = timeit(lambda: F.conv2d(xc, wc, stride=s, padding=pad))
    cd = timeit(lambda: torch.ops.aten.convolution_backward(dyc, xc, wc, None, [s, s], [pad, pad], [1, 1], False, [0, 0], 1, [True, False, False]))
    cw = timeit(lambda: torch.ops.aten.convolution_backward(dyc, xc, wc, None, [s, s], [pad, pad], [1, 1], False, [0, 0], 1, [False, True, False]))
    fl = 2.0 * B * y.shape[1] * y.shape[2] * Co * k * k * Ci / 1e9
    print("%3dx%-3d %4d->%-4d k%d s%d x%d  %8.3f %8.3f %8.3f | %8.3f %8.3f %8.3f | %5.0f %5.0f %5.0f" % (H, H, Ci, Co, k, s, cnt, tf, td, tw, cf, cd, cw, fl / tf, fl / td, fl / tw))
    for key, v in zip(("of", "od", "ow", "cf", "cd", "cw"), (tf, td, tw, cf, cd, cw)):
        tot[key] += v * cnt
print("TOTAL (weighted by layer count, ms): ours f %.2f d %.2f w %.2f = %.2f | cudnn f %.2f d %.2f w %.2f = %.2f" % (
    tot["of"], tot["od"], tot["ow"], tot["of"] + tot["od"] + tot["ow"], tot["cf"], tot["cd"], tot["cw"], tot["cf"] + tot["cd"] + tot["cw"]))
