"""GPU numerics check of the tcgen05 GEMM / implicit-GEMM conv kernels against torch fp32 references."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from distributed_tensorflow_framework_b200.ops import raw

torch.manual_seed(0)
dev = "cuda"
fails = 0


def report(name, got, ref, tol=2e-2):
    global fails
    got = got.float(); ref = ref.float()
    err = (got - ref).abs().max().item()
    scale = ref.abs().max().item() + 1e-6
    bad = not (err / scale < tol) or torch.isnan(got).any().item()
    if bad:
        fails += 1
    print("%-52s max_err %.4e  ref_max %.3e  rel %.3e  %s" % (name, err, scale, err / scale, "FAIL" if bad else "ok"), flush=True)


def rnd(*shape):
    return torch.randn(*shape, device=dev).to(torch.bfloat16)


def ref_conv(x, w_hwio, stride, KH, KW):
    n, H, W, C = x.shape
    OH, pt = raw.same_pad(H, KH, stride); OW, pl = raw.same_pad(W, KW, stride)
    pb = max((OH - 1) * stride + KH - H, 0) - pt; pr = max((OW - 1) * stride + KW - W, 0) - pl
    xp = F.pad(x.float().permute(0, 3, 1, 2), (pl, pr, pt, pb))
    return F.conv2d(xp, w_hwio.float().permute(3, 2, 0, 1), stride=stride)  # NCHW


# ---------------- dense GEMM
for (M, N, K) in [(128, 128, 64), (256, 128, 256), (200, 192, 136), (128, 384, 4096), (1024, 64, 576), (4096, 256, 128)]:
    a = rnd(M, K); b = rnd(K, N); bias = torch.randn(N, device=dev)
    ref = a.float() @ b.float()
    report("gemm %dx%dx%d" % (M, N, K), raw.gemm(a, b), ref)
    report("gemm+bias+relu fp32out %dx%dx%d" % (M, N, K), raw.gemm(a, b, bias, relu=True, out_fp32=True), torch.relu(ref + bias))
    at = a.t().contiguous()
    report("gemm A^T %dx%dx%d" % (M, N, K), raw.gemm(at, b, a_transposed=True), ref)
    acc = torch.zeros(M, N, device=dev)
    raw.gemm(at, b, a_transposed=True, out=acc, accumulate=True)
    report("gemm A^T splitK atomic %dx%dx%d" % (M, N, K), acc, ref)
torch.cuda.synchronize()

# ---------------- conv family
cases = [  # n, H, W, Cin, Cout, KH, KW, stride
    (2, 16, 16, 64, 64, 3, 3, 1),
    (4, 32, 32, 8, 64, 5, 5, 1),
    (4, 16, 16, 64, 64, 5, 5, 1),
    (2, 28, 28, 128, 256, 3, 3, 2),
    (2, 14, 14, 256, 512, 1, 1, 1),
    (2, 56, 56, 64, 128, 1, 1, 2),
    (2, 30, 30, 16, 32, 7, 7, 2),
    (3, 17, 19, 32, 64, 3, 3, 2),
]
for (n, H, W, Ci, Co, KH, KW, s) in cases:
    tag = "n%d %dx%d c%d->%d k%dx%d s%d" % (n, H, W, Ci, Co, KH, KW, s)
    x = rnd(n, H, W, Ci); w = (torch.randn(KH, KW, Ci, Co, device=dev) * 0.05).to(torch.bfloat16)
    bias = torch.randn(Co, device=dev)
    xf = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    wf = w.float().permute(3, 2, 0, 1).requires_grad_(True)
    OH, pt = raw.same_pad(H, KH, s); OW, pl = raw.same_pad(W, KW, s)
    pb = max((OH - 1) * s + KH - H, 0) - pt; pr = max((OW - 1) * s + KW - W, 0) - pl
    yref = F.conv2d(F.pad(xf, (pl, pr, pt, pb)), wf, stride=s)
    y = raw.conv_fprop(x, w.reshape(KH * KW * Ci, Co), bias, KH, KW, s, relu=True)
    report("fprop " + tag, y, torch.relu(yref + bias.view(1, -1, 1, 1)).permute(0, 2, 3, 1))
    dy = rnd(n, OH, OW, Co)
    yref.backward(dy.float().permute(0, 3, 1, 2))
    w_t = w.permute(0, 1, 3, 2).contiguous().reshape(KH * KW * Co, Ci)
    dx = raw.conv_dgrad(dy, w_t, (H, W), KH, KW, s)
    report("dgrad " + tag, dx, xf.grad.permute(0, 2, 3, 1))
    dw = raw.conv_wgrad(x, dy, KH, KW, s)
    report("wgrad " + tag, dw.view(KH, KW, Ci, Co), wf.grad.permute(2, 3, 1, 0), tol=3e-2)
torch.cuda.synchronize()

# ---------------- timing (large GEMM and a ResNet-like conv), CUDA events
def timeit(fn, iters=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters

M = N = K = 4096
a = rnd(M, K); b = rnd(K, N)
t = timeit(lambda: raw.gemm(a, b))
print("gemm 4096^3: %.3f ms  %.1f TFLOP/s   (torch.matmul: %.3f ms)" % (t, 2 * M * N * K / t / 1e9, timeit(lambda: a @ b)))
x = rnd(128, 56, 56, 64); w = (torch.randn(3, 3, 64, 64, device=dev) * 0.05).to(torch.bfloat16)
w2 = w.reshape(576, 64)
t = timeit(lambda: raw.conv_fprop(x, w2, None, 3, 3, 1))
fl = 2 * 128 * 56 * 56 * 64 * 576
xc = x.permute(0, 3, 1, 2).contiguous(memory_format=torch.channels_last); wc = w.permute(3, 2, 0, 1).contiguous(memory_format=torch.channels_last)
tt = timeit(lambda: F.conv2d(xc, wc, padding=1))
print("conv3x3 128x56x56x64->64 fprop: %.3f ms %.1f TFLOP/s (cudnn via torch: %.3f ms)" % (t, fl / t / 1e9, tt))
dy = rnd(128, 56, 56, 64)
t = timeit(lambda: raw.conv_wgrad(x, dy, 3, 3, 1))
print("conv3x3 wgrad: %.3f ms %.1f TFLOP/s" % (t, fl / t / 1e9))
print("FAILS", fails)
sys.exit(1 if fails else 0)
