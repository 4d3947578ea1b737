import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_framework_b200.ops import raw, _lib
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def timeit(fn, iters=5):
    fn(); fn(); torch.cuda.synchronize(); tot = 0.0
    for _ in range(iters):
        flush.fill_(1)
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); tot += e0.elapsed_time(e1)
    return tot / iters
SHAPES = [(802816, 256), (802816, 64), (200704, 512), (50176, 1024), (12544, 2048)]
if len(sys.argv) > 2:
    _lib.lib().dtf_set_tuning(0, int(sys.argv[1])); _lib.lib().dtf_set_tuning(1, int(sys.argv[2]))
    print("tuning: colreduce blocks/SM %s, elementwise blocks/SM %s" % (sys.argv[1], sys.argv[2]))
for (R, N) in SHAPES:
    x = torch.randn(R, N, device="cuda").to(torch.bfloat16); dy = torch.randn_like(x); y = torch.relu(x); res = torch.randn_like(x)
    nb = x.numel() * 2 / 1e6
    stats = torch.zeros(_lib.lib().dtf_colreduce_blocks(R, N), 2, N, device="cuda"); mean = torch.zeros(N, device="cuda"); rstd = torch.ones(N, device="cuda")
    scale = torch.ones(N, device="cuda"); shift = torch.zeros(N, device="cuda"); out = torch.empty_like(x); coef = torch.empty(3, N, device="cuda")
    sp = _lib.stream_ptr
    t1 = timeit(lambda: _lib.call("dtf_bn_stats", x, stats, R, N, sp()))
    t2 = timeit(lambda: _lib.call("dtf_bn_apply", x, res, out, scale, shift, R, N, 1, sp()))
    t3 = timeit(lambda: _lib.call("dtf_bn_bwd_reduce", dy, None, x, None, stats, mean, rstd, scale, shift, R, N, 2, sp()))
    t4 = timeit(lambda: _lib.call("dtf_bn_bwd_apply", dy, x, out, stats, mean, rstd, None, coef, None, None, scale, shift, R, N, sp()))
    t5 = timeit(lambda: out.copy_(x))
    print("R=%7d N=%5d (%6.1f MB)  stats %.3f ms %5.0f GB/s | apply %.3f %5.0f | bwd_reduce %.3f %5.0f | bwd_apply %.3f %5.0f | torch copy %.3f %5.0f" % (
        R, N, nb, t1, nb / t1, t2, 3 * nb / t2, t3, 2 * nb / t3, t4, 3 * nb / t4, t5, 2 * nb / t5))
