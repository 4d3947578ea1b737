"""One dense GEMM shape (default: the 56x56 64->256 1x1 convolution of ResNet-50 at batch 256) for ncu captures."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_framework_b200.ops import raw
M = int(sys.argv[1]) if len(sys.argv) > 1 else 802816
N = int(sys.argv[2]) if len(sys.argv) > 2 else 256
K = int(sys.argv[3]) if len(sys.argv) > 3 else 64
stats = len(sys.argv) > 4 and sys.argv[4] == "stats"
a = torch.randn(M, K, device="cuda").to(torch.bfloat16); b = (torch.randn(K, N, device="cuda") * 0.1).to(torch.bfloat16)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
ts = []
for _ in range(6):
    flush.fill_(1)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); raw.gemm(a, b, stats={} if stats else None); e1.record(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
print("M %d N %d K %d stats %s: %.4f ms (min of 6), %.2f TB/s" % (M, N, K, stats, min(ts), (M * K + M * N + K * N) * 2 / min(ts) / 1e9))
