import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_framework_b200.ops import raw
M, N, K = [int(v) for v in sys.argv[1:4]] if len(sys.argv) > 3 else (200704, 64, 256)
a = torch.randn(M, K, device="cuda").to(torch.bfloat16); b = torch.randn(K, N, device="cuda").to(torch.bfloat16)
for _ in range(4):
    raw.gemm(a, b)
torch.cuda.synchronize(); print("done")
