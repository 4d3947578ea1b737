import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_framework_b200.ops import raw
M = N = K = 4096
a = torch.randn(M, K, device="cuda").to(torch.bfloat16); b = torch.randn(K, N, device="cuda").to(torch.bfloat16)
for _ in range(4):
    raw.gemm(a, b)
torch.cuda.synchronize(); print("done")
