"""Batch-norm forward + backward (residual + ReLU) on the ResNet-50 stage-1 tensor [256*56*56, 256] for ncu captures,
then one K-RS-OPT collective over a 25.6 M-parameter arena on a single-rank communicator (kernel structure only)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import distributed_tensorflow_framework_b200 as dtf
from distributed_tensorflow_framework_b200.ops import raw
from distributed_tensorflow_framework_b200.parallel.comm import Communicator, SymmetricHeap
R, N = 256 * 56 * 56, 256
x = torch.randn(R, N, device="cuda").to(torch.bfloat16); res = torch.randn(R, N, device="cuda").to(torch.bfloat16)
dy = torch.randn(R, N, device="cuda").to(torch.bfloat16)
gamma, beta = torch.ones(N, device="cuda"), torch.zeros(N, device="cuda")
for _ in range(2):
    y, mean, rstd, sc, sh, mask = raw.bn_train_fwd(x, gamma, beta, None, None, 1e-3, 0.1, res, True, want_mask=True)
    dx, dz, dg, db = raw.bn_bwd(dy, mask, x, mean, rstd, gamma, True)
    y2, mean2, rstd2, sc2, sh2 = raw.bn_train_fwd(x, gamma, beta, None, None, 1e-3, 0.1, None, True)
    dx2, _, _, _ = raw.bn_bwd(dy, None, x, mean2, rstd2, gamma, True, mask_scale=sc2, mask_shift=sh2)
torch.cuda.synchronize()
n = 25_600_000 // 64 * 64
heap = SymmetricHeap(n * 16 + (4 << 20), torch.device("cuda", 0))
grad = heap.alloc("grad", n, torch.float32); shadow = heap.alloc("shadow", n, torch.bfloat16)
master = heap.alloc("master", n, torch.float32); mom = heap.alloc("slot_m", n, torch.float32)
comm = Communicator(heap, 0, 1)
opt = dtf.MomentumOptimizer(0.01, 0.9)
for _ in range(2):
    comm.rs_optim_ag(heap.offset_of("grad"), [0, n], opt, 1, master, mom, None, heap.offset_of("shadow"))
torch.cuda.synchronize(); print("done")
