import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import distributed_tensorflow_framework_b200 as dtf
from distributed_tensorflow_framework_b200 import models
torch.manual_seed(1)
x = torch.rand(64, 32, 32, 3).pin_memory(); y = torch.randint(0, 10, (64,)).pin_memory()
def mk(g): return dtf.Trainer(dtf.Net(models.Cifar10CNN()), dtf.SoftmaxCrossEntropyLoss(), dtf.MomentumOptimizer(0.02, 0.9), device="cuda:0", seed=3, use_cuda_graph=g)
e1, e2, g = mk(False), mk(False), mk(True)
l1 = [float(e1.step((x, y))) for _ in range(40)]
l2 = [float(e2.step((x, y))) for _ in range(40)]
lg = [float(g.step((x, y))) for _ in range(40)]
for i in range(0, 40, 3):
    print(i, "%.4f %.4f %.4f" % (l1[i], l2[i], lg[i]))
print("global_step", e1.store.global_step, g.store.global_step)
