"""Kernel-time breakdown of one training step (torch.profiler / CUPTI; for analysis only, never a bench number)."""
import os, sys, argparse
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import distributed_tensorflow_framework_b200 as dtf
from distributed_tensorflow_framework_b200 import models
import bench

ap = argparse.ArgumentParser()
ap.add_argument("--config", default="resnet50")
ap.add_argument("--batch", type=int, default=0)
ap.add_argument("--steps", type=int, default=3)
ap.add_argument("--rows", type=int, default=40)
a = ap.parse_args()
cfg = bench.CONFIGS[a.config]
batch = a.batch or cfg["batch"]
loss = dtf.SoftmaxCrossEntropyLoss() if cfg["loss"] == "xent" else dtf.MSELoss()
tr = dtf.Trainer(dtf.Net(getattr(models, cfg["model"])()), loss, dtf.MomentumOptimizer(0.01, 0.9), device="cuda:0")
x, y = bench.make_batch(cfg, batch, 1)
xd, yd = x.cuda(), y.cuda()
for _ in range(3):
    tr.step_device([(xd, yd)])
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(a.steps):
        tr.step_device([(xd, yd)])
    torch.cuda.synchronize()
ev = prof.key_averages()
rows = sorted(ev, key=lambda e: -e.device_time_total)
tot = sum(e.device_time_total for e in ev)
print("total device time per step: %.3f ms" % (tot / a.steps / 1e3))
for e in rows[:a.rows]:
    print("%8.3f ms %5.1f%% n=%4d  %s" % (e.device_time_total / a.steps / 1e3, 100.0 * e.device_time_total / tot, e.count // a.steps, e.key[:110]))
