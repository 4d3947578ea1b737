"""Kernel-level timing (torch.profiler) of one conv call: python scripts/prof_one.py B H Ci Co k s [fprop|dgrad|wgrad]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
from distributed_tensorflow_framework_b200.ops import raw
B, H, Ci, Co, k, s = [int(v) for v in sys.argv[1:7]]
mode = sys.argv[7] if len(sys.argv) > 7 else "fprop"
x = torch.randn(B, H, H, Ci, device="cuda").to(torch.bfloat16)
w = (torch.randn(k, k, Ci, Co, device="cuda") * 0.05).to(torch.bfloat16)
w2 = w.reshape(k * k * Ci, Co); wt = w.permute(0, 1, 3, 2).contiguous().reshape(k * k * Co, Ci)
y = raw.conv_fprop(x, w2, None, k, k, s); dy = torch.randn_like(y); dw = torch.zeros(k * k * Ci, Co, device="cuda")
def run():
    if mode == "fprop": raw.conv_fprop(x, w2, None, k, k, s)
    elif mode == "dgrad": raw.conv_dgrad(dy, wt, (H, H), k, k, s)
    else: raw.conv_wgrad(x, dy, k, k, s, out=dw)
for _ in range(3): run()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(5): run()
    torch.cuda.synchronize()
for e in sorted(prof.key_averages(), key=lambda e: -e.device_time_total)[:8]:
    print("%9.1f us x%d  %s" % (e.device_time_total / e.count, e.count // 5, e.key[:90]))
