import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline"))
import torch, torch.nn.functional as F
import torch_baseline as tb
torch.backends.cudnn.benchmark = True
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
model = tb.build("ResNet50").cuda().to(memory_format=torch.channels_last)
opt = torch.optim.SGD(model.parameters(), lr=0.01, momentum=0.9)
x = torch.rand(B, 3, 224, 224, device="cuda").contiguous(memory_format=torch.channels_last); y = torch.randint(0, 1000, (B,), device="cuda")
def step():
    with torch.autocast("cuda", dtype=torch.bfloat16):
        loss = F.cross_entropy(model(x).float(), y)
    opt.zero_grad(set_to_none=False); loss.backward(); opt.step()
for _ in range(5): step()
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(3): step()
    torch.cuda.synchronize()
ev = prof.key_averages(); tot = sum(e.device_time_total for e in ev)
print("torch baseline: total device time per step %.3f ms" % (tot / 3 / 1e3))
groups = {}
for e in ev:
    k = e.key
    g = "conv/gemm" if any(t in k for t in ("cudnn", "cutlass", "gemm", "xmma", "nvjet", "conv", "sm90", "sm100", "implicit")) else ("bn" if "batch_norm" in k or "bn_" in k else "other")
    groups[g] = groups.get(g, 0) + e.device_time_total
for g, v in sorted(groups.items(), key=lambda t: -t[1]): print("  %-10s %.3f ms" % (g, v / 3 / 1e3))
for e in sorted(ev, key=lambda e: -e.device_time_total)[:14]:
    print("%8.3f ms n=%4d %s" % (e.device_time_total / 3 / 1e3, e.count // 3, e.key[:100]))
