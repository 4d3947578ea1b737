"""Per-variable gradient comparison of one step on the CUDA kernels against the CPU fp32 oracle (identical weights)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import distributed_tensorflow_framework_b200 as dtf
from distributed_tensorflow_framework_b200 import models

name = sys.argv[1] if len(sys.argv) > 1 else "ResNet18"
batch = int(sys.argv[2]) if len(sys.argv) > 2 else 8
size = int(sys.argv[3]) if len(sys.argv) > 3 else 64
torch.manual_seed(4)
cls = getattr(models, name)
ncls = 10 if name == "Cifar10CNN" else 16
if name != "Cifar10CNN":
    class Small(cls):
        num_classes = 16
else:
    Small, size = cls, 32
x = torch.rand(batch, size, size, 3).to(torch.bfloat16).float()
y = torch.randint(0, ncls, (batch,))
mk = lambda d: dtf.Trainer(dtf.Net(Small()), dtf.SoftmaxCrossEntropyLoss(), dtf.GradientDescentOptimizer(0.0), device=d)
cpu, gpu = mk("cpu"), mk("cuda:0")
cpu.build(x, y); gpu.build(x, y)
gpu.store.load_state_dict(cpu.store.state_dict())
gpu.store.shadow_arena.copy_(gpu.store.master_arena)
gpu.store.refresh_transposed_shadows()
with torch.no_grad():
    for v in cpu.store.trainable_variables():
        if v.kind in ("conv_w", "fc_w"):
            v.master.copy_(v.master.to(torch.bfloat16).float())
from distributed_tensorflow_framework_b200.engine import layers
layers.emulate_bf16(os.environ.get("EMU", "1") == "1")
lc, tg = cpu._forward_backward([(x, y)])
layers.emulate_bf16(False)
cg = {v.name: g for g, v in tg[0]}
gpu.engine.begin_step(1)
lg, _ = gpu._forward_backward([(x.cuda(), y.cuda())])
torch.cuda.synchronize()
print("loss cpu %.6f gpu %.6f" % (float(lc), float(lg)))
for v in gpu.store.trainable_variables():
    gc = cg[v.name].float(); gg = v.logical(v.grad.detach()).float().cpu()
    if float(gc.norm()) < 1e-9:
        continue
    rel = float((gg - gc).norm() / gc.norm())
    cos = float(torch.dot(gg.flatten(), gc.flatten()) / (gg.norm() * gc.norm() + 1e-20))
    print("%-44s rel %.4f  |g|gpu/|g|cpu %.4f  cos %.4f  n=%d" % (v.name, rel, float(gg.norm() / gc.norm()), cos, gc.numel()))
