"""Layer-by-layer comparison (activations and data gradients) of the CUDA path against the bf16-emulating CPU oracle."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import distributed_tensorflow_framework_b200 as dtf
from distributed_tensorflow_framework_b200 import models
from distributed_tensorflow_framework_b200.engine import layers
from distributed_tensorflow_framework_b200.models import resnet as resnet_mod

name = sys.argv[1] if len(sys.argv) > 1 else "ResNet18"
batch = int(sys.argv[2]) if len(sys.argv) > 2 else 8
size = int(sys.argv[3]) if len(sys.argv) > 3 else 64
torch.manual_seed(4)
if name.startswith("Shallow"):
    class Small(resnet_mod._ResNet):
        num_classes = 16
        blocks = (1, 2)
        bottleneck = name.endswith("B")
else:
    cls = getattr(models, name)
    class Small(cls):
        num_classes = 16
resnet_mod.set_relu(os.environ.get("RELU", "1") == "1")
x = torch.rand(batch, size, size, 3).to(torch.bfloat16).float()
y = torch.randint(0, 16, (batch,))
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

rec = None
def wrap(fname):
    orig = getattr(layers, fname)
    def f(*a, **k):
        out = orig(*a, **k)
        t = out[0] if isinstance(out, tuple) else out
        tag = "%s:%s" % (fname, a[0] if isinstance(a[0], str) else "")
        from distributed_tensorflow_framework_b200.engine.variables import current_scope
        tag = current_scope() + "/" + tag
        idx = len(rec)
        rec.append([tag, t.detach().float().cpu(), None])
        if t.requires_grad:
            t.register_hook(lambda g, i=idx: rec[i].__setitem__(2, g.detach().float().cpu()))
        return out
    setattr(layers, fname, f)
    setattr(resnet_mod.L, fname, f)
for fn in ("conv_nonacti", "batch_norm_channel", "pool", "global_avg_pool", "fc_nonacti"):
    wrap(fn)

rec = []
layers.emulate_bf16(True)
lc, tgc = cpu._forward_backward([(x, y)])
layers.emulate_bf16(False)
rc = rec
rec = []
gpu.engine.begin_step(1)
lg, _ = gpu._forward_backward([(x.cuda(), y.cuda())])
torch.cuda.synchronize()
rg = rec
print("loss", float(lc), float(lg))
def rel(a, b):
    if a is None or b is None:
        return float("nan")
    if a.shape != b.shape:
        b = b[..., :a.shape[-1]] if b.shape[-1] > a.shape[-1] else b
    return float((a - b).norm() / (a.norm() + 1e-20))
for (t1, a1, g1), (t2, a2, g2) in zip(rc, rg):
    print("%-60s act %.5f  grad %.5f" % (t1, rel(a1, a2), rel(g1, g2)))

print("---- per variable")
cg = {v.name: g for g, v in tgc[0]}
for v in gpu.store.trainable_variables():
    gc = cg[v.name].float(); gg = v.logical(v.grad.detach()).float().cpu()
    if float(gc.norm()) < 1e-9:
        continue
    print("%-44s rel %.4f  ratio %.4f  cos %.5f" % (v.name, float((gg - gc).norm() / gc.norm()), float(gg.norm() / gc.norm()),
          float(torch.dot(gg.flatten(), gc.flatten()) / (gg.norm() * gc.norm() + 1e-20))))
