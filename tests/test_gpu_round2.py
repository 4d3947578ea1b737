"""Single-GPU tests of the round-2 engine paths: device-resident learning rate / Adam bias correction under CUDA-graph
replay, capture without side effects, staged H2D with prefetch, and a per-variable gradient comparison of one
ResNet step against the CPU fp32 oracle."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _trainer(model_cls, device, opt, **kw):
    import distributed_tensorflow_framework_b200 as dtf
    return dtf.Trainer(dtf.Net(model_cls()), dtf.SoftmaxCrossEntropyLoss(), opt, device=device, **kw)


def _batch(n=32, seed=0):
    g = torch.Generator().manual_seed(seed)
    return torch.rand(n, 32, 32, 3, generator=g).pin_memory(), torch.randint(0, 10, (n,), generator=g).pin_memory()


def test_adam_under_graph_replay_matches_eager():
    import distributed_tensorflow_framework_b200 as dtf
    from distributed_tensorflow_framework_b200 import models
    x, y = _batch()
    eager = _trainer(models.Cifar10CNN, "cuda:0", dtf.AdamOptimizer(1e-3), seed=3)
    graph = _trainer(models.Cifar10CNN, "cuda:0", dtf.AdamOptimizer(1e-3), seed=3, use_cuda_graph=True)
    le = [float(eager.step((x, y))) for _ in range(12)]
    lg = [float(graph.step((x, y))) for _ in range(12)]
    assert graph._graph is not None and graph.store.global_step == 12 == eager.store.global_step
    assert max(abs(a - b) for a, b in zip(le, lg)) < 5e-2, (le, lg)
    # bias correction follows the real step: a frozen lr_t (step 1: lr * sqrt(1-b2)/(1-b1) = 0.316 lr ... ) would
    # leave the two runs far apart after 12 steps
    d = (eager.store.master_arena - graph.store.master_arena).abs().mean()
    assert float(d) < 2e-4, float(d)
    hyper = graph.engine.hyper.cpu()
    assert int(hyper[2:3].view(torch.int32)) == 12
    import math
    want = 1e-3 * math.sqrt(1 - 0.999 ** 12) / (1 - 0.9 ** 12)
    assert abs(float(hyper[1]) - want) < 1e-8 * 1e3


def test_learning_rate_schedule_takes_effect_under_graph_replay():
    import distributed_tensorflow_framework_b200 as dtf
    from distributed_tensorflow_framework_b200 import models

    class Sched(object):
        learning_rate = 0.05

    sched = Sched()
    x, y = _batch(seed=1)
    tr = _trainer(models.Cifar10CNN, "cuda:0", dtf.GradientDescentOptimizer(sched), use_cuda_graph=True)
    for _ in range(3):
        tr.step((x, y))
    torch.cuda.synchronize()
    w0 = tr.store.master_arena.clone()
    tr.step((x, y))
    torch.cuda.synchronize()
    w1 = tr.store.master_arena.clone()
    assert float((w1 - w0).abs().max()) > 0
    sched.learning_rate = 0.0  # the captured graph must read the new rate from device memory
    tr.step((x, y))
    torch.cuda.synchronize()
    assert torch.equal(tr.store.master_arena, w1)
    sched.learning_rate = 0.05
    tr.step((x, y))
    torch.cuda.synchronize()
    assert float((tr.store.master_arena - w1).abs().max()) > 0


def test_capture_applies_the_first_batch_once():
    import distributed_tensorflow_framework_b200 as dtf
    from distributed_tensorflow_framework_b200 import models
    x, y = _batch(seed=2)
    eager = _trainer(models.Cifar10CNN, "cuda:0", dtf.MomentumOptimizer(0.05, 0.9), seed=7)
    graph = _trainer(models.Cifar10CNN, "cuda:0", dtf.MomentumOptimizer(0.05, 0.9), seed=7, use_cuda_graph=True)
    eager.step((x, y))
    graph.step((x, y))
    torch.cuda.synchronize()
    assert graph.store.global_step == 1 == eager.store.global_step
    dm = float((eager.store.master_arena - graph.store.master_arena).abs().max())
    ref = float((eager.store.master_arena).abs().max())
    assert dm < 2e-3 * ref + 1e-4, dm  # three applications of the batch would differ by ~2 lr g
    dv = float((eager.store.slot_m - graph.store.slot_m).abs().max())
    assert dv < 0.05 * float(eager.store.slot_m.abs().max()) + 1e-5, dv


def test_prefetched_steps_equal_plain_steps():
    import distributed_tensorflow_framework_b200 as dtf
    from distributed_tensorflow_framework_b200 import models
    batches = [_batch(seed=10 + i) for i in range(4)]
    a = _trainer(models.Cifar10CNN, "cuda:0", dtf.GradientDescentOptimizer(0.05), seed=1, use_cuda_graph=True)
    b = _trainer(models.Cifar10CNN, "cuda:0", dtf.GradientDescentOptimizer(0.05), seed=1, use_cuda_graph=True)
    la, lb = [], []
    for i in range(8):
        la.append(float(a.step(batches[i % 4])))
    b.prefetch(batches[0])
    for i in range(8):
        loss = b.step(batches[i % 4])
        b.prefetch(batches[(i + 1) % 4])  # H2D of the next batch overlaps this step
        lb.append(float(loss))
    assert max(abs(p - q) for p, q in zip(la, lb)) < 2e-2, (la, lb)
    assert b._copy_stream is not None


def _oracle_step(model_cls, batch, size, classes=16, damp=None):
    """One forward/backward with IDENTICAL bf16-representable weights on the CUDA kernels and on the CPU oracle (fp32
    arithmetic, activations / data gradients rounded where the kernels store bf16). Returns the two losses and
    [(relative L2 error, cosine, name)] per trainable variable."""
    import distributed_tensorflow_framework_b200 as dtf
    from distributed_tensorflow_framework_b200.engine import layers
    x = torch.rand(batch, size, size, 3).to(torch.bfloat16).float()
    y = torch.randint(0, classes, (batch,))
    cpu = _trainer(model_cls, "cpu", dtf.GradientDescentOptimizer(0.0))
    gpu = _trainer(model_cls, "cuda:0", dtf.GradientDescentOptimizer(0.0))
    cpu.build(x, y)
    gpu.build(x, y)
    if damp is not None:  # scale of the last batch norm of every residual branch (see the full-depth test)
        with torch.no_grad():
            names = set(cpu.store.vars)
            for v in cpu.store.trainable_variables():
                last = "conv3/bn/gamma" if v.name.replace("conv2/bn/gamma", "conv3/bn/gamma") in names else "conv2/bn/gamma"
                if v.name.startswith("stage") and v.name.endswith(last):
                    v.master.fill_(damp)
    gpu.store.load_state_dict(cpu.store.state_dict())
    gpu.store.shadow_arena.copy_(gpu.store.master_arena)
    gpu.store.refresh_transposed_shadows()
    with torch.no_grad():
        for v in cpu.store.trainable_variables():
            if v.kind in ("conv_w", "fc_w"):
                v.master.copy_(v.master.to(torch.bfloat16).float())
    old = layers.emulate_bf16(True)
    try:
        lc, tower_grads = cpu._forward_backward([(x, y)])
    finally:
        layers.emulate_bf16(old)
    cg = {v.name: g for g, v in tower_grads[0]}
    gpu.engine.begin_step(1)
    lg, _ = gpu._forward_backward([(x.cuda(), y.cuda())])
    torch.cuda.synchronize()
    out = []
    for v in gpu.store.trainable_variables():
        gc = cg[v.name].float()
        gg = v.logical(v.grad.detach()).float().cpu()
        if float(gc.norm()) < 1e-6:
            continue
        denom = float(gc.norm())
        if v.name.endswith("/beta") and v.name[:-4] + "gamma" in cg:
            # the gradient of a batch-norm shift that feeds another batch norm is (nearly) zero by construction -- a
            # sum of thousands of cancelling terms; measure its error on the scale of the layer's gamma gradient
            denom = max(denom, float(cg[v.name[:-4] + "gamma"].float().norm()))
        rel = float((gg - gc).norm() / denom)
        if denom > float(gc.norm()) * 4:
            out.append((rel, 1.0, v.name))
            continue
        cos = float(torch.dot(gg.flatten(), gc.flatten()) / (gg.norm() * gc.norm() + 1e-20))
        out.append((rel, cos, v.name))
    out.sort(reverse=True)
    return float(lc), float(lg), out


@pytest.mark.parametrize("bottleneck", [False, True])
@pytest.mark.parametrize("relu", [False, True])
def test_residual_stack_gradients_match_cpu_oracle_per_variable(bottleneck, relu):
    """A SHALLOW residual network that exercises every composite path of the ResNets -- 7x7 stride-2 stem (tap
    packing), max pool, identity and projection shortcuts with the forked input gradient, the stride-2 3x3
    (space-to-depth) and strided 1x1 layers, batch norm with residual, global pooling, the FC head -- against the CPU
    oracle, per variable.
    relu=False (activations switched off, models.resnet.set_relu): no ReLU decision can differ between the two
    implementations, so the comparison is tight and a wrong scale or a dropped term in a convolution / batch-norm /
    fork backward stands out. relu=True: two correct bf16 implementations already disagree on ~0.1 % of the ReLU
    decisions per layer (each costs sqrt(fraction) in relative gradient error), so the bound is the measured noise
    floor of this depth; the masked batch-norm backward itself is checked per op in test_gpu_ops.py."""
    from distributed_tensorflow_framework_b200.models import resnet
    torch.manual_seed(5)

    class Shallow(resnet._ResNet):
        num_classes = 16
        blocks = (1, 2)

    Shallow.bottleneck = bottleneck
    old = resnet.set_relu(relu)
    try:
        lc, lg, errs = _oracle_step(Shallow, 16, 64)
    finally:
        resnet.set_relu(old)
    print("shallow bottleneck=%s relu=%s: loss %.5f / %.5f, worst %s" % (
        bottleneck, relu, lc, lg, [(round(r, 4), n) for r, _, n in errs[:4]]))
    assert abs(lc - lg) < 5e-3 * max(1.0, abs(lc)), (lc, lg)
    bound = 0.2 if relu else 0.04
    assert errs[0][0] < bound, errs[:6]
    assert sorted(r for r, _, _ in errs)[len(errs) // 2] < bound * 0.75


@pytest.mark.parametrize("name", ["ResNet18", "ResNet50"])
def test_full_depth_resnet_gradients_against_cpu_oracle(name):
    """Full-depth nets. A randomly initialised batch-norm ResNet amplifies any perturbation by ~1.2x per layer (the
    known gradient explosion of BN networks at initialisation): measured here, two correct bf16 implementations that
    agree to 3e-5 after the stem disagree by 35 % in the last stage of ResNet-50, and their gradients are unrelated
    (cosine 0.3). The comparison is therefore made at the usual "zero-init residual" operating point: the last batch
    norm of every residual branch gets gamma = 0.2, which makes the blocks near-identity and the network
    well-conditioned, while every kernel still runs on full-size non-trivial data."""
    from distributed_tensorflow_framework_b200 import models
    torch.manual_seed(4)
    cls = getattr(models, name)

    class Small(cls):
        num_classes = 16

    lc, lg, errs = _oracle_step(Small, 8, 64, damp=0.2)
    print("%s: loss %.5f / %.5f, worst %s" % (name, lc, lg, [(round(r, 4), round(c, 4), n) for r, c, n in errs[:4]]))
    assert abs(lc - lg) < 1e-2 * max(1.0, abs(lc)), (lc, lg)
    for rel, cos, vname in errs:
        assert cos > 0.9 and rel < 0.45, (vname, rel, cos)
    assert sorted(r for r, _, _ in errs)[len(errs) // 2] < 0.35
