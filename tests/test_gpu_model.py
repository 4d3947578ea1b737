"""End-to-end model tests on the CUDA path: CPU-oracle parity of one training step, convergence, CUDA-graph replay."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _make(model_cls, device, opt=None, **kw):
    import distributed_tensorflow_framework_b200 as dtf
    loss = kw.pop("loss", None) or dtf.SoftmaxCrossEntropyLoss()
    return dtf.Trainer(dtf.Net(model_cls()), loss, opt or dtf.GradientDescentOptimizer(0.05), device=device, **kw)


def test_cifar_step_matches_cpu_oracle():
    from distributed_tensorflow_framework_b200 import models
    torch.manual_seed(0)
    x = torch.rand(16, 32, 32, 3)
    y = torch.randint(0, 10, (16,))
    cpu = _make(models.Cifar10CNN, "cpu")
    gpu = _make(models.Cifar10CNN, "cuda:0")
    cpu.build(x, y)
    gpu.build(x, y)
    gpu.store.load_state_dict(cpu.store.state_dict())
    gpu.store.shadow_arena.copy_(gpu.store.master_arena)
    gpu.store.refresh_transposed_shadows()
    before = {k: v.clone() for k, v in cpu.store.state_dict().items()}
    lc = float(cpu.step((x, y)))
    lg = float(gpu.step((x, y)))
    assert abs(lc - lg) < 3e-2 * max(1.0, abs(lc)), (lc, lg)
    after_c, after_g = cpu.store.state_dict(), gpu.store.state_dict()
    for k in before:
        dc = (after_c[k] - before[k]).flatten()
        dg = (after_g[k] - before[k]).flatten()
        if dc.norm() < 1e-7:
            continue
        cos = torch.dot(dc, dg) / (dc.norm() * dg.norm() + 1e-12)
        assert cos > 0.98, (k, float(cos))
        assert abs(float(dg.norm() / dc.norm()) - 1.0) < 0.1, k


def test_cifar_converges_and_graph_matches_eager():
    import distributed_tensorflow_framework_b200 as dtf
    from distributed_tensorflow_framework_b200 import models
    torch.manual_seed(1)
    x = torch.rand(64, 32, 32, 3).pin_memory()
    y = torch.randint(0, 10, (64,)).pin_memory()
    eager = _make(models.Cifar10CNN, "cuda:0", dtf.MomentumOptimizer(0.01, 0.9), seed=3)
    graph = _make(models.Cifar10CNN, "cuda:0", dtf.MomentumOptimizer(0.01, 0.9), seed=3, use_cuda_graph=True)
    le = [float(eager.step((x, y))) for _ in range(30)]
    lg = [float(graph.step((x, y))) for _ in range(30)]
    assert le[-1] < 0.9 * le[0], le
    # the two warm-up steps of the capture are rolled back: replay i is training step i, like the eager run
    for i in (0, 5, 12, 20, 29):
        assert abs(lg[i] - le[i]) < 0.03, (i, lg[i], le[i])
    assert graph.store.global_step == 30
    assert graph._graph is not None


@pytest.mark.parametrize("name", ["ResNet18", "ResNet50"])
def test_resnet_trains(name):
    import distributed_tensorflow_framework_b200 as dtf
    from distributed_tensorflow_framework_b200 import models
    torch.manual_seed(2)
    cls = getattr(models, name)

    class Small(cls):
        num_classes = 16

    x = torch.rand(8, 64, 64, 3)
    y = torch.randint(0, 16, (8,))
    tr = _make(Small, "cuda:0", dtf.MomentumOptimizer(0.05, 0.9))
    losses = [float(tr.step((x, y))) for _ in range(25)]
    assert all(l == l for l in losses)
    assert min(losses[-5:]) < 0.7 * losses[0], losses
    # inference mode (running statistics; the convolutions that feed statistics to their batch norm in training emit
    # none here) against the CPU path with the same variables
    from distributed_tensorflow_framework_b200.engine import layers, variables as V
    cpu = _make(Small, "cpu")
    cpu.build(x, y)
    cpu.store.load_state_dict(tr.store.state_dict())
    layers.set_training(False)
    try:
        with torch.no_grad():
            with V.using_store(tr.store):
                lg = tr.net.process(x.cuda()).float().cpu()
            with V.using_store(cpu.store):
                lc = cpu.net.process(x).float()
    finally:
        layers.set_training(True)
    assert lg.shape == lc.shape == (8, 16) and bool(torch.isfinite(lg).all())
    assert (lg - lc).abs().max() < 0.08 * lc.abs().max() + 0.05, ((lg - lc).abs().max(), lc.abs().max())


def test_dehaze_fcn_trains_and_matches_cpu_forward():
    import distributed_tensorflow_framework_b200 as dtf
    from distributed_tensorflow_framework_b200 import models
    torch.manual_seed(3)
    x = torch.rand(2, 32, 32, 3)
    t = torch.rand(2, 32, 32, 3)
    cpu = _make(models.DehazeFCN, "cpu", loss=dtf.MSELoss())
    gpu = _make(models.DehazeFCN, "cuda:0", dtf.AdamOptimizer(1e-3), loss=dtf.MSELoss())
    cpu.build(x, t)
    gpu.build(x, t)
    gpu.store.load_state_dict(cpu.store.state_dict())
    gpu.store.shadow_arena.copy_(gpu.store.master_arena)
    gpu.store.refresh_transposed_shadows()
    yc = cpu.forward(x)
    yg = gpu.forward(x).float().cpu()
    assert yg.shape == yc.shape == (2, 32, 32, 3)
    assert (yc - yg).abs().max() < 0.05 * (yc.abs().max() + 1e-3) + 2e-2
    losses = [float(gpu.step((x, t))) for _ in range(30)]
    assert losses[-1] < losses[0], losses
