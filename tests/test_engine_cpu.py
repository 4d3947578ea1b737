"""CPU engine (the oracle path): layer semantics vs manual TF-style math, tower gradient averaging, optimizers."""
import hypothesis.strategies as st
import pytest
import torch
import torch.nn.functional as F
from hypothesis import given, settings

import distributed_tensorflow_framework_b200 as dtf
from distributed_tensorflow_framework_b200.engine import layers as L
from distributed_tensorflow_framework_b200.engine import variables as V
from distributed_tensorflow_framework_b200.ops.raw import same_pad


@pytest.fixture()
def store():
    s = V.VariableStore("cpu", seed=1)
    with V.using_store(s):
        yield s


@given(st.integers(1, 40), st.integers(1, 7), st.integers(1, 3))
@settings(max_examples=60, deadline=None)
def test_same_padding_rule(n, k, s):
    out, before = same_pad(n, k, s)
    assert out == -(-n // s)
    total = max((out - 1) * s + k - n, 0)
    assert before == total // 2 and total - before >= before  # extra padding goes to the bottom / right


def test_conv_same_semantics_and_variable_names(store):
    x = torch.randn(2, 7, 9, 3)
    y = L.conv("conv1", x, 3, 4, kernel_size=[3, 3], stride=[1, 2, 2, 1])
    assert y.shape == (2, 4, 5, 4) and (y >= 0).all()
    assert set(store.vars) == {"conv1/weights", "conv1/biases"}
    w = store.vars["conv1/weights"].master
    assert w.shape == (3, 3, 3, 4) and w.abs().max() <= 0.1 + 1e-6  # trunc-normal(0.05) clipped at 2 sigma
    # manual reference: pad (top 1,bottom 1) x (left 1,right 1) for 7x9 / k3 / s2
    xp = F.pad(x.permute(0, 3, 1, 2), (1, 1, 1, 1))
    ref = F.relu(F.conv2d(xp, w.permute(3, 2, 0, 1), stride=2)).permute(0, 2, 3, 1)
    assert torch.allclose(y, ref, atol=1e-5)
    y2 = L.conv_eval("conv1", x, 3, 4, kernel_size=[3, 3], stride=[1, 2, 2, 1])  # *_eval reuses by name
    assert torch.equal(y, y2) and len(store.vars) == 2
    L.conv_nonacti("c2", x, 3, 4)
    assert {"c2/weights_nonacti", "c2/biases_nonacti"} <= set(store.vars)
    with pytest.raises(ValueError):
        L.conv("conv1", x, 3, 8)  # same name, different shape


def test_even_kernel_asymmetric_padding(store):
    x = torch.randn(1, 4, 4, 1)
    y = L.conv_nonacti("c", x, 1, 1, kernel_size=[2, 2])
    w = store.vars["c/weights_nonacti"].master
    ref = F.conv2d(F.pad(x.permute(0, 3, 1, 2), (0, 1, 0, 1)), w.permute(3, 2, 0, 1)).permute(0, 2, 3, 1)
    assert torch.allclose(y, ref, atol=1e-6)


@pytest.mark.parametrize("stride,k,hw", [(1, 3, (6, 6)), (2, 3, (8, 8)), (2, 3, (7, 9)), (2, 1, (8, 8)), (2, 5, (10, 6))])
def test_deconv_is_the_adjoint_of_same_conv(store, stride, k, hw):
    cin_t, cout_t = 3, 5  # transposed conv: x has 5 channels, output has 3
    oh, _ = same_pad(hw[0], k, stride)
    ow, _ = same_pad(hw[1], k, stride)
    x = torch.randn(2, oh, ow, cout_t)
    y = L.deconv("up", x, cin_t, cout_t, output_shape=[2, hw[0], hw[1], cin_t], kernel_size=[k, k],
                 stride=[1, stride, stride, 1])
    assert y.shape == (2, hw[0], hw[1], cin_t)
    w = store.vars["up/weights"].master  # [k, k, cin_t, cout_t] = HWIO of the forward conv (cin_t -> cout_t)
    probe = torch.randn(2, hw[0], hw[1], cin_t, requires_grad=True)
    pt, pl = same_pad(hw[0], k, stride)[1], same_pad(hw[1], k, stride)[1]
    pb = max((oh - 1) * stride + k - hw[0], 0) - pt
    pr = max((ow - 1) * stride + k - hw[1], 0) - pl
    fwd = F.conv2d(F.pad(probe.permute(0, 3, 1, 2), (pl, pr, pt, pb)), w.permute(3, 2, 0, 1), stride=stride)
    (fwd * x.permute(0, 3, 1, 2)).sum().backward()
    assert torch.allclose(y, probe.grad, atol=1e-4)


def test_pool_avg_excludes_padding_and_max(store):
    x = torch.ones(1, 3, 3, 1)
    y = L.pool("p", x, kernel=[1, 2, 2, 1], stride=[1, 2, 2, 1], is_max_pool=False)
    assert y.shape == (1, 2, 2, 1) and torch.allclose(y, torch.ones_like(y))  # TF SAME avg divides by valid count
    x = torch.arange(16.).view(1, 4, 4, 1)
    assert L.pool("p", x).flatten().tolist() == [5., 7., 13., 15.]


def test_literal_batch_norm_and_lrn_and_fc(store):
    x = torch.randn(16, 4, 4, 3)
    y = L.batch_norm(x)
    ref = (x - x.mean(0, keepdim=True)) / torch.sqrt(x.var(0, unbiased=False, keepdim=True) + 1e-3)
    assert torch.allclose(y, ref, atol=1e-5) and not store.vars  # no scale / offset variables (distribute_tools.py:168-180)
    z = torch.randn(2, 3, 3, 16)
    out = L.norm("n", z, lsize=4)
    sq = F.pad(z ** 2, (4, 4))
    s = sum(sq[..., i:i + 16] for i in range(9))
    assert torch.allclose(out, z / (1.0 + (0.001 / 9.0) * s) ** 0.75, atol=1e-5)
    dtf.FLAGS.lrn_compat_noop = True
    try:
        assert L.norm("n", z) is z  # the reference's literal behaviour (result discarded)
    finally:
        dtf.FLAGS.lrn_compat_noop = False
    f = L.FC_layer("fc", x, 7)
    assert f.shape == (16, 7) and (f >= 0).all() and store.vars["fc/weights"].shape == (48, 7)
    lim = (6.0 / (48 + 7)) ** 0.5
    assert store.vars["fc/weights"].master.abs().max() <= lim  # Xavier uniform


def test_batch_norm_channel_train_and_eval(store):
    x = torch.randn(8, 5, 5, 6) * 3 + 2
    y = L.batch_norm_channel("bn", x)
    flat = y.reshape(-1, 6)
    assert flat.mean(0).abs().max() < 1e-4 and (flat.var(0, unbiased=False) - 1).abs().max() < 1e-2
    assert store.vars["bn/moving_mean"].master.abs().max() > 0 and not store.vars["bn/moving_mean"].trainable
    L.set_training(False)
    try:
        y_eval = L.batch_norm_channel("bn", x)
    finally:
        L.set_training(True)
    assert not torch.allclose(y, y_eval)


@given(st.integers(1, 4), st.integers(1, 3))
@settings(max_examples=20, deadline=None)
def test_average_gradients_is_per_variable_mean(n_towers, n_vars):
    torch.manual_seed(n_towers * 10 + n_vars)
    vars_ = ["v%d" % i for i in range(n_vars)]
    tower_grads = [[(torch.randn(3, 2), v) for v in vars_] for _ in range(n_towers)]
    avg = dtf.Tower.average_gradients(tower_grads)
    assert [v for _, v in avg] == vars_
    for i, (g, _) in enumerate(avg):
        ref = torch.stack([tower_grads[t][i][0] for t in range(n_towers)]).mean(0)
        assert torch.allclose(g, ref, atol=1e-6)


@pytest.mark.parametrize("opt_name", ["sgd", "momentum", "adam"])
def test_cpu_optimizers_match_torch(store, opt_name):
    v = store.get_variable("w", [4, 3], V.XavierInitializer())
    ref = v.master.detach().clone().requires_grad_(True)
    ours = {"sgd": dtf.GradientDescentOptimizer(0.1), "momentum": dtf.MomentumOptimizer(0.1, 0.9),
            "adam": dtf.AdamOptimizer(0.01)}[opt_name]
    theirs = {"sgd": torch.optim.SGD([ref], 0.1), "momentum": torch.optim.SGD([ref], 0.1, momentum=0.9),
              "adam": torch.optim.Adam([ref], 0.01)}[opt_name]
    for _ in range(4):
        g = torch.randn(4, 3)
        ref.grad = g.clone()
        theirs.step()
        ours.apply_gradients([(g, v)], store=store)
    assert torch.allclose(v.master, ref, atol=1e-5)
    assert store.global_step == 4
    slots = ours.slot_tensors(store)
    want = {"sgd": set(), "momentum": {"w/Momentum"}, "adam": {"w/Adam", "w/Adam_1", "beta1_power", "beta2_power"}}
    assert set(slots) == want[opt_name]  # TF slot names = checkpoint keys


def test_weight_decay_term_joins_the_losses_collection(store):
    L._variable_with_weight_decay("w", [3, 3], stddev=0.1, wd=0.5)
    (name, term), = store.losses
    assert name == "w/weight_loss"
    assert torch.allclose(term, 0.5 * 0.5 * (store.vars["w"].master ** 2).sum())


def test_bf16_emulating_oracle_rounds_activations_and_gradients():
    """layers.emulate_bf16: layer outputs become bf16-representable, the data gradient entering a layer is rounded,
    parameter gradients stay fp32, and the mode is off by default."""
    import torch
    import distributed_tensorflow_framework_b200 as dtf
    from distributed_tensorflow_framework_b200 import models
    from distributed_tensorflow_framework_b200.engine import layers
    torch.manual_seed(0)
    x = torch.rand(4, 32, 32, 3)
    y = torch.randint(0, 10, (4,))
    tr = dtf.Trainer(dtf.Net(models.Cifar10CNN()), dtf.SoftmaxCrossEntropyLoss(), dtf.GradientDescentOptimizer(0.0))
    tr.build(x, y)
    l_plain, g_plain = tr._forward_backward([(x, y)])
    assert layers.emulate_bf16(True) is False
    try:
        with V.using_store(tr.store):
            h = layers.conv("conv1", x, 3, 64, [5, 5])
        assert torch.equal(h, h.to(torch.bfloat16).float())
        l_emu, g_emu = tr._forward_backward([(x, y)])
    finally:
        assert layers.emulate_bf16(False) is True
    assert abs(float(l_plain) - float(l_emu)) < 2e-2
    for (ga, va), (gb, vb) in zip(g_plain[0], g_emu[0]):
        assert va is vb and ga.dtype == torch.float32 == gb.dtype
        if float(ga.norm()) > 1e-6:
            assert float((ga - gb).norm() / ga.norm()) < 0.25, va.name
    assert any(not torch.equal(ga, gb) for (ga, _), (gb, _) in zip(g_plain[0], g_emu[0]))
