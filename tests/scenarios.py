"""User-level scenarios shared by the CPU tier (tests/test_scenarios_cpu.py) and the GPU tier
(tests/test_gpu_scenarios.py): every scenario goes through the IoC container, like an experiment script would."""
import os

import numpy as np
import torch

import distributed_tensorflow_framework_b200 as dtf
from distributed_tensorflow_framework_b200 import models
from distributed_tensorflow_framework_b200.api import annotations as ann
from distributed_tensorflow_framework_b200.engine import layers, variables as V
from distributed_tensorflow_framework_b200.io import tensorbundle as tb
from distributed_tensorflow_framework_b200.plugins import input as Input
from distributed_tensorflow_framework_b200.plugins.learningrate import LearningRate
from distributed_tensorflow_framework_b200.runtime import hooks as H


class TinyResNet18(models.ResNet18):
    num_classes = 16


class TinyResNet50(models.ResNet50):
    num_classes = 16


class TinyImageLoader(models.ImageNetSyntheticLoader):
    image_shape = (64, 64, 3)
    num_classes = 16
    pool_batches = 2
    eval_batches = 2


class TinyMLP(dtf.Model):
    def inference(self, x, *a, **k):
        x = dtf.tools.FC_layer("fc1", x, 64)
        return dtf.tools.fc_nonacti("logits", x, 4)


def _blob(i):
    """Sample i of a 4-class toy problem: class = i % 4, features = one-hot-ish blob plus deterministic noise."""
    c = i % 4
    v = np.full((16,), 0.1, np.float32)
    v[4 * c:4 * c + 4] = 1.0
    v += 0.05 * np.sin(np.arange(16, dtype=np.float32) * (i + 1))
    return v, np.int64(c)


class BlobPlaceholderLoader(Input.PlaceholderDataLoader):
    """Placeholder kind: a queue of sample descriptors, decoded per batch (distribute_input.py:118-160)."""

    def _put_names_dict_into_queue(self, q, *a, **k):
        for i in range(64):
            q.put({"index": i})

    def decode_data_from_path_name(self, paths, *a, **k):
        v, c = _blob(paths["index"])
        return {"raw_data": v, "ground_truth": c}

    def load_eval_batch(self, *a, **k):
        n = getattr(self, "_served", 0)
        if n >= 2:
            self._served = 0
            return None
        self._served = n + 1
        xs, ys = zip(*[_blob(1000 + n * self.batch_size + j) for j in range(self.batch_size)])
        return torch.from_numpy(np.stack(xs)), torch.from_numpy(np.stack(ys))


class BlobPathLoader(Input.DataPathDataLoader):
    """Data-path kind: a name queue, a reader and a parser (distribute_input.py:162-209)."""

    def _create_reader(self):
        return lambda name: int(name.rsplit("_", 1)[1])

    def _parse_raw_data(self, value):
        v, c = _blob(value)
        return [v, c]

    def create_name_queue(self, data_dir):
        return Input.FilenameQueue(["%s/sample_%d" % (data_dir, i) for i in range(64)])

    load_eval_batch = BlobPlaceholderLoader.load_eval_batch


def experiment(model_dir, gpu, mode="Train", model="TinyResNet18", loader="TinyImageLoader", epochs=1, samples=64,
               batch=8, opt=None, lr_handler=None, port=29741):
    @ann.current_model(model=model)
    @ann.optimizer(optimizer=opt or dtf.MomentumOptimizer(0.02, 0.9))
    @ann.loss(loss="SoftmaxCrossEntropyLoss")
    @ann.current_mode(mode=mode)
    @ann.current_input(input=loader)
    @ann.gpu_num(gpu_num=1 if gpu else 0)
    @ann.ps_hosts(ps_hosts="")
    @ann.worker_hosts(worker_hosts="127.0.0.1:%d" % port)
    @ann.job_name(job_name="worker")
    @ann.task_index(task_index=0)
    @ann.batch_size(batch_size=batch)
    @ann.sample_number(sample_number=samples)
    @ann.epoch_num(epoch_num=epochs)
    @ann.model_dir(model_dir=model_dir)
    @ann.data_dir(data_dir="blobs")
    def main():
        pass

    if lr_handler is not None:
        main = ann.learning_rate(learning_rate=lr_handler)(main)
    return main


def resnet_train_eval_and_cross_device_restore(tmp_path, gpu, model="TinyResNet18"):
    """Train a batch-norm ResNet through the Train operator, evaluate it with the Eval operator (running statistics
    restored from the checkpoint), then load the checkpoint into a CPU store and compare inference logits."""
    md = str(tmp_path / "m")
    os.makedirs(md)
    op = dtf.run_main(experiment(md, gpu, model=model, epochs=2))
    assert op.trainer.device.type == ("cuda" if gpu else "cpu")
    assert op.trainer.store.global_step == 16 and op.final_loss == op.final_loss
    prefix = tb.latest_checkpoint(md)
    with tb.BundleReader(prefix) as r:
        keys = set(r.keys())
        assert "conv1/bn/moving_mean" in keys and "conv1/bn/moving_variance" in keys and "conv1/bn/gamma" in keys
        mm = torch.from_numpy(r.get_tensor("conv1/bn/moving_mean"))
        assert float(mm.abs().max()) > 0  # running statistics were updated by training and saved
    ev = dtf.run_main(experiment(md, gpu, mode="Eval", model=model))
    res = ev.results
    assert res["global_step"] == 16 and res["batches"] == 2 and res["loss"] == res["loss"]
    assert 0.0 <= res["accuracy"] <= 1.0
    # the same checkpoint in a fresh CPU store gives the logits of the device that trained it
    loader = TinyImageLoader.__new__(TinyImageLoader)
    loader.batch_size = 8
    x, _ = loader.next_host_batch()
    x = x.clone()
    cpu = dtf.Trainer(dtf.Net(globals()[model]()), dtf.SoftmaxCrossEntropyLoss(), dtf.GradientDescentOptimizer(0.1),
                      device="cpu")
    cpu.build(x, torch.zeros(8, dtype=torch.long))
    dtf.Saver(cpu.store).restore(prefix)
    layers.set_training(False)
    try:
        with torch.no_grad():
            with V.using_store(cpu.store):
                lc = cpu.net.process(x).float()
            with V.using_store(op.trainer.store):
                lt = op.trainer.net.process(x.to(op.trainer.device)).float().cpu()
    finally:
        layers.set_training(True)
    assert (lc - lt).abs().max() < 0.08 * lc.abs().max() + 0.05, (float((lc - lt).abs().max()), float(lc.abs().max()))
    return res


def loader_kinds_learn_the_toy_problem(tmp_path, gpu):
    """Placeholder and data-path loaders through Train + Eval: the 4-class blobs are separable, so accuracy rises."""
    out = {}
    for kind, loader in (("placeholder", "BlobPlaceholderLoader"), ("datapath", "BlobPathLoader")):
        md = str(tmp_path / kind)
        os.makedirs(md)
        dtf.run_main(experiment(md, gpu, model="TinyMLP", loader=loader, epochs=6, samples=64, batch=16,
                                opt=dtf.AdamOptimizer(0.01)))
        ev = dtf.run_main(experiment(md, gpu, mode="Eval", model="TinyMLP", loader=loader, batch=16))
        assert ev.results["global_step"] == 24 and ev.results["accuracy"] >= 0.9, (kind, ev.results)
        out[kind] = ev.results
    return out


def adam_with_decaying_learning_rate(tmp_path, gpu):
    """Adam + a LearningRate handler that decays during the run: the optimizer follows it (under CUDA-graph replay on
    the GPU), the value is persisted with the checkpoint, and a second run resumes with the decayed value."""
    md = str(tmp_path / "m")
    os.makedirs(md)
    handler = LearningRate(0.01, os.path.join(md, "lr.json"), decay_factor=0.1)
    seen = []

    class Decay(H.SessionRunHook):
        def after_run(self, ctx, values):
            seen.append((ctx.session.global_step, op.optimizer.learning_rate))
            if ctx.session.global_step == 4:
                handler.decay()

    op = dtf.build_operator(experiment(md, gpu, model="TinyMLP", loader="BlobPlaceholderLoader", epochs=2, samples=64,
                                       batch=16, opt=dtf.AdamOptimizer(123.0), lr_handler=handler))
    op.extra_hooks = [Decay()]
    op.run()
    assert [round(lr, 6) for _, lr in seen] == [0.01] * 4 + [0.001] * 4, seen
    assert op.final_loss == op.final_loss
    again = LearningRate(0.01, os.path.join(md, "lr.json"), decay_factor=0.1)
    assert abs(again.learning_rate - 0.001) < 1e-9
    return seen


# ---------------------------------------------------------------------------------------------------------------------
# The whole layer library on awkward shapes: rectangular images, channel counts that are not multiples of 8, odd
# spatial sizes, 1/3/5/7 kernels, strides 1 and 2, literal batch norm, LRN, transposed convolutions.

class KitchenSinkClassifier(dtf.Model):
    def inference(self, x, *a, **k):
        T = dtf.tools
        x = T.conv("c1", x, 3, 20, kernel_size=[5, 5])                                  # 24x40x20
        x = T.pool("p1", x, kernel=[1, 3, 3, 1], stride=[1, 2, 2, 1])                   # 12x20
        x = T.norm("n1", x)
        x = T.conv_nonacti("c2", x, 20, 36, kernel_size=[3, 3], stride=[1, 2, 2, 1])    # 6x10x36
        x = T.acti_layer(x)
        x = T.batch_norm(x)
        x = T.conv_nonacti("c3", x, 36, 28, kernel_size=[1, 1])
        x = layers.batch_norm_channel("bn3", x, relu=True)                              # per-channel, 28 channels
        x = T.conv("c4", x, 28, 24, kernel_size=[7, 7])
        x = T.pool("p2", x, is_max_pool=False)                                          # 3x5x24
        x = T.FC_layer("f1", x, 50)
        return T.fc_nonacti("logits", x, 7)


class KitchenSinkImage(dtf.Model):
    def inference(self, x, *a, **k):
        T = dtf.tools
        n, h, w = x.shape[0], x.shape[1], x.shape[2]
        e = T.conv("c1", x, 3, 12, kernel_size=[3, 3])                                  # 21x27x12
        d = T.conv("c2", e, 12, 40, kernel_size=[3, 3], stride=[1, 2, 2, 1])            # 11x14x40
        d = T.conv("c3", d, 40, 44, kernel_size=[3, 3])
        u = T.deconv("u1", d, 12, 44, output_shape=[n, h, w, 12], stride=[1, 2, 2, 1])  # back to 21x27x12
        u = T.acti_layer(u)
        u = T.add_relu(u, e, relu=False)
        return T.conv_nonacti("out", u, 12, 3, kernel_size=[1, 1])


def _one_step_parity(model_cls, loss, x, y, cos_min=0.97, ratio_tol=0.15):
    """One SGD step on the CPU oracle and on CUDA from identical variables: loss, forward output and the update of every
    variable (direction and size) must agree within bf16 noise."""
    cpu = dtf.Trainer(dtf.Net(model_cls()), loss, dtf.GradientDescentOptimizer(0.05), device="cpu")
    gpu = dtf.Trainer(dtf.Net(model_cls()), loss, dtf.GradientDescentOptimizer(0.05), device="cuda:0")
    cpu.build(x, y)
    gpu.build(x, y)
    gpu.store.load_state_dict(cpu.store.state_dict())
    gpu.store.shadow_arena.copy_(gpu.store.master_arena)
    gpu.store.refresh_transposed_shadows()
    fc, fg = cpu.forward(x).float(), gpu.forward(x).float().cpu()
    assert fc.shape == fg.shape
    assert (fc - fg).abs().max() < 0.04 * fc.abs().max() + 0.02, (float((fc - fg).abs().max()), float(fc.abs().max()))
    before = {k: v.clone() for k, v in cpu.store.state_dict().items()}
    lc, lg = float(cpu.step((x, y))), float(gpu.step((x, y)))
    assert abs(lc - lg) < 0.03 * max(1.0, abs(lc)), (lc, lg)
    after_c, after_g = cpu.store.state_dict(), gpu.store.state_dict()
    checked = 0
    # a bias in front of a batch norm has a zero gradient up to rounding: updates far below the largest one are only
    # required to be small on both devices
    tiny = 1e-3 * max(float((after_c[k] - before[k]).norm()) for k in before)
    for k in before:
        dc = (after_c[k] - before[k]).flatten().double()
        dg = (after_g[k].cpu() - before[k]).flatten().double()
        if dc.norm() < tiny:
            assert dg.norm() < 20 * tiny + 1e-6, (k, float(dc.norm()), float(dg.norm()))
            continue
        cos = float(torch.dot(dc, dg) / (dc.norm() * dg.norm() + 1e-30))
        ratio = float(dg.norm() / dc.norm())
        assert cos > cos_min and abs(ratio - 1.0) < ratio_tol, (k, cos, ratio)
        checked += 1
    return checked


def kitchen_sink_parity():
    torch.manual_seed(11)
    # (the first layers sit behind two batch norms over a batch of 6, an LRN and a max pool: bf16 noise is amplified
    # there, so this integration check is looser than the per-layer one below)
    n = _one_step_parity(KitchenSinkClassifier, dtf.SoftmaxCrossEntropyLoss(), torch.rand(6, 24, 40, 3),
                         torch.randint(0, 7, (6,)), cos_min=0.95, ratio_tol=0.3)
    assert n >= 14
    n = _one_step_parity(KitchenSinkImage, dtf.MSELoss(), torch.rand(3, 21, 27, 3), torch.rand(3, 21, 27, 3))
    assert n >= 8


# Every layer on its own behind a 1x1 convolution that produces C channels: forward, loss and the update of that
# convolution (i.e. the layer's backward) against the CPU path.
_T = dtf.tools
LAYER_CASES = {
    "pool_max_3x3s2": lambda x: _T.pool("p", x, kernel=[1, 3, 3, 1], stride=[1, 2, 2, 1]),
    "pool_max_2x2": lambda x: _T.pool("p", x),
    "pool_avg_2x2": lambda x: _T.pool("p", x, is_max_pool=False),
    "lrn": lambda x: _T.norm("n", x),
    "batch_norm_literal": lambda x: _T.batch_norm(x),
    "acti": lambda x: _T.acti_layer(x - 0.5),
    "add_relu": lambda x: _T.add_relu(x - 0.5, 0.5 * x, relu=True),
    "gap": lambda x: _T.global_avg_pool(x),
    "bn_channel": lambda x: layers.batch_norm_channel("bn", x, relu=True),
    "bn_channel_res": lambda x: layers.batch_norm_channel("bn", x, relu=True, residual=0.5 - x),
    "conv3x3": lambda x: _T.conv("c", x, x.shape[-1], 12),
    "conv5x5s2_nonacti": lambda x: _T.conv_nonacti("c", x, x.shape[-1], 20, kernel_size=[5, 5], stride=[1, 2, 2, 1]),
    "conv1x1": lambda x: _T.conv("c", x, x.shape[-1], 36, kernel_size=[1, 1]),
    "deconv_s2": lambda x: _T.deconv("d", x, 12, x.shape[-1], stride=[1, 2, 2, 1],
                                     output_shape=[x.shape[0], 2 * x.shape[1], 2 * x.shape[2], 12]),
    "fc": lambda x: _T.FC_layer("f", x, 30),
    "fc_nonacti": lambda x: _T.fc_nonacti("f", x, 7),
}


def layer_case_parity(name, n, h, w, C):
    fn = LAYER_CASES[name]

    class M(dtf.Model):
        def inference(self, x, *a, **k):
            return fn(_T.conv_nonacti("pre", x, 3, C, kernel_size=[1, 1]))

    torch.manual_seed(0)
    x = torch.rand(n, h, w, 3)
    probe = dtf.Trainer(dtf.Net(M()), dtf.MSELoss(), dtf.GradientDescentOptimizer(0.05), device="cpu")
    probe.build(x, None)
    shape = probe.forward(x).shape
    return _one_step_parity(M, dtf.MSELoss(), x, torch.rand(*shape)), tuple(shape)


def in_process_towers_equal_one_big_batch(device, use_graph=False):
    """`gpu_num = 2` towers inside one process (distribute_train.py:116-139 with fewer devices than towers): the mean
    of the tower gradients is the gradient of the concatenated batch for a net without batch statistics, so two towers
    of 8 images must move the variables like one tower of 16."""
    torch.manual_seed(5)
    xs = [torch.rand(8, 32, 32, 3) for _ in range(2)]
    ys = [torch.randint(0, 10, (8,)) for _ in range(2)]
    kw = {"use_cuda_graph": True} if use_graph else {}
    one = dtf.Trainer(dtf.Net(models.Cifar10CNN()), dtf.SoftmaxCrossEntropyLoss(), dtf.MomentumOptimizer(0.05, 0.9),
                      device=device, **kw)
    two = dtf.Trainer(dtf.Net(models.Cifar10CNN()), dtf.SoftmaxCrossEntropyLoss(), dtf.MomentumOptimizer(0.05, 0.9),
                      device=device, gpu_num=2, **kw)
    one.build(xs[0], ys[0])
    two.build(xs[0], ys[0])
    two.store.load_state_dict(one.store.state_dict())
    if str(device) != "cpu":
        two.store.shadow_arena.copy_(two.store.master_arena)
        two.store.refresh_transposed_shadows()
    before = {k: v.clone().cpu() for k, v in one.store.state_dict().items()}
    for _ in range(3):  # three steps: momentum slots take part, graph replay (if any) is exercised
        l1 = float(one.step((torch.cat(xs), torch.cat(ys))))
        l2 = float(two.step([(xs[0], ys[0]), (xs[1], ys[1])]))
        assert abs(l1 - l2) < 0.02 * max(1.0, abs(l1)), (l1, l2)
    a1, a2 = one.store.state_dict(), two.store.state_dict()
    for k in before:
        d1 = (a1[k].cpu() - before[k]).flatten().double()
        d2 = (a2[k].cpu() - before[k]).flatten().double()
        if d1.norm() < 1e-9:
            continue
        cos = float(torch.dot(d1, d2) / (d1.norm() * d2.norm() + 1e-30))
        assert cos > 0.98 and abs(float(d2.norm() / d1.norm()) - 1.0) < 0.1, (k, cos, float(d2.norm() / d1.norm()))
    assert one.store.global_step == two.store.global_step == 3


# ---------------------------------------------------------------------------------------------------------------------
# User code written with plain torch ops between the library's layers, a user-defined loss and advice hooks.

class UserOpsModel(dtf.Model):
    """Two branches joined with torch.cat (a skip connection), arithmetic and a sigmoid written in torch."""

    def inference(self, x, *a, **k):
        T = dtf.tools
        s = T.conv("stem", x, 3, 16)
        a1 = T.conv("a1", s, 16, 16)
        b1 = T.pool("b_pool", s)
        b1 = T.deconv("b_up", T.conv("b1", b1, 16, 24), 8, 24, output_shape=[s.shape[0], s.shape[1], s.shape[2], 8],
                      stride=[1, 2, 2, 1])
        j = torch.cat([a1, b1.to(a1.dtype) * 0.5], dim=-1)                       # 24 channels
        y = T.conv_nonacti("head", j, 24, 3, kernel_size=[1, 1])
        return torch.sigmoid(y.float())


class L1PlusTVLoss(dtf.Loss):
    """A user loss in torch: mean absolute error plus a small total-variation term on the prediction."""

    def loss(self, predict, ground_truth):
        p = predict.float()
        l1 = (p - ground_truth.float()).abs().mean()
        tv = (p[:, 1:] - p[:, :-1]).abs().mean() + (p[:, :, 1:] - p[:, :, :-1]).abs().mean()
        return l1 + 0.1 * tv


def _centre(raw, gt, args, kwargs):
    return raw * 2.0 - 1.0, gt


def user_torch_ops_loss_and_advice_parity():
    torch.manual_seed(7)
    x, t = torch.rand(4, 16, 20, 3), torch.rand(4, 16, 20, 3)
    mk = lambda dev: dtf.Trainer(dtf.Net(UserOpsModel()), L1PlusTVLoss(), dtf.GradientDescentOptimizer(1.0),  # noqa
                                 device=dev, pre_process_fn=_centre)
    cpu, gpu = mk("cpu"), mk("cuda:0")
    cpu.build(x, t)
    gpu.build(x, t)
    gpu.store.load_state_dict(cpu.store.state_dict())
    gpu.store.shadow_arena.copy_(gpu.store.master_arena)
    gpu.store.refresh_transposed_shadows()
    before = {k: v.clone() for k, v in cpu.store.state_dict().items()}
    lc, lg = float(cpu.step((x, t))), float(gpu.step((x, t)))
    assert abs(lc - lg) < 0.02 * max(1.0, abs(lc)), (lc, lg)
    ac, ag = cpu.store.state_dict(), gpu.store.state_dict()
    tiny = 1e-3 * max(float((ac[k] - before[k]).norm()) for k in before)
    n = 0
    for k in before:
        dc = (ac[k] - before[k]).flatten().double()
        dg = (ag[k].cpu() - before[k]).flatten().double()
        if dc.norm() < tiny:
            continue
        cos = float(torch.dot(dc, dg) / (dc.norm() * dg.norm() + 1e-30))
        assert cos > 0.95 and abs(float(dg.norm() / dc.norm()) - 1.0) < 0.25, (k, cos, float(dg.norm() / dc.norm()))
        n += 1
    assert n >= 8, n
    losses = [float(gpu.step((x, t))) for _ in range(20)]
    assert losses[-1] < lg, (lg, losses)


def graph_replay_with_a_smaller_batch_in_between(opt_factory, abs_tol=1e-5):
    """CUDA-graph training where one batch has another size (the last batch of an epoch): that step runs eagerly,
    replays continue afterwards, and the result equals eager training on the same sequence."""
    torch.manual_seed(9)
    big = [(torch.rand(16, 32, 32, 3), torch.randint(0, 10, (16,))) for _ in range(3)]
    small = (torch.rand(6, 32, 32, 3), torch.randint(0, 10, (6,)))
    seq = [big[0], big[1], small, big[2], big[0], small, big[1]]
    mk = lambda g: dtf.Trainer(dtf.Net(models.Cifar10CNN()), dtf.SoftmaxCrossEntropyLoss(), opt_factory(),  # noqa
                               device="cuda:0", use_cuda_graph=g)
    eager, graph = mk(False), mk(True)
    eager.build(*big[0])
    graph.build(*big[0])
    graph.store.load_state_dict(eager.store.state_dict())
    graph.store.shadow_arena.copy_(graph.store.master_arena)
    graph.store.refresh_transposed_shadows()
    for b in seq:
        le, lg = float(eager.step(b)), float(graph.step(b))
        assert abs(le - lg) < 5e-3 * max(1.0, abs(le)), (le, lg)
    assert graph._graph is not None and eager.store.global_step == graph.store.global_step == len(seq)
    se, sg = eager.store.state_dict(), graph.store.state_dict()
    for k in se:
        d = float((se[k] - sg[k]).abs().max())
        assert d <= 2e-3 * float(se[k].abs().max()) + abs_tol, (k, d)


# ---------------------------------------------------------------------------------------------------------------------
# weight() / bias() handed to user code (distribute_tools.py:211-224): the variables are used by plain torch ops.

class RawVariableModel(dtf.Model):
    def inference(self, x, *a, **k):
        T = dtf.tools
        h = T.global_avg_pool(T.conv("c1", x, 3, 48))                      # [N, 48]
        with dtf.variable_scope("proj"):
            W, b = T.weight([48, 12]), T.bias([12])
        z = torch.tanh(h.float() @ W.value().float() + b.value().float())   # user math in torch
        with dtf.variable_scope("gate"):
            g = T.weight([12])
        z = z * torch.sigmoid(g.value().float()) + z * g.value().float()    # the same variable used twice
        return T.fc_nonacti("logits", z, 5)


def raw_variables_in_user_code_train(device):
    torch.manual_seed(13)
    x, y = torch.rand(8, 12, 12, 3), torch.randint(0, 5, (8,))
    tr = dtf.Trainer(dtf.Net(RawVariableModel()), dtf.SoftmaxCrossEntropyLoss(), dtf.GradientDescentOptimizer(0.5),
                     device=device)
    tr.build(x, y)
    return tr, x, y


def raw_variables_parity():
    cpu, x, y = raw_variables_in_user_code_train("cpu")
    gpu, _, _ = raw_variables_in_user_code_train("cuda:0")
    gpu.store.load_state_dict(cpu.store.state_dict())
    gpu.store.shadow_arena.copy_(gpu.store.master_arena)
    gpu.store.refresh_transposed_shadows()
    before = {k: v.clone() for k, v in cpu.store.state_dict().items()}
    lc, lg = float(cpu.step((x, y))), float(gpu.step((x, y)))
    assert abs(lc - lg) < 0.02 * max(1.0, abs(lc)), (lc, lg)
    ac, ag = cpu.store.state_dict(), gpu.store.state_dict()
    for k in ("proj/weights", "proj/biases", "gate/weights", "c1/weights", "logits/weights"):
        dc = (ac[k] - before[k]).flatten().double()
        dg = (ag[k].cpu() - before[k]).flatten().double()
        assert dc.norm() > 1e-6, k  # the user's variables do receive gradients
        cos = float(torch.dot(dc, dg) / (dc.norm() * dg.norm() + 1e-30))
        assert cos > 0.97 and abs(float(dg.norm() / dc.norm()) - 1.0) < 0.15, (k, cos, float(dg.norm() / dc.norm()))


def warm_start_then_resume(tmp_path, gpu):
    """init_fn warm-starts a new model_dir from another run's checkpoint (tf.train.Scaffold(init_fn=...),
    distribute_train.py:169-175); a later run on the same model_dir resumes from ITS checkpoint and does not run the
    init_fn again."""
    pre, md = str(tmp_path / "pretrained"), str(tmp_path / "finetune")
    os.makedirs(pre)
    os.makedirs(md)
    dtf.run_main(experiment(pre, gpu, model="TinyMLP", loader="BlobPlaceholderLoader", epochs=4, samples=64, batch=16,
                            opt=dtf.AdamOptimizer(0.01)))
    src = tb.latest_checkpoint(pre)
    with tb.BundleReader(src) as r:
        w_src = torch.from_numpy(r.get_tensor("fc1/weights"))
    calls = []

    def init_fn(scaffold, sess):
        calls.append(sess.global_step)
        dtf.Saver(sess.store).restore(src)   # parameters only: no optimizer passed, slots start fresh
        sess.store.global_step = 0

    seen = {}

    class FirstStep(H.SessionRunHook):
        def after_create_session(self, session):
            seen["w0"] = session.store.vars["fc1/weights"].master.detach().cpu().clone()
            seen["step0"] = session.global_step

    def run(epochs):
        main = ann.get_advice(init_fn=init_fn)(experiment(md, gpu, model="TinyMLP", loader="BlobPlaceholderLoader",
                                                          epochs=epochs, samples=64, batch=16,
                                                          opt=dtf.AdamOptimizer(0.001)))
        op = dtf.build_operator(main)
        op.extra_hooks = [FirstStep()]
        op.run()
        return op

    op = run(1)
    assert calls == [0] and seen["step0"] == 0
    assert torch.allclose(seen["w0"], w_src, atol=0), "the first run starts from the pretrained weights"
    assert op.trainer.store.global_step == 4
    w_after = op.trainer.store.vars["fc1/weights"].master.detach().cpu().clone()
    op2 = run(2)
    assert calls == [0], "a resumed run must not be re-initialised"
    assert seen["step0"] == 4 and torch.allclose(seen["w0"], w_after, atol=0)
    assert op2.trainer.store.global_step == 8


# ---------------------------------------------------------------------------------------------------------------------
# The first model most users write: a LeNet on 28x28x1 images, numpy batches, uint8 pixels, int32 labels.

class LeNet(dtf.Model):
    def inference(self, x, *a, **k):
        T = dtf.tools
        x = T.pool("p1", T.conv("c1", x, 1, 32, kernel_size=[5, 5]))      # 14x14x32
        x = T.pool("p2", T.conv("c2", x, 32, 64, kernel_size=[5, 5]))     # 7x7x64
        return T.fc_nonacti("logits", T.FC_layer("f1", x, 128), 10)


def _digits(n, seed):
    """Synthetic 'digits': class k lights up a k-dependent stripe pattern; uint8 pixels, int32 labels, numpy."""
    rng = np.random.RandomState(seed)
    y = rng.randint(0, 10, size=(n,)).astype(np.int32)
    x = rng.randint(0, 40, size=(n, 28, 28, 1)).astype(np.uint8)
    for i, c in enumerate(y):
        x[i, 2 * c + 2:2 * c + 6, :, 0] += 150
        x[i, :, 2 * c + 3:2 * c + 5, 0] += 60
    return x, y


def lenet_on_numpy_uint8_batches_learns(device, use_graph=False):
    kw = {"use_cuda_graph": True} if use_graph else {}
    tr = dtf.Trainer(dtf.Net(LeNet()), dtf.SoftmaxCrossEntropyLoss(), dtf.AdamOptimizer(2e-3), device=device, **kw)
    pool = [_digits(32, s) for s in range(4)]
    losses = []
    for i in range(40):
        b = pool[i % 4]
        tr.prefetch(pool[(i + 1) % 4])
        losses.append(float(tr.step(b)))
    assert all(l == l for l in losses) and losses[-1] < 0.5 * losses[0], (losses[0], losses[-1])
    x, y = _digits(64, 99)
    layers.set_training(False)
    try:
        with torch.no_grad(), V.using_store(tr.store):
            logits = tr.net.process(torch.from_numpy(x).to(tr.device)).float().cpu()
    finally:
        layers.set_training(True)
    acc = float((logits.argmax(1).numpy() == y).mean())
    assert acc > 0.8, acc
    return acc


def resnet_tiny_batches(device):
    """Batch 2 at 32x32 (the last stage is 1x1: batch norm over two values per channel), then inference on batch 1
    and on batch 3."""
    torch.manual_seed(21)
    tr = dtf.Trainer(dtf.Net(TinyResNet18()), dtf.SoftmaxCrossEntropyLoss(), dtf.MomentumOptimizer(0.01, 0.9),
                     device=device)
    x, y = torch.rand(2, 32, 32, 3), torch.randint(0, 16, (2,))
    losses = [float(tr.step((x, y))) for _ in range(4)]
    assert all(l == l and l < 1e4 for l in losses), losses
    layers.set_training(False)
    try:
        with torch.no_grad(), V.using_store(tr.store):
            for n in (1, 3):
                out = tr.net.process(torch.rand(n, 32, 32, 3).to(tr.device)).float().cpu()
                assert out.shape == (n, 16) and bool(torch.isfinite(out).all())
    finally:
        layers.set_training(True)
