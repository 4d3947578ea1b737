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
    cpu = dtf.Trainer(dtf.Net(getattr(__import__(__name__, fromlist=["x"]), model)()), dtf.SoftmaxCrossEntropyLoss(),
                      dtf.GradientDescentOptimizer(0.1), device="cpu")
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
