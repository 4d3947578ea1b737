"""CPU tier for the round-2 framework paths: Estimator over Trainer / Saver / session, the learning-rate handler wired
through the container into the optimizer and the checkpoint, the collective checkpoint hook (every rank saves at the
agreed step, only the chief writes), the control-plane `agree`, tower spawning decision, watchdog hook."""
import json
import os

import pytest
import torch

import distributed_tensorflow_framework_b200 as dtf
from distributed_tensorflow_framework_b200 import models
from distributed_tensorflow_framework_b200.api import annotations as ann
from distributed_tensorflow_framework_b200.api import flags
from distributed_tensorflow_framework_b200.io import tensorbundle as tb
from distributed_tensorflow_framework_b200.plugins.learningrate import LearningRate
from distributed_tensorflow_framework_b200.runtime import container, hooks as H
from distributed_tensorflow_framework_b200.runtime.cluster import ClusterSpec, Server
from distributed_tensorflow_framework_b200.runtime.estimator import Estimator, EstimatorSpec
from distributed_tensorflow_framework_b200.runtime.session import WorkerLostError


def _cifar_batches(n, batch=8, seed=0):
    g = torch.Generator().manual_seed(seed)
    return [(torch.rand(batch, 32, 32, 3, generator=g), torch.randint(0, 10, (batch,), generator=g)) for _ in range(n)]


def test_estimator_trains_checkpoints_resumes_and_evaluates(tmp_path):
    model_dir = str(tmp_path)

    def model_fn(features, labels, mode, params):
        return EstimatorSpec(net=models.Cifar10CNN(), loss=dtf.SoftmaxCrossEntropyLoss(),
                             optimizer=dtf.MomentumOptimizer(params["lr"], 0.9))

    data = _cifar_batches(2)
    est = Estimator(model_fn, model_dir=model_dir, params={"lr": 0.01}, device="cpu")
    est.train(lambda: iter(data * 3), steps=6)
    prefix = tb.latest_checkpoint(model_dir)
    assert prefix is not None and prefix.endswith("model.ckpt-6")
    with tb.BundleReader(prefix) as r:
        assert int(r.get_tensor("global_step").reshape(-1)[0]) == 6 and "conv1/weights" in r.keys()
        assert "conv1/weights/Momentum" in r.keys()
    ev = est.evaluate(lambda: iter(data))
    assert ev["global_step"] == 6 and ev["loss"] == ev["loss"] and 0.0 <= ev["accuracy"] <= 1.0
    # a second estimator on the same model_dir resumes: global_step continues from the checkpoint
    est2 = Estimator(model_fn, model_dir=model_dir, params={"lr": 0.01}, device="cpu")
    est2.train(lambda: iter(data * 2), steps=3)
    assert tb.latest_checkpoint(model_dir).endswith("model.ckpt-9")
    preds = list(est2.predict(lambda: iter(data[:1])))
    assert preds[0].shape == (8, 10)


def test_learning_rate_handler_reaches_optimizer_and_checkpoint(tmp_path):
    """@learning_rate(...) is read by the container, the optimizer follows the handler's current value at every
    step, and the JSON file is written with the checkpoint (distribute_learningrate.py:12-50)."""
    flags.FLAGS.reset()
    lr_path = str(tmp_path / "lr.json")
    handler = LearningRate(0.5, lr_path, decay_factor=0.1)
    model_dir = str(tmp_path / "m")
    os.makedirs(model_dir)
    seen = []

    class _Probe(H.SessionRunHook):
        def after_run(self, ctx, values):
            seen.append(op.optimizer.learning_rate)
            if ctx.session.global_step == 2:
                handler.decay()

    @ann.current_model(model=models.Cifar10CNN)
    @ann.optimizer(optimizer=dtf.GradientDescentOptimizer(123.0))
    @ann.loss(loss="SoftmaxCrossEntropyLoss")
    @ann.current_mode(mode="Train")
    @ann.current_input(input=models.Cifar10SyntheticLoader)
    @ann.gpu_num(gpu_num=0)
    @ann.batch_size(batch_size=4)
    @ann.sample_number(sample_number=16)
    @ann.epoch_num(epoch_num=1)
    @ann.model_dir(model_dir=model_dir)
    @ann.data_dir(data_dir="")
    @ann.learning_rate(learning_rate=handler)
    def main(argv=None):
        pass

    op = container.build_operator(main)
    assert op.learning_rate is handler and op.optimizer._lr is handler  # a constant lr is replaced by the handler
    op.extra_hooks = [_Probe()]
    op.run()
    assert seen[:2] == [0.5, 0.5] and abs(seen[2] - 0.05) < 1e-12 and abs(seen[3] - 0.05) < 1e-12
    assert abs(json.load(open(lr_path))["learning_rate"] - 0.05) < 1e-12  # persisted with the final checkpoint
    assert abs(LearningRate(0.5, lr_path).learning_rate - 0.05) < 1e-12  # a restarted run picks it up
    flags.FLAGS.reset()


class _FakeSaver(object):
    def __init__(self):
        self.calls = []

    def save(self, d, step, write=True):
        self.calls.append((step, write))


class _Sess(object):
    def __init__(self):
        self.global_step = 0


def test_collective_checkpoint_hook_saves_on_every_rank_at_the_agreed_step():
    """Wall-clock cadence in collective mode: the chief's decision is adopted by every rank at probe steps only; the
    non-chief never writes. Step cadence needs no agreement."""
    decisions = {}

    def agree_chief(step, mine):
        decisions[step] = mine
        return mine

    def agree_worker(step, mine):
        assert mine is False  # a worker never decides
        return decisions[step]

    sv0, sv1 = _FakeSaver(), _FakeSaver()
    h0 = H.CheckpointSaverHook(sv0, "d", save_secs=0, write=True, collective=True, agree_fn=agree_chief, probe_every=5)
    h1 = H.CheckpointSaverHook(sv1, "d", save_secs=0, write=False, collective=True, agree_fn=agree_worker,
                               probe_every=5)
    s0, s1 = _Sess(), _Sess()
    h0.after_create_session(s0)
    h1.after_create_session(s1)
    for step in range(1, 13):
        s0.global_step = s1.global_step = step
        h0.after_run(H.RunContext(s0), None)
        h1.after_run(H.RunContext(s1), None)
    assert [c[0] for c in sv0.calls] == [5, 10] == [c[0] for c in sv1.calls]
    assert all(w for _, w in sv0.calls) and not any(w for _, w in sv1.calls)
    h0.end(s0)
    h1.end(s1)
    assert sv0.calls[-1] == (12, True) and sv1.calls[-1] == (12, False)
    # step cadence: deterministic on every rank
    sv = _FakeSaver()
    h = H.CheckpointSaverHook(sv, "d", save_secs=None, save_steps=4, write=False, collective=True)
    s = _Sess()
    h.after_create_session(s)
    for step in range(1, 10):
        s.global_step = step
        h.after_run(H.RunContext(s), None)
    assert [c[0] for c in sv.calls] == [4, 8]


def test_server_agree_without_store_and_spawn_decision(monkeypatch):
    srv = Server(ClusterSpec({"ps": "", "worker": "127.0.0.1:29999"}), "worker", 0)
    assert srv.control_store() is None and srv.agree("ckpt", 3, True, True) is True

    def main():
        pass
    main.gpu_num, main.mode = 4, "Train"
    monkeypatch.delenv("WORLD_SIZE", raising=False)
    monkeypatch.delenv("TORCHELASTIC_RUN_ID", raising=False)
    monkeypatch.delenv("DTF_SPAWNED", raising=False)
    assert container._should_spawn(main) == 0  # no GPUs here: towers run in-process
    monkeypatch.setattr(torch.cuda, "is_available", lambda: True)
    monkeypatch.setattr(torch.cuda, "device_count", lambda: 8)
    assert container._should_spawn(main) == 4
    main.worker_hosts = "a:1,b:2"  # a real multi-worker cluster is launched by the user, not spawned
    assert container._should_spawn(main) == 0
    main.worker_hosts = "a:1"
    monkeypatch.setenv("DTF_SPAWNED", "1")
    assert container._should_spawn(main) == 0


def test_watchdog_hook_raises_worker_lost_from_the_error_record():
    class _Comm(object):
        world = 2

        def __init__(self):
            self.bad = False

        def check(self):
            if self.bad:
                raise WorkerLostError("barrier wait timed out")

        def heartbeat(self, beat):
            return torch.tensor([beat, beat])

    c = _Comm()
    h = H.HeartbeatHook(c, every_n_steps=2, timeout_s=1e9)
    s = _Sess()
    s.global_step = 2
    h.after_run(H.RunContext(s), None)
    c.bad = True
    with pytest.raises(WorkerLostError):
        h.after_run(H.RunContext(s), None)


def test_sync_mode_annotation_is_read(tmp_path):
    from distributed_tensorflow_framework_b200.runtime.train import _sync_mode
    flags.FLAGS.reset()
    assert _sync_mode("async") == "async" and _sync_mode(None) == "sync"
    flags.FLAGS.sync_mode = "backup"
    assert _sync_mode("async") == "backup"  # the flag wins
    flags.FLAGS.reset()
    flags.FLAGS.sync_replicas = False
    assert _sync_mode(None) == "async"
    flags.FLAGS.reset()
