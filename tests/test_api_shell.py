"""Annotation registry, flags, constants, plugin ABCs, learning-rate persistence, hooks, placement (CPU unit tier)."""
import json
import os
import time

import pytest
import torch

import distributed_tensorflow_framework_b200 as dtf
from distributed_tensorflow_framework_b200.api import annotations as ann
from distributed_tensorflow_framework_b200.api import constants, flags
from distributed_tensorflow_framework_b200.plugins import input as Input
from distributed_tensorflow_framework_b200.plugins import loss as Loss
from distributed_tensorflow_framework_b200.runtime import hooks as H
from distributed_tensorflow_framework_b200.runtime import placement


def test_each_annotation_sets_exactly_one_whitelisted_attribute():
    @ann.current_model(model="M", bogus=1)
    @ann.current_input(input="I")
    @ann.current_mode(mode="Train")
    @ann.current_feature(features={"a": 1})
    @ann.gpu_num(gpu_num=4)
    @ann.ps_hosts(ps_host="127.0.0.1: 22")       # reference typo (distribute.py:37) is tolerated
    @ann.worker_hosts(worker_hosts="a:1, b:2")
    @ann.job_name(job_name="worker")
    @ann.task_index(task_index=1)
    @ann.batch_size(batch_size=35)
    @ann.sample_number(sample_number=100)
    @ann.epoch_num(epoch_num=2)
    @ann.model_dir(model_dir="/tmp")
    @ann.data_dir(data_dir="/data")
    @ann.optimizer(optimizer=None)
    @ann.loss(loss="L")
    def main():
        pass

    g = ann.get_value_from_annotation
    assert (g(main, "model"), g(main, "input"), g(main, "mode"), g(main, "gpu_num")) == ("M", "I", "Train", 4)
    assert g(main, "ps_hosts") == "127.0.0.1: 22" and g(main, "features") == {"a": 1}
    assert g(main, "batch_size") == 35 and g(main, "optimizer") is None and g(main, "loss") == "L"
    assert not hasattr(main, "bogus")
    with pytest.raises(ValueError):
        g(main, "nonexistent")


def test_get_instance_uses_new_without_init_and_registry():
    class Probe(object):
        def __init__(self):
            raise AssertionError("__init__ must not run (reference uses cls.__new__, distribute_annotations.py:318-333)")

    class Holder(object):
        pass

    mod = type("M", (), {"Probe": Probe})

    @ann.current_model(model="Probe")
    def main():
        pass

    assert isinstance(ann.get_instance_from_annotation(main, "model", mod), Probe)
    ann.register_plugin(Holder, name="RegisteredHolder")

    @ann.loss(loss="RegisteredHolder")
    def main2():
        pass

    assert isinstance(ann.get_instance_from_annotation(main2, "loss", None), Holder)

    @ann.loss(loss="Nope")
    def main3():
        pass

    with pytest.raises(ValueError):
        ann.get_instance_from_annotation(main3, "loss", mod)


def test_get_advice_accepts_both_spellings_and_rejects_non_callables():
    fn = lambda *a: "x"  # noqa: E731

    @ann.get_advice(pre_fn=fn, post_processs_fn=fn, init_fn=None)
    @ann.parse_data_dir(parse_data_dir_fn=lambda d: d + "/train")
    class Op(object):
        pass

    op = Op()
    assert op.pre_fn() == "x" and op.post_process_fn(1) == "x" and op.parse_data_dir_fn("/d") == "/d/train"
    with pytest.raises(TypeError):
        ann.get_advice(pre_fn=print("evaluated at import like distribute_train.py:22") or 3)(Op)


def test_flag_defaults_match_reference_and_parse():
    flags.FLAGS.reset()
    F = flags.FLAGS
    assert (F.use_fp16, F.job_name, F.ps_hosts, F.worker_hosts, F.task_index) == (False, "", "", "", None)
    assert F.replicas_to_aggregate is None and F.log_device_placement is False
    assert (F.input_image_height, F.input_image_width, F.sample_number) == (224, 224, 100000)
    assert F.train_learning_rate == 0.001 and F.learning_rate_json == "./learning_rate.json"
    assert getattr(F, "intra_op_parallelism_threads ") == 0  # trailing-space spelling of the reference still resolves
    rest = flags.parse(["--job_name=ps", "--task_index", "3", "--replicas_to_aggregate=2", "--unknown=1"])
    assert F.job_name == "ps" and F.task_index == 3 and F.replicas_to_aggregate == 2 and rest == ["--unknown=1"]
    flags.FLAGS.reset()
    assert constants.TOWER_NAME == "tower" and constants.MIN_FRACTION_OF_EXAMPLE_IN_QUEUE == 0.05
    assert constants.initial_learning_rate() == 0.001


def test_learning_rate_json_roundtrip(tmp_path):
    path = str(tmp_path / "lr.json")
    lr = dtf.LearningRate(0.01, path, decay_factor=0.5)
    assert lr.learning_rate == 0.01 and lr.load() == 0.01  # file missing: guarded
    lr.save(0.02)
    assert json.load(open(path)) == {"learning_rate": 0.02}
    assert dtf.LearningRate(0.5, path).learning_rate == 0.02  # an existing file wins over the initial value
    lr.update(0.004)
    assert lr.decay() == 0.002
    opt = dtf.GradientDescentOptimizer(lr)
    assert opt.learning_rate == 0.002


def test_net_process_composes_aspects():
    calls = []

    class M(dtf.Model):
        def inference(self, x, *a, **k):
            calls.append("inference")
            return x * 2

    class N(dtf.Net):
        def pre_process(self, x, *a, **k):
            calls.append("pre")
            return x + 1

        def post_process(self, y, *a, **k):
            calls.append("post")
            return y - 1

    assert N(M()).process(3) == 7 and calls == ["pre", "inference", "post"]
    assert ann.registered("M") is M  # subclasses of Model self-register
    with pytest.raises(TypeError):
        dtf.Model()  # abstract


def test_builtin_losses_cpu():
    logits = torch.randn(8, 5)
    labels = torch.randint(0, 5, (8,))
    assert torch.allclose(Loss.SoftmaxCrossEntropyLoss().loss(logits, labels),
                          torch.nn.functional.cross_entropy(logits, labels))
    a, b = torch.randn(2, 4, 4, 3), torch.randn(2, 4, 4, 3)
    assert torch.allclose(Loss.MSELoss().loss(a, b), ((a - b) ** 2).mean())


def test_examples_per_second_hook_arithmetic(monkeypatch):
    with pytest.raises(ValueError):
        H.ExamplesPerSecondHook(32, every_n_steps=10, every_n_secs=5)
    with pytest.raises(ValueError):
        H.ExamplesPerSecondHook(32, every_n_steps=None, every_n_secs=None)
    now = [100.0]
    monkeypatch.setattr(time, "time", lambda: now[0])
    hook = H.ExamplesPerSecondHook(batch_size=32, every_n_steps=10)

    class S(object):
        global_step = 0

    ctx = H.RunContext(S())
    for step in range(0, 31):
        S.global_step = step
        hook.after_run(ctx, None)
        now[0] += 0.5
    # triggers at steps 0, 10, 20, 30: each interval is 10 steps in 5 s -> 2 steps/s -> 64 examples/s
    assert hook.current_examples_per_sec == pytest.approx(64.0)
    assert hook.average_examples_per_sec == pytest.approx(64.0)


def test_stop_at_step_and_timer():
    t = H.SecondOrStepTimer(every_steps=5)
    assert t.should_trigger_for_step(0)
    t.update_last_triggered_step(0)
    assert not t.should_trigger_for_step(4) and t.should_trigger_for_step(5)
    with pytest.raises(ValueError):
        H.StopAtStepHook()

    class S(object):
        global_step = 7

    h = H.StopAtStepHook(last_step=7)
    ctx = H.RunContext(S())
    h.after_run(ctx, None)
    assert ctx.stop_requested


def test_local_device_setter_round_robin_and_owner_map():
    chooser = placement.local_device_setter(num_devices=3, ps_device_type="gpu", worker_device="/gpu:0")
    got = [chooser(("VariableV2", "v%d" % i)) for i in range(5)] + [chooser(("Conv2D", "c"))]
    assert got == ["/gpu:0", "/gpu:1", "/gpu:2", "/gpu:0", "/gpu:1", "/gpu:0"]
    with pytest.raises(TypeError):
        placement.local_device_setter(ps_strategy=3)
    rds = placement.replica_device_setter(2, worker_device="/job:worker/task:1")
    assert [rds(("VariableV2", i)) for i in range(3)] == ["/job:ps/task:0", "/job:ps/task:1", "/job:ps/task:0"]
    assert rds(("MatMul", 0)) == "/job:worker/task:1"

    class V(object):
        def __init__(self, n):
            self.numel = n

    assert placement.owner_map([V(1)] * 5, 2) == [0, 1, 0, 1, 0]
    assert placement.owner_map([V(100), V(1), V(1), V(1)], 2, policy="bytes") == [0, 1, 1, 1]


def test_loader_kind_dispatch_and_cluster_spec():
    assert Input.TFRecordDataLoader.type == "TFRecordDataLoader"
    assert Input.PlaceholderDataLoader.type == "PlaceholderDataLoader"
    assert Input.DataPathDataLoader.type == "DataPathDataLoader"
    c = dtf.ClusterSpec({"ps": "127.0.0.1: 22", "worker": "127.0.0.1:23, 127.0.0.2: 24"})
    assert c.num_tasks("worker") == 2 and c.job_tasks("ps") == ["127.0.0.1:22"]  # whitespace stripped (Q4)
    assert c.task_address("worker", 1) == "127.0.0.2:24"


def test_run_config_uid_is_stable():
    a = dtf.RunConfig(model_dir="/m", cluster_spec=dtf.ClusterSpec({"worker": "b:1,a:2", "ps": "p:1"}), foo=1,
                      save_summary_steps=5)
    b = dtf.RunConfig(model_dir="/m", cluster_spec=dtf.ClusterSpec({"ps": "p:1", "worker": "b:1,a:2"}), foo=1,
                      save_summary_steps=99)
    assert a.uid() == b.uid() and "_foo=1" in a.uid()


def test_fault_injection_spec_and_nvtx_noop():
    from distributed_tensorflow_framework_b200.utils import fault, nvtx
    assert fault.parse("") is None and fault.parse("1:20") == (1, 20)
    with pytest.raises(ValueError):
        fault.parse("oops")
    died = []
    assert not fault.maybe_kill(0, 20, "1:20", _exit=died.append)
    assert fault.maybe_kill(1, 20, "1:20", _exit=died.append) and died == [17]
    with nvtx.range("unit"):
        nvtx.mark("m")
