"""Runtime tier on CPU: sync / backup-worker / async protocol on the in-process fake backend, the Train operator end
to end (BASELINE.json config 1), TF1-layout checkpoints + resume, the Eval operator, ps / worker processes over
loopback with gloo, input loaders, experiment / estimator wrappers."""
import os
import struct
import subprocess
import sys
import threading
import time

import numpy as np
import pytest
import torch

import distributed_tensorflow_framework_b200 as dtf
from distributed_tensorflow_framework_b200 import models
from distributed_tensorflow_framework_b200.api import annotations as ann
from distributed_tensorflow_framework_b200.engine import variables as V
from distributed_tensorflow_framework_b200.io import tensorbundle as tb
from distributed_tensorflow_framework_b200.io import tfrecord as tfr
from distributed_tensorflow_framework_b200.plugins import input as Input
from distributed_tensorflow_framework_b200.runtime.sync_replicas import (AsyncReplicasOptimizer,
                                                                         InProcessParameterServer,
                                                                         SyncReplicasOptimizer)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ps(n_workers, r2a, lr=1.0):
    store = V.VariableStore("cpu")
    with V.using_store(store):
        v = store.get_variable("w", [4], V.ConstantInitializer(0.0))
    ps = InProcessParameterServer(store, dtf.GradientDescentOptimizer(lr), r2a, n_workers)
    return store, v, ps


def test_sync_replicas_all_workers_mean_and_token_barrier():
    store, v, ps = _ps(3, 3)
    ps.start_chief_queue_runner()
    steps_seen = [[] for _ in range(3)]

    def worker(i):
        opt = SyncReplicasOptimizer(dtf.GradientDescentOptimizer(1.0), 3, 3, backend=ps)
        for _ in range(4):
            g = torch.full((4,), float(i + 1))
            steps_seen[i].append(opt.apply_gradients([(g, v)]))

    ts = [threading.Thread(target=worker, args=(i,), daemon=True) for i in range(3)]
    [t.start() for t in ts]
    [t.join(10) for t in ts]
    ps.stop()
    assert ps.global_step == 4 and all(s == [1, 2, 3, 4] for s in steps_seen)  # tokens carry the new global step
    assert torch.allclose(v.master, torch.full((4,), -4 * 2.0))  # mean(1,2,3) = 2 per step, lr 1
    assert ps.tokens_per_step == 3


def test_backup_workers_drop_the_stale_straggler():
    """Deterministic schedule on the fake backend: aggregate the 2 freshest of 3 workers."""
    store, v, ps = _ps(3, 2)
    g = lambda val: [(torch.full((4,), val), v)]  # noqa: E731
    assert ps.tokens_per_step == 3  # max(total_num_replicas, replicas_to_aggregate)
    # step 1: workers A and B arrive, the straggler C does not
    assert ps.push(g(1.0), local_step=0) == 1 and ps.push(g(3.0), local_step=0) == 1
    assert ps.aggregate_and_apply(timeout=1.0) == 1
    assert torch.allclose(v.master, torch.full((4,), -2.0))  # mean(1, 3)
    assert [ps.take_token(1.0) for _ in range(3)] == [1, 1, 1]  # one token per replica, carrying the new step
    # C finally pushes the gradient it computed from step-0 weights: stale, dropped, nothing changes
    assert ps.push(g(1000.0), local_step=0) == 0 and ps.accumulators["w"].dropped == 1
    with pytest.raises(TimeoutError):
        ps.aggregate_and_apply(timeout=0.1)
    assert ps.global_step == 1 and torch.allclose(v.master, torch.full((4,), -2.0))
    # step 2: C has caught up (its token said step 1); now A is the slow one. All three arrive before the chief
    # looks: take_grad averages everything that was accumulated (>= replicas_to_aggregate), like TF.
    for val in (5.0, 7.0, 9.0):
        assert ps.push(g(val), local_step=1) == 1
    assert ps.aggregate_and_apply(timeout=1.0) == 2
    assert torch.allclose(v.master, torch.full((4,), -2.0 - 7.0))


def test_async_mode_applies_every_push_and_bounds_staleness():
    store, v, ps = _ps(2, 1)
    a = AsyncReplicasOptimizer(dtf.GradientDescentOptimizer(1.0), 2, backend=ps, max_staleness=1)
    assert a.make_session_run_hook(True) is None
    for i in range(3):
        a.apply_gradients([(torch.ones(4), v)], computed_at_step=ps.global_step)
    assert ps.global_step == 3 and torch.allclose(v.master, torch.full((4,), -3.0))
    a.apply_gradients([(torch.ones(4), v)], computed_at_step=0)  # 3 steps stale > max_staleness
    assert a.dropped == 1 and ps.global_step == 3


def _cifar_main(model_dir, epochs=1, samples=128, batch=16, gpu_num=0, opt=None, **advice):
    @ann.current_model(model="Cifar10CNN")
    @ann.optimizer(optimizer=opt or dtf.MomentumOptimizer(0.02, 0.9))
    @ann.loss(loss="SoftmaxCrossEntropyLoss")
    @ann.current_mode(mode="Train")
    @ann.current_input(input="Cifar10SyntheticLoader")
    @ann.gpu_num(gpu_num=gpu_num)
    @ann.ps_hosts(ps_hosts="")
    @ann.worker_hosts(worker_hosts="127.0.0.1:29617")
    @ann.job_name(job_name="worker")
    @ann.task_index(task_index=0)
    @ann.batch_size(batch_size=batch)
    @ann.sample_number(sample_number=samples)
    @ann.epoch_num(epoch_num=epochs)
    @ann.model_dir(model_dir=model_dir)
    @ann.data_dir(data_dir="")
    def main():
        pass

    for k, fn in advice.items():
        setattr(main, k, fn)
    return main


def test_train_operator_end_to_end_checkpoint_resume_and_eval(tmp_path):
    model_dir = str(tmp_path / "model")
    with pytest.raises(ValueError):
        dtf.build_operator(_cifar_main(model_dir))  # model_dir must exist (distribute.py:74-75)
    os.makedirs(model_dir)
    events = []
    main = _cifar_main(model_dir, epochs=2, samples=128, batch=16,
                       pre_fn=lambda a, k: events.append("pre") or "ctx",
                       post_fn=lambda a, k: events.append("post"),
                       init_fn=lambda scaffold, sess: events.append("init"),
                       pre_process_fn=lambda raw, gt, a, k: (raw * 1.0, gt),
                       post_process_fn=lambda logits: logits)
    op = dtf.build_operator(main)
    assert op.input_mode == dtf.InputOptions.PLACEHOLDER and isinstance(op.net, dtf.Net) and op.gpu_num == 0
    op.run()
    assert events == ["pre", "init", "post"]
    total_step = 2 * (128 // 16)
    assert op.trainer.store.global_step == total_step == 16
    first_losses = op.final_loss
    # TF1 Saver layout on disk, keys = variable names + slots + global_step
    latest = tb.latest_checkpoint(model_dir)
    assert latest.endswith("model.ckpt-16") and os.path.exists(latest + ".data-00000-of-00001")
    with tb.BundleReader(latest) as r:
        keys = set(r.keys())
        assert {"conv1/weights", "conv1/biases", "local3/weights", "softmax_linear/biases", "global_step",
                "conv1/weights/Momentum"} <= keys
        assert int(r.get_tensor("global_step").reshape(-1)[0]) == 16 and r.get_tensor("conv1/weights").shape == (5, 5, 3, 64)
        saved_w = r.get_tensor("conv2/weights")
    assert any(f.startswith("events.out.tfevents") for f in os.listdir(model_dir))
    # resume: a second run restores, continues the global step and stops at the new total
    main2 = _cifar_main(model_dir, epochs=3, samples=128, batch=16)
    op2 = dtf.build_operator(main2)
    op2.run()
    assert op2.trainer.store.global_step == 24
    assert tb.latest_checkpoint(model_dir).endswith("model.ckpt-24")
    assert op2.final_loss < 2.4 and first_losses == first_losses
    # Eval operator restores the same checkpoint and reports loss / accuracy
    main2.mode = "Eval"
    ev = dtf.build_operator(main2)
    res = ev.run()
    assert res["global_step"] == 24 and res["batches"] == 4 and 0.0 <= res["accuracy"] <= 1.0 and res["loss"] > 0
    assert os.path.isdir(os.path.join(model_dir, "eval"))
    assert saved_w.shape == (5, 5, 64, 64)


def test_multi_tower_cpu_equals_mean_of_tower_gradients():
    torch.manual_seed(0)
    xa, ya = torch.rand(4, 32, 32, 3), torch.randint(0, 10, (4,))
    xb, yb = torch.rand(4, 32, 32, 3), torch.randint(0, 10, (4,))
    t2 = dtf.Trainer(dtf.Net(models.Cifar10CNN()), dtf.SoftmaxCrossEntropyLoss(), dtf.GradientDescentOptimizer(0.1),
                     gpu_num=2, seed=4)
    t1 = dtf.Trainer(dtf.Net(models.Cifar10CNN()), dtf.SoftmaxCrossEntropyLoss(), dtf.GradientDescentOptimizer(0.1),
                     gpu_num=1, seed=4)
    t2.step([(xa, ya), (xb, yb)])
    t1.step((torch.cat([xa, xb]), torch.cat([ya, yb])))  # mean over towers == gradient of the concatenated batch
    for k, v in t1.store.state_dict().items():
        assert torch.allclose(v, t2.store.state_dict()[k], atol=1e-5), k


def _free_ports(n):
    import socket
    socks, ports = [], []
    for _ in range(n):
        sk = socket.socket()
        sk.bind(("127.0.0.1", 0))
        socks.append(sk)
        ports.append(sk.getsockname()[1])
    for sk in socks:
        sk.close()
    return ports


def test_ps_and_two_workers_over_loopback(tmp_path):
    p_ps, p_w0, p_w1 = _free_ports(3)
    script = tmp_path / "job.py"
    script.write_text('''
import os, sys
sys.path.insert(0, %r)
import torch
import distributed_tensorflow_framework_b200 as dtf
from distributed_tensorflow_framework_b200 import annotations as ann, flags, models
flags.parse()
md = os.path.join(%r, "m")
os.makedirs(md, exist_ok=True)
@ann.current_model(model="Cifar10CNN")
@ann.optimizer(optimizer=dtf.GradientDescentOptimizer(0.05))
@ann.loss(loss="SoftmaxCrossEntropyLoss")
@ann.current_mode(mode="Train")
@ann.current_input(input="Cifar10SyntheticLoader")
@ann.gpu_num(gpu_num=0)
@ann.ps_hosts(ps_hosts="127.0.0.1:%d")
@ann.worker_hosts(worker_hosts="127.0.0.1:%d, 127.0.0.1:%d")
@ann.job_name(job_name=flags.FLAGS.job_name)
@ann.task_index(task_index=flags.FLAGS.task_index)
@ann.batch_size(batch_size=8)
@ann.sample_number(sample_number=32)
@ann.epoch_num(epoch_num=1)
@ann.model_dir(model_dir=md)
@ann.data_dir(data_dir="")
def main(): pass
op = dtf.run_main(main)
if flags.FLAGS.job_name == "worker":
    w = op.trainer.store.vars["local4/weights"].master
    print("RESULT", flags.FLAGS.task_index, op.trainer.store.global_step, "%%.8f" %% float(w.double().sum()), flush=True)
else:
    print("PS_EXIT", flush=True)
''' % (ROOT, str(tmp_path), p_ps, p_w0, p_w1))
    env = dict(os.environ, OMP_NUM_THREADS="2")
    import time as _time
    logs = [tmp_path / ("out_%d.log" % i) for i in range(3)]
    files = [open(l, "w") for l in logs]
    # output goes to files (a full pipe can never block a rank) and the three ranks share ONE deadline
    procs = [subprocess.Popen([sys.executable, str(script), "--job_name=%s" % j, "--task_index=%d" % t],
                              stdout=f, stderr=subprocess.STDOUT, text=True, env=env)
             for f, (j, t) in zip(files, (("ps", 0), ("worker", 0), ("worker", 1)))]
    try:
        deadline = _time.time() + 240
        while _time.time() < deadline and any(p.poll() is None for p in procs):
            _time.sleep(0.2)
    finally:
        for p in procs:
            if p.poll() is None:
                p.kill()
                p.wait()
        for f in files:
            f.close()
    outs = [l.read_text() for l in logs]
    assert all(p.returncode == 0 for p in procs), "\\n".join(o[-1500:] for o in outs)
    assert "PS_EXIT" in outs[0]  # the ps left after both workers signalled the done queue
    r0 = [l for l in outs[1].splitlines() if l.startswith("RESULT")][0].split()
    r1 = [l for l in outs[2].splitlines() if l.startswith("RESULT")][0].split()
    assert r0[2] == r1[2] == "4" and r0[3] == r1[3]  # same global step, identical replicas after sync training


class _RecordLoader(Input.TFRecordDataLoader):
    def _decode_raw_data(self, raw_features, height, width, *args, **kwargs):
        img = np.frombuffer(raw_features["image_raw"], dtype=np.uint8).reshape(height, width, 3)
        return [img.astype(np.float32) / 255.0, np.int64(raw_features["label"])]


def test_tfrecord_loader_with_user_decode_hook(tmp_path):
    path = str(tmp_path / "train.tfrecord")
    with tfr.TFRecordWriter(path) as w:
        for i in range(20):
            img = np.full((8, 8, 3), i, dtype=np.uint8)
            w.write(tfr.make_example({"image_raw": img.tobytes(), "label": [i]}))
    dtf.FLAGS.input_image_height, dtf.FLAGS.input_image_width = 8, 8
    try:
        loader = _RecordLoader.__new__(_RecordLoader)
        loader.batch_size, loader.sample_number = 5, 20
        loader.features = {"image_raw": tfr.FixedLenFeature([], "string"), "label": tfr.FixedLenFeature([], "int64")}
        with pytest.raises(RuntimeError):
            loader.load_train_batch(None)
        q = Input.FilenameQueue([path])
        seen = set()
        for _ in range(8):
            raw, gt = loader.load_train_batch(q)
            assert raw.shape == (5, 8, 8, 3) and gt.shape == (5,)
            assert torch.allclose(raw[:, 0, 0, 0] * 255.0, gt.float())  # image i carries label i
            seen |= set(gt.tolist())
        assert len(seen) > 10
    finally:
        dtf.FLAGS.reset()


class _LegacyLoader(Input.TFRecordDataLoader):
    # the reference's name-mangled hook spelling (distribute_input.py:85) still resolves
    def _TFRecordDataLoader__decode_raw_data(self, raw_features, height, width, *a, **k):
        return [np.zeros((2,), np.float32), np.int64(raw_features["label"])]


def test_legacy_mangled_hook_and_placeholder_and_datapath(tmp_path):
    assert Input._find_hook(_LegacyLoader.__new__(_LegacyLoader), ["TFRecordDataLoader"],
                            "_decode_raw_data") is not None

    class PH(Input.PlaceholderDataLoader):
        def _put_names_dict_into_queue(self, q, *a, **k):
            for i in range(6):
                q.put({"i": i})

        def decode_data_from_path_name(self, paths, *a, **k):
            return {"raw_data": np.full((2, 2), paths["i"], np.float32), "ground_truth": np.int64(paths["i"])}

    ph = PH.__new__(PH)
    ph.batch_size = 4
    q = ph.load_queue_for_placeholder("dir")
    raw, gt = ph.load_placeholder_data(q)
    assert gt.tolist() == [0, 1, 2, 3] and raw.shape == (4, 2, 2) and q.qsize() == 6  # paths re-queued at the tail
    assert ph.load_placeholder_data(q)[1].tolist() == [4, 5, 0, 1]

    class DP(Input.DataPathDataLoader):
        def _create_reader(self):
            return lambda name: int(name.split("_")[1])

        def _parse_raw_data(self, value):
            return [np.full((3,), value, np.float32), np.int64(value)]

        def create_name_queue(self, data_dir):
            return Input.FilenameQueue(["s_%d" % i for i in range(10)])

    dp = DP.__new__(DP)
    dp.batch_size, dp.sample_number = 4, 10
    raw, gt = dp.load_train_batch(dp.create_name_queue("x"))
    assert raw.shape == (4, 3) and torch.allclose(raw[:, 0], gt.float())
    with pytest.raises(AssertionError):
        dp.load_train_batch(None)


def test_experiment_and_estimator_wrappers():
    calls = []
    with pytest.raises(ValueError):
        dtf.DistributeExperiment("Bogus")
    exp = dtf.DistributeExperiment("Train", train_fn=lambda loader, mode, pre, post: calls.append((type(loader).__name__, mode)) or 7,
                                   train_dataloader="Cifar10SyntheticLoader")
    assert exp.run() == 7 and calls == [("Cifar10SyntheticLoader", dtf.InputOptions.PLACEHOLDER)]
    with pytest.raises(ValueError):
        dtf.DistributeExperiment("Train", train_fn=lambda *a: 0, train_dataloader=_RecordLoader)  # needs features
    ev = dtf.DistributeExperiment("Eval", eval_fn=lambda loader, pre, post: "evaluated",
                                  eval_dataloader=models.Cifar10SyntheticLoader)
    assert ev.run() == "evaluated"
    est = dtf.DistributeEstimator(model_fn=lambda f, l, mode, p: (mode, f + p["k"]),
                                  eval_model_fn=lambda f, l, mode, p: ("custom_eval", f), params={"k": 10},
                                  input_class="Cifar10SyntheticLoader")
    assert isinstance(est.input_class, dtf.Dataloader)
    data = lambda: iter([(1, None), (2, None)])  # noqa: E731
    assert est.train(data, steps=2) == ("train", 12)
    assert est.evaluate(data) == [("custom_eval", 1), ("custom_eval", 2)]
    with pytest.raises(AssertionError):
        dtf.DistributeEstimator(lambda *a: 0, input_class=object())


def test_session_recovers_from_lost_worker(tmp_path):
    from distributed_tensorflow_framework_b200.runtime.session import MonitoredTrainingSession, WorkerLostError
    store = V.VariableStore("cpu")
    with V.using_store(store):
        v = store.get_variable("w", [2], V.ConstantInitializer(1.0))
    opt = dtf.GradientDescentOptimizer(0.5)
    fail = {"at": 3}

    def step():
        if store.global_step == fail["at"]:
            fail["at"] = -1
            with torch.no_grad():
                v.master.fill_(999.0)  # state corrupted by the failed step
            raise WorkerLostError("rank 1 silent")
        opt.apply_gradients([(torch.ones(2), v)], store=store)
        return {"loss": 0.0}

    with MonitoredTrainingSession(is_chief=True, checkpoint_dir=str(tmp_path), store=store, optimizer=opt,
                                  save_checkpoint_secs=None, save_checkpoint_steps=1,
                                  hooks=[dtf.StopAtStepHook(last_step=5)]) as sess:
        while not sess.should_stop():
            sess.run(step)
    assert store.global_step == 5 and torch.allclose(v.master, torch.full((2,), 1.0 - 2.5))


def test_metrics_jsonl_hook_writes_one_record_per_interval(tmp_path):
    import json
    from distributed_tensorflow_framework_b200.runtime import hooks as H

    class FakeSession(object):
        global_step = 0

    sess, path = FakeSession(), str(tmp_path / "metrics.jsonl")
    hook = H.MetricsJsonlHook(path, batch_size=32, every_n_steps=5, device_ms_fn=lambda: 1.5, extra={"rank": 0})
    hook.begin()
    hook.after_create_session(sess)
    ctx = H.RunContext(sess)
    for step in range(1, 21):
        sess.global_step = step
        hook.after_run(ctx, {"loss": 2.0 / step, "summaries": None})
    hook.end(sess)
    recs = [json.loads(l) for l in open(path)]
    assert [r["step"] for r in recs] == [5, 10, 15, 20]
    assert all(r["rank"] == 0 and r["device_ms_per_step"] == 1.5 and r["examples_per_sec"] > 0 for r in recs)
    assert abs(recs[-1]["loss"] - 0.1) < 1e-9
