"""User-level scenarios on the CPU (oracle) path; the same bodies run on the CUDA path in test_gpu_scenarios.py."""
import scenarios


def test_resnet_train_eval_and_checkpoint_restore(tmp_path):
    scenarios.resnet_train_eval_and_cross_device_restore(tmp_path, gpu=False)


def test_placeholder_and_datapath_loaders_learn(tmp_path):
    scenarios.loader_kinds_learn_the_toy_problem(tmp_path, gpu=False)


def test_adam_follows_a_decaying_learning_rate_handler(tmp_path):
    scenarios.adam_with_decaying_learning_rate(tmp_path, gpu=False)


def test_two_in_process_towers_equal_one_tower_on_the_concatenated_batch():
    scenarios.in_process_towers_equal_one_big_batch("cpu")


def test_restoring_into_a_model_of_another_shape_names_the_variable(tmp_path):
    import pytest
    import torch
    import distributed_tensorflow_framework_b200 as dtf
    from distributed_tensorflow_framework_b200 import models

    class TwentyClasses(models.Cifar10CNN):
        num_classes = 20

    x, y = torch.rand(4, 32, 32, 3), torch.randint(0, 10, (4,))
    a = dtf.Trainer(dtf.Net(models.Cifar10CNN()), dtf.SoftmaxCrossEntropyLoss(), dtf.GradientDescentOptimizer(0.1),
                    device="cpu")
    a.step((x, y))
    prefix = dtf.Saver(a.store).save(str(tmp_path), 1)
    b = dtf.Trainer(dtf.Net(TwentyClasses()), dtf.SoftmaxCrossEntropyLoss(), dtf.GradientDescentOptimizer(0.1),
                    device="cpu")
    b.build(x, y)
    if tuple(b.store.vars["softmax_linear/weights"].shape) == tuple(a.store.vars["softmax_linear/weights"].shape):
        pytest.skip("the example model has a fixed head")
    with pytest.raises(ValueError, match="softmax_linear/weights.*checkpoint holds shape"):
        dtf.Saver(b.store).restore(prefix)


def test_weight_and_bias_helpers_used_by_plain_torch_code_receive_gradients():
    tr, x, y = scenarios.raw_variables_in_user_code_train("cpu")
    before = {k: v.clone() for k, v in tr.store.state_dict().items()}
    losses = [float(tr.step((x, y))) for _ in range(15)]
    after = tr.store.state_dict()
    for k in ("proj/weights", "proj/biases", "gate/weights"):
        assert float((after[k] - before[k]).abs().max()) > 0, k
    assert losses[-1] < losses[0], losses


def test_init_fn_warm_starts_a_fresh_run_and_is_skipped_on_resume(tmp_path):
    scenarios.warm_start_then_resume(tmp_path, gpu=False)


def test_tfrecord_inputs_may_be_a_file_a_list_a_glob_or_a_directory(tmp_path):
    import pytest
    from distributed_tensorflow_framework_b200.runtime.train import _record_files
    for n in ("b.tfrecord", "a.tfrecord", "c.tfrecords", "part.tfrecord-00001-of-00002", "notes.txt"):
        (tmp_path / n).write_bytes(b"")
    d = str(tmp_path)
    assert _record_files(d + "/a.tfrecord") == [d + "/a.tfrecord"]
    assert _record_files([d + "/a.tfrecord", d + "/b.tfrecord"]) == [d + "/a.tfrecord", d + "/b.tfrecord"]
    assert [f.rsplit("/", 1)[1] for f in _record_files(d + "/*.tfrecord")] == ["a.tfrecord", "b.tfrecord"]
    assert [f.rsplit("/", 1)[1] for f in _record_files(d)] == ["a.tfrecord", "b.tfrecord", "c.tfrecords",
                                                              "part.tfrecord-00001-of-00002"]
    with pytest.raises(ValueError):
        _record_files(d + "/*.nothing")


def test_trainer_accepts_numpy_batches_and_list_pairs():
    import numpy as np
    import torch
    import distributed_tensorflow_framework_b200 as dtf
    from distributed_tensorflow_framework_b200 import models
    rng = np.random.RandomState(0)
    x, y = rng.rand(4, 32, 32, 3), rng.randint(0, 10, size=(4,)).astype(np.int32)   # float64 images, int32 labels
    tr = dtf.Trainer(dtf.Net(models.Cifar10CNN()), dtf.SoftmaxCrossEntropyLoss(), dtf.GradientDescentOptimizer(0.05),
                     device="cpu")
    l0 = float(tr.step([x, y]))            # a list pair, numpy
    l1 = float(tr.step((torch.from_numpy(x).float(), torch.from_numpy(y).long())))
    assert l0 == l0 and l1 == l1 and tr.store.global_step == 2
    two = dtf.Trainer(dtf.Net(models.Cifar10CNN()), dtf.SoftmaxCrossEntropyLoss(), dtf.GradientDescentOptimizer(0.05),
                      device="cpu", gpu_num=2)
    losses = [float(two.step([(x, y), [x, y]])) for _ in range(2)]   # one pair per tower, tuple or list
    assert all(l == l for l in losses) and two.store.global_step == 2


def test_lenet_on_numpy_uint8_batches_learns():
    scenarios.lenet_on_numpy_uint8_batches_learns("cpu")


def test_resnet_on_tiny_batches():
    scenarios.resnet_tiny_batches("cpu")


def test_inspect_checkpoint_tool_lists_tensors_and_events(tmp_path, capsys):
    import os
    import distributed_tensorflow_framework_b200 as dtf
    from distributed_tensorflow_framework_b200.cli import inspect_checkpoint
    md = str(tmp_path / "m")
    os.makedirs(md)
    dtf.run_main(scenarios.experiment(md, False, model="TinyMLP", loader="BlobPlaceholderLoader", epochs=1, samples=64,
                                      batch=16, opt=dtf.AdamOptimizer(0.01)))
    assert inspect_checkpoint.main([md, "--events"]) == 0
    out = capsys.readouterr().out
    assert "model.ckpt-4" in out and "fc1/weights" in out and "fc1/weights/Adam_1" in out and "[16, 64]" in out
    assert "global_step" in out and "int64" in out and "event file" in out
    assert inspect_checkpoint.main([md, "--tensor_name", "global_step"]) == 0
    assert "\n4\n" in capsys.readouterr().out


def test_checkpoint_converts_to_a_torch_dict_and_back(tmp_path):
    import os
    import numpy as np
    import torch
    import distributed_tensorflow_framework_b200 as dtf
    from distributed_tensorflow_framework_b200.cli import convert_checkpoint
    from distributed_tensorflow_framework_b200.io import tensorbundle as tb
    md = str(tmp_path / "m")
    os.makedirs(md)
    op = dtf.run_main(scenarios.experiment(md, False, model="TinyMLP", loader="BlobPlaceholderLoader", epochs=1,
                                           samples=64, batch=16, opt=dtf.MomentumOptimizer(0.05, 0.9)))
    pt = str(tmp_path / "model.pt")
    assert convert_checkpoint.main([md, pt]) == 0
    state = torch.load(pt)
    assert int(state["global_step"]) == 4 and state["global_step"].shape == ()
    assert torch.equal(state["fc1/weights"], op.trainer.store.vars["fc1/weights"].master.detach())
    state["extra/bf16"] = torch.arange(6, dtype=torch.float32).reshape(2, 3).to(torch.bfloat16)
    torch.save(state, pt)
    new_dir = tmp_path / "converted"
    new_dir.mkdir()
    assert convert_checkpoint.main([pt, str(new_dir / "model.ckpt-4")]) == 0
    assert tb.latest_checkpoint(str(new_dir)).endswith("model.ckpt-4")
    # the converted checkpoint resumes a run
    fresh = dtf.Trainer(dtf.Net(scenarios.TinyMLP()), dtf.SoftmaxCrossEntropyLoss(), dtf.MomentumOptimizer(0.05, 0.9),
                        device="cpu")
    fresh.build(torch.zeros(2, 16), torch.zeros(2, dtype=torch.long))
    assert dtf.Saver(fresh.store, fresh.optimizer).restore(tb.latest_checkpoint(str(new_dir))) == 4
    assert torch.equal(fresh.store.vars["logits/weights"].master.detach(), state["logits/weights"])
    with tb.BundleReader(tb.latest_checkpoint(str(new_dir))) as r:
        assert r.dtype_of("extra/bf16") == tb.DT_BFLOAT16
        bits = r.get_tensor("extra/bf16")
        assert np.array_equal((bits.astype(np.uint32) << 16).view(np.float32), np.arange(6, dtype=np.float32).reshape(2, 3))
    back = str(tmp_path / "back.pt")
    assert convert_checkpoint.main([str(new_dir), back]) == 0
    again = torch.load(back)
    assert again["extra/bf16"].dtype == torch.bfloat16 and torch.equal(again["extra/bf16"], state["extra/bf16"])
    assert sorted(again) == sorted(state)
