"""User-level scenarios on the CUDA path (tests/scenarios.py): batch-norm ResNets through Train -> checkpoint -> Eval ->
CPU restore, the placeholder / data-path loader kinds, Adam with a decaying learning-rate handler under CUDA-graph
replay, and the example scripts as a user starts them."""
import os
import subprocess
import sys

import pytest

import scenarios

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("model", ["TinyResNet18", "TinyResNet50"])
def test_resnet_train_eval_and_cross_device_restore(tmp_path, model):
    scenarios.resnet_train_eval_and_cross_device_restore(tmp_path, gpu=True, model=model)


def test_placeholder_and_datapath_loaders_learn(tmp_path):
    scenarios.loader_kinds_learn_the_toy_problem(tmp_path, gpu=True)


def test_adam_follows_a_decaying_learning_rate_handler(tmp_path):
    scenarios.adam_with_decaying_learning_rate(tmp_path, gpu=True)


def _run(args, timeout=240):
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "TORCHELASTIC_RUN_ID")}
    p = subprocess.run([sys.executable] + args, cwd=ROOT, env=env, capture_output=True, text=True, timeout=timeout)
    assert p.returncode == 0, (p.stdout + p.stderr)[-3000:]
    return p.stdout + p.stderr


def test_dehaze_tfrecord_example_on_cuda(tmp_path):
    md, rd = tmp_path / "m", tmp_path / "r"
    md.mkdir()
    rd.mkdir()
    out = _run(["examples/dehaze_tfrecord.py", "--gpu_num_per_worker=1", "--image_size=64", "--samples=128",
                "--epochs=3", "--batch=8", "--model_dir=%s" % md, "--records_dir=%s" % rd])
    line = [l for l in out.splitlines() if l.startswith("EVAL step 48:")]
    assert line, out[-2000:]
    mse = float(line[0].split("mse")[1].split()[0])
    assert mse < 0.12, line  # an untrained net sits near 0.35 on these pairs


def test_estimator_example_on_cuda(tmp_path):
    out = _run(["examples/estimator.py", "--model=ResNet18", "--image_size=64", "--steps=30", "--lr=0.01",
                "--model_dir=%s" % tmp_path])
    assert "trained 30 steps" in out and "resumed to step 60" in out and "predict: [" in out


def test_layer_library_on_awkward_shapes_matches_the_cpu_oracle():
    scenarios.kitchen_sink_parity()


@pytest.mark.parametrize("geom", [(4, 9, 13, 20), (3, 8, 8, 3), (2, 7, 7, 36)])
def test_every_layer_with_channel_counts_that_are_not_multiples_of_8(geom):
    failures = []
    for name in scenarios.LAYER_CASES:
        try:
            scenarios.layer_case_parity(name, *geom)
        except Exception as e:  # collect all of them: one line per failing layer
            failures.append("%s: %s: %s" % (name, type(e).__name__, str(e)[:160]))
    assert not failures, "\n".join(failures)


@pytest.mark.parametrize("graph", [False, True])
def test_two_in_process_towers_equal_one_tower_on_the_concatenated_batch(graph):
    scenarios.in_process_towers_equal_one_big_batch("cuda:0", use_graph=graph)


def test_user_torch_ops_custom_loss_and_advice_match_the_cpu_path():
    scenarios.user_torch_ops_loss_and_advice_parity()


@pytest.mark.parametrize("opt", ["momentum", "adam"])
def test_graph_replay_survives_a_batch_of_another_size(opt):
    import distributed_tensorflow_framework_b200 as dtf
    # Adam moves every weight by ~lr per step whatever the size of its gradient, so the run-to-run noise of the
    # filter-gradient reductions (atomic adds) shows up as a few lr of absolute difference after seven steps
    if opt == "momentum":
        scenarios.graph_replay_with_a_smaller_batch_in_between(lambda: dtf.MomentumOptimizer(0.02, 0.9))
    else:
        scenarios.graph_replay_with_a_smaller_batch_in_between(lambda: dtf.AdamOptimizer(1e-3), abs_tol=3e-3)


def test_weight_and_bias_helpers_used_by_plain_torch_code_match_the_cpu_path():
    scenarios.raw_variables_parity()
    tr, x, y = scenarios.raw_variables_in_user_code_train("cuda:0")
    losses = [float(tr.step((x, y))) for _ in range(15)]
    assert losses[-1] < losses[0], losses


def test_init_fn_warm_starts_a_fresh_run_and_is_skipped_on_resume(tmp_path):
    scenarios.warm_start_then_resume(tmp_path, gpu=True)


@pytest.mark.parametrize("graph", [False, True])
def test_lenet_on_numpy_uint8_batches_learns(graph):
    scenarios.lenet_on_numpy_uint8_batches_learns("cuda:0", use_graph=graph)


def test_resnet_on_tiny_batches():
    scenarios.resnet_tiny_batches("cuda:0")
