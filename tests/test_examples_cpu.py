"""The scripts under examples/ run end to end on CPU (each in its own process, like a user would start them)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, timeout=300):
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "TORCHELASTIC_RUN_ID")}
    env["CUDA_VISIBLE_DEVICES"] = ""
    p = subprocess.run([sys.executable] + args, cwd=ROOT, env=env, capture_output=True, text=True, timeout=timeout)
    assert p.returncode == 0, (p.stdout + p.stderr)[-3000:]
    return p.stdout + p.stderr


def test_entry_script_example(tmp_path):
    out = _run(["examples/distribute.py", "--job_name=worker", "--task_index=0", "--samples=128",
                "--model_dir=%s" % tmp_path])
    assert "Saving checkpoints for 4" in out
    assert any(f.startswith("model.ckpt-4") for f in os.listdir(tmp_path))


def test_dehaze_tfrecord_example_trains_and_evaluates(tmp_path):
    model_dir, rec_dir = tmp_path / "m", tmp_path / "r"
    model_dir.mkdir()
    rec_dir.mkdir()
    out = _run(["examples/dehaze_tfrecord.py", "--image_size=16", "--samples=32", "--eval_samples=8", "--epochs=1",
                "--batch=4", "--model_dir=%s" % model_dir, "--records_dir=%s" % rec_dir])
    assert "EVAL step 8:" in out and "over 2 batches" in out
    assert sorted(os.listdir(rec_dir)) == ["eval.tfrecord", "train.tfrecord"]
    assert (model_dir / "lr.json").exists() and (model_dir / "checkpoint").exists()
    assert (model_dir / "eval").is_dir()  # evaluation scalars for TensorBoard


def test_estimator_example_trains_resumes_and_predicts(tmp_path):
    out = _run(["examples/estimator.py", "--steps=6", "--batch=8", "--model_dir=%s" % tmp_path])
    assert "trained 6 steps" in out and "resumed to step 12" in out and "predict: [" in out


@pytest.mark.timeout(400)
def test_dehaze_example_under_a_two_rank_launcher(tmp_path):
    """torchrun, gloo, two ranks: both train (one shared TFRecord file), rank 0 writes the checkpoint, both evaluate
    their share of the batches and agree on the result."""
    model_dir, rec_dir = tmp_path / "m", tmp_path / "r"
    model_dir.mkdir()
    rec_dir.mkdir()
    out = _run(["-m", "torch.distributed.run", "--standalone", "--local-addr", "127.0.0.1", "--nproc-per-node", "2",
                "examples/dehaze_tfrecord.py", "--image_size=16", "--samples=32", "--eval_samples=8", "--epochs=1",
                "--batch=4", "--model_dir=%s" % model_dir, "--records_dir=%s" % rec_dir], timeout=380)
    assert out.count("eval @ step 8: batches = 2") == 2 and "EVAL step 8:" in out
    assert "freshly initialised" not in out  # no rank evaluates before the chief's final checkpoint exists


@pytest.mark.timeout(300)
def test_local_cluster_script_runs_a_ps_and_two_workers(tmp_path):
    import socket
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        base = s.getsockname()[1]
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "TORCHELASTIC_RUN_ID")}
    env.update(CUDA_VISIBLE_DEVICES="", DTF_BASE_PORT=str(base), DTF_MODEL_DIR=str(tmp_path))
    p = subprocess.run(["bash", "examples/launch_local_cluster.sh", "2", "--samples=128"], cwd=ROOT, env=env,
                       capture_output=True, text=True, timeout=280)
    out = p.stdout + p.stderr
    assert p.returncode == 0 and "cluster finished (exit 0)" in out, out[-3000:]
    assert any(f.startswith("model.ckpt-4") for f in os.listdir(tmp_path))
