"""Peer-memory collectives and the Train / Eval operators across real GPUs (one process per GPU via torchrun).
Parametrised over 2 / 4 / 8 ranks; a world size the box does not have is skipped."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _torchrun(world, script, *args, timeout=900, extra=()):
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(_port())] + list(extra) + \
          [os.path.join(HERE, script)] + [str(a) for a in args]
    env = dict(os.environ)
    env.pop("DTF_KILL_RANK_AT_STEP", None)
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=timeout, env=env)
    # keep the ranks' own output (torchrun's failure summary hides it at the end of stdout)
    out_dir = os.path.join(os.path.dirname(HERE), "gpurun_out")
    if os.path.isdir(out_dir):
        tag = "_".join([os.path.splitext(script)[0].strip("_"), str(world)] + [os.path.basename(str(a)) for a in args])
        with open(os.path.join(out_dir, "mr_%s.log" % tag), "a") as f:
            f.write(p.stdout + "\n==== rc %d\n" % p.returncode)
    p.worker_tail = "\n".join(l for l in p.stdout.splitlines() if not l.startswith(("E  ", "Traceback", "  File "))
                              and "torch/distributed" not in l)[-4000:]
    return p


@pytest.mark.parametrize("world", [2, 4, 8])
def test_collectives_and_data_parallel_trainer(world):
    p = _torchrun(world, "_comm_worker.py")
    assert p.returncode == 0 and "COMM_WORKER_OK" in p.stdout, p.worker_tail


@pytest.mark.parametrize("world", [2, 4])
def test_owner_sharded_fp32_backup_async_graph_and_timeouts(world):
    p = _torchrun(world, "_comm_worker2.py")
    assert p.returncode == 0 and "COMM_WORKER2_OK" in p.stdout, p.worker_tail


@pytest.mark.parametrize("sync", ["sync", "backup", "async"])
def test_train_operator_save_resume_eval_multi_rank(tmp_path, sync):
    world = 2
    md = str(tmp_path / sync)
    os.makedirs(md)
    p = _torchrun(world, "_train_worker.py", md, sync, 2)
    assert p.returncode == 0 and "TRAIN_WORKER_OK 20" in p.stdout, p.worker_tail
    assert "Saving checkpoints for 5 " in p.stdout or "Saving checkpoints for 5" in p.stdout, p.worker_tail
    # resume: a second launch restores step 20 on every rank and continues to 30
    p = _torchrun(world, "_train_worker.py", md, sync, 3)
    assert p.returncode == 0 and "TRAIN_WORKER_OK 30" in p.stdout, p.worker_tail
    assert "Restoring parameters from" in p.stdout


def test_kill_one_rank_and_resume_from_checkpoint(tmp_path):
    """DTF_KILL_RANK_AT_STEP kills the last rank at step 13 of the first attempt. The survivor's collective runs into
    its time bound (error record, WorkerLostError) instead of hanging; torchrun restarts the group, which restores
    the step-10 checkpoint and finishes with global_step continuing to 20."""
    md = str(tmp_path / "kill")
    os.makedirs(md)
    p = _torchrun(2, "_train_worker.py", md, "sync", 2, "kill", extra=("--max-restarts", "1"), timeout=900)
    assert p.returncode == 0 and "TRAIN_WORKER_OK 20 restart=1" in p.stdout, p.worker_tail
    assert "Restoring parameters from" in p.stdout and "model.ckpt-10" in p.stdout


def test_gpu_num_two_without_a_launcher_spawns_one_rank_per_gpu(tmp_path):
    """`@gpu_num(2)` and no torchrun: the container starts one process per GPU; both train (global batch =
    batch_size x gpu_num as in the reference), the chief writes the checkpoint."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    md = str(tmp_path / "spawn")
    os.makedirs(md)
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "TORCHELASTIC_RUN_ID",
                                                            "MASTER_ADDR", "MASTER_PORT", "DTF_SPAWNED")}
    p = subprocess.run([sys.executable, os.path.join(HERE, "_spawn_main.py"), md], stdout=subprocess.PIPE,
                       stderr=subprocess.STDOUT, text=True, timeout=600, env=env)
    out = p.stdout
    assert p.returncode == 0 and "SPAWN_PARENT_OK" in out, out[-3000:]
    assert "SPAWN_RANK_OK rank=0 world=2 device=cuda:0 step=10 towers=1" in out, out[-3000:]
    assert "SPAWN_RANK_OK rank=1 world=2 device=cuda:1 step=10 towers=1" in out, out[-3000:]
    from distributed_tensorflow_framework_b200.io import tensorbundle as tb
    assert tb.latest_checkpoint(md).endswith("model.ckpt-10")
