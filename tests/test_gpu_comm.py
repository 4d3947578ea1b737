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


def test_ps_and_two_gpu_workers_launched_by_hand(tmp_path):
    """The reference's own launch mode on real GPUs: three processes started with --job_name / --task_index (no
    launcher): a ps task and two workers, one GPU each. The declared ps makes the workers train in parameter-server
    mode (owner-sharded masters, push = reduce-scatter + owner update); both end with identical replicas, the chief
    writes one data shard per owner, the ps leaves through the done queue."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    ports = [_port() for _ in range(3)]
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
@ann.optimizer(optimizer=dtf.MomentumOptimizer(0.02, 0.9))
@ann.loss(loss="SoftmaxCrossEntropyLoss")
@ann.current_mode(mode="Train")
@ann.current_input(input="Cifar10SyntheticLoader")
@ann.gpu_num(gpu_num=0 if flags.FLAGS.job_name == "ps" else 1)
@ann.ps_hosts(ps_hosts="127.0.0.1:%d")
@ann.worker_hosts(worker_hosts="127.0.0.1:%d, 127.0.0.1:%d")
@ann.job_name(job_name=flags.FLAGS.job_name)
@ann.task_index(task_index=flags.FLAGS.task_index)
@ann.batch_size(batch_size=32)
@ann.sample_number(sample_number=640)
@ann.epoch_num(epoch_num=1)
@ann.model_dir(model_dir=md)
@ann.data_dir(data_dir="")
def main(): pass
op = dtf.run_main(main)
if flags.FLAGS.job_name == "worker":
    tr = op.trainer
    full = tr.engine.gather_masters(slots=False) if tr.engine.needs_gather() else None
    w = tr.store.vars["local4/weights"].master
    print("RESULT", flags.FLAGS.task_index, tr.store.global_step, tr.device, tr.algo,
          "%%.6f" %% float(w.double().sum()), "%%.4f" %% op.final_loss, flush=True)
else:
    print("PS_EXIT", flush=True)
''' % (os.path.dirname(HERE), str(tmp_path), ports[0], ports[1], ports[2]))
    import time
    logs = [tmp_path / ("out_%d.log" % i) for i in range(3)]
    files = [open(l, "w") for l in logs]
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "TORCHELASTIC_RUN_ID")}
    procs = [subprocess.Popen([sys.executable, str(script), "--job_name=%s" % j, "--task_index=%d" % t],
                              stdout=f, stderr=subprocess.STDOUT, text=True, env=env)
             for f, (j, t) in zip(files, (("ps", 0), ("worker", 0), ("worker", 1)))]
    try:
        deadline = time.time() + 240
        while time.time() < deadline and any(p.poll() is None for p in procs):
            time.sleep(0.2)
    finally:
        for p in procs:
            if p.poll() is None:
                p.kill()
                p.wait()
        for f in files:
            f.close()
    outs = [l.read_text() for l in logs]
    out_dir = os.path.join(os.path.dirname(HERE), "gpurun_out")
    if os.path.isdir(out_dir):
        with open(os.path.join(out_dir, "mr_ps_two_gpu_workers.log"), "w") as f:
            f.write("\n==== next process\n".join(outs))
    assert all(p.returncode == 0 for p in procs), "\n".join(o[-1500:] for o in outs)
    assert "PS_EXIT" in outs[0]
    r0 = [l for l in outs[1].splitlines() if l.startswith("RESULT")][0].split()
    r1 = [l for l in outs[2].splitlines() if l.startswith("RESULT")][0].split()
    assert r0[2] == r1[2] == "20" and r0[3] == "cuda:0" and r1[3] == "cuda:1", (r0, r1)
    assert r0[4] == r1[4] == "twoshot", (r0, r1)  # declared ps tasks select the parameter-server collectives
    assert r0[5] == r1[5], (r0, r1)  # identical replicas
    files_m = sorted(os.listdir(tmp_path / "m"))
    assert any(f.startswith("model.ckpt-20.data-00000-of-00002") for f in files_m), files_m
    assert any(f.startswith("model.ckpt-20.data-00001-of-00002") for f in files_m), files_m
