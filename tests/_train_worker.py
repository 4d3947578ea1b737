"""Multi-rank worker for tests/test_gpu_comm.py: the Train / Eval OPERATORS through the IoC container under torchrun
(one rank per GPU). argv: <model_dir> <mode: sync|backup|async> <epochs> [kill]. Prints TRAIN_WORKER_OK <step> on
rank 0. With `kill`, DTF_KILL_RANK_AT_STEP makes one rank die in the first attempt; torchrun restarts the group, which
must resume from the last checkpoint."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import distributed_tensorflow_framework_b200 as dtf
from distributed_tensorflow_framework_b200 import models  # noqa: F401
from distributed_tensorflow_framework_b200.api import annotations as ann
from distributed_tensorflow_framework_b200.api import flags
from distributed_tensorflow_framework_b200.io import tensorbundle as tb


def build_main(model_dir, mode, epochs, sync):
    @ann.current_model(model="Cifar10CNN")
    @ann.optimizer(optimizer=dtf.MomentumOptimizer(0.01, 0.9))
    @ann.loss(loss="SoftmaxCrossEntropyLoss")
    @ann.current_mode(mode=mode)
    @ann.current_input(input="Cifar10SyntheticLoader")
    @ann.gpu_num(gpu_num=1)
    @ann.batch_size(batch_size=32)
    @ann.sample_number(sample_number=320)
    @ann.epoch_num(epoch_num=epochs)
    @ann.model_dir(model_dir=model_dir)
    @ann.data_dir(data_dir="")
    @ann.sync_mode(sync_mode=sync)
    def main():
        pass
    return main


def main():
    model_dir, sync, epochs = sys.argv[1], sys.argv[2], int(sys.argv[3])
    kill = len(sys.argv) > 4 and sys.argv[4] == "kill"
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    restart = int(os.environ.get("TORCHELASTIC_RESTART_COUNT", "0"))
    if kill and restart == 0:
        os.environ["DTF_KILL_RANK_AT_STEP"] = "%d:13" % (world - 1)
        os.environ["DTF_COMM_TIMEOUT_MS"] = "3000"
    else:
        os.environ.pop("DTF_KILL_RANK_AT_STEP", None)
    flags.FLAGS.reset()
    flags.FLAGS.save_checkpoint_steps = 5
    flags.FLAGS.use_cuda_graph = True
    if sync != "async":
        flags.FLAGS.comm_algo = "twoshot"  # owner-sharded masters: the checkpoint needs the collective gather
    if sync == "backup":
        flags.FLAGS.replicas_to_aggregate = world - 1
    op = dtf.run_main(build_main(model_dir, "Train", epochs, sync))
    total = epochs * 10
    step = op.trainer.store.global_step
    assert step == total, (step, total)
    eng = op.trainer.engine
    assert eng.world == world and eng.needs_gather()
    if sync == "backup":
        assert eng.backup_r == world - 1
    if sync == "async":
        assert eng.algo == "async"
    else:
        assert eng.replicas_equal(), "replicas diverged"
    torch.cuda.synchronize()
    import torch.distributed as dist
    dist.barrier()
    if rank == 0:
        prefix = tb.latest_checkpoint(model_dir)
        assert prefix is not None and prefix.endswith("model.ckpt-%d" % total), prefix
        shards = sorted(f for f in os.listdir(model_dir) if f.startswith(os.path.basename(prefix) + ".data-"))
        assert len(shards) == world, shards  # one data shard per owner rank
        with tb.BundleReader(prefix) as r:
            keys = set(r.keys())
            assert {"conv1/weights", "local4/biases", "softmax_linear/weights/Momentum", "global_step"} <= keys
            # the gathered masters: every variable (owned by any rank) equals what this rank now holds
            st = op.trainer.store
            for name in ("conv1/weights", "local3/weights", "softmax_linear/biases"):
                v = st.vars[name]
                saved = torch.from_numpy(r.get_tensor(name))
                assert torch.equal(saved, v.logical(v.master.detach()).cpu()), name
                if sync != "async":  # the bf16 weights every rank computes with are the rounded masters
                    assert torch.equal(saved.to(torch.bfloat16), v.logical(v.shadow.detach()).cpu()), name
    dist.barrier()
    # the Eval operator on every rank: each evaluates its share of the batches, the sums are combined
    ev = dtf.run_main(build_main(model_dir, "Eval", epochs, sync))
    assert ev.results["global_step"] == total and ev.results["loss"] > 0 and ev.results["batches"] >= world
    if rank == 0:
        print("TRAIN_WORKER_OK %d restart=%d loss=%.4f eval_batches=%d" % (step, restart, op.final_loss,
                                                                          ev.results["batches"]), flush=True)
    dist.barrier()


if __name__ == "__main__":
    main()
