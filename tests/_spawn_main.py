"""@gpu_num(2) WITHOUT a launcher (tests/test_gpu_comm.py): run_main spawns one process per GPU (the reference's
in-graph towers, /root/reference/distribute_train.py:116-139). argv: <model_dir>."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import distributed_tensorflow_framework_b200 as dtf
from distributed_tensorflow_framework_b200 import models  # noqa: F401
from distributed_tensorflow_framework_b200.api import annotations as ann
from distributed_tensorflow_framework_b200.api import flags

MODEL_DIR = sys.argv[1]


@ann.current_model(model="Cifar10CNN")
@ann.optimizer(optimizer=dtf.MomentumOptimizer(0.01, 0.9))
@ann.loss(loss="SoftmaxCrossEntropyLoss")
@ann.current_mode(mode="Train")
@ann.current_input(input="Cifar10SyntheticLoader")
@ann.gpu_num(gpu_num=2)
@ann.batch_size(batch_size=32)
@ann.sample_number(sample_number=320)
@ann.epoch_num(epoch_num=1)
@ann.model_dir(model_dir=MODEL_DIR)
@ann.data_dir(data_dir="")
def main(argv=None):
    pass


if __name__ == "__main__":
    flags.FLAGS.reset()
    flags.FLAGS.save_checkpoint_steps = 5
    op = dtf.run_main(main)
    if op is not None:  # a spawned tower
        import torch
        eng = op.trainer.engine
        print("SPAWN_RANK_OK rank=%s world=%d device=%s step=%d towers=%d" % (
            os.environ.get("RANK"), eng.world, op.trainer.device, op.trainer.store.global_step, op.trainer.num_towers),
            flush=True)
    else:
        print("SPAWN_PARENT_OK", flush=True)
