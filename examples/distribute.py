"""Runnable equivalent of the reference entry script (/root/reference/distribute.py): the whole experiment is the
decorator stack on `main`.

  single process, CPU (BASELINE.json config 1):
      python examples/distribute.py --job_name=worker --task_index=0
  parameter server + two workers on loopback (three shells):
      python examples/distribute.py --job_name=ps     --task_index=0 --ps_hosts=127.0.0.1:23456 --worker_hosts=127.0.0.1:23457,127.0.0.1:23458
      python examples/distribute.py --job_name=worker --task_index=0 --ps_hosts=... --worker_hosts=...
      python examples/distribute.py --job_name=worker --task_index=1 --ps_hosts=... --worker_hosts=...
  8 GPUs, one process per GPU:
      torchrun --standalone --nproc-per-node 8 examples/distribute.py --gpu_num_per_worker=1
"""
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import distributed_tensorflow_framework_b200 as dtf  # noqa: E402
from distributed_tensorflow_framework_b200 import annotations, flags  # noqa: E402
from distributed_tensorflow_framework_b200 import models  # noqa: E402,F401  (registers the example plugins)

flags.FLAGS.define("gpu_num_per_worker", int, 0, "towers per worker process (0 = CPU)")
flags.FLAGS.define("model_dir", str, "", "checkpoint directory (default: a temp dir)")
flags.FLAGS.define("epochs", int, 1, "epochs")
flags.FLAGS.define("samples", int, 640, "samples per epoch")
flags.parse()
_model_dir = flags.FLAGS.model_dir or tempfile.mkdtemp(prefix="dtf_model_")


@annotations.current_model(model="Cifar10CNN")
@annotations.optimizer(optimizer=dtf.AdamOptimizer(0.001))
@annotations.loss(loss="SoftmaxCrossEntropyLoss")
@annotations.current_mode(mode="Train")
@annotations.current_input(input="Cifar10SyntheticLoader")
@annotations.gpu_num(gpu_num=flags.FLAGS.gpu_num_per_worker)
@annotations.ps_hosts(ps_hosts=flags.FLAGS.ps_hosts)
@annotations.worker_hosts(worker_hosts=flags.FLAGS.worker_hosts or "127.0.0.1:29511")
@annotations.job_name(job_name=flags.FLAGS.job_name or "worker")
@annotations.task_index(task_index=flags.FLAGS.task_index or 0)
@annotations.batch_size(batch_size=32)
@annotations.sample_number(sample_number=flags.FLAGS.samples)
@annotations.epoch_num(epoch_num=flags.FLAGS.epochs)
@annotations.model_dir(model_dir=_model_dir)
@annotations.data_dir(data_dir="")
def main(argv=None):
    return dtf.run_main(main)


if __name__ == "__main__":
    main(sys.argv)
