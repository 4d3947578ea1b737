"""Image-to-image experiment over TFRecord files: the workload the reference's entry script describes
(/root/reference/distribute.py:30-35: features `hazed_image_raw` / `clear_image_raw`, a GMAN-style dehazing net, MSE
loss) with a user-written TFRecord loader (the `_decode_raw_data` hook of /root/reference/distribute_input.py:71-106),
advice hooks, a learning-rate handler, then an evaluation pass over the checkpoint in the same script.

  python examples/dehaze_tfrecord.py                       # CPU, tiny images, ~1 minute
  python examples/dehaze_tfrecord.py --gpu_num_per_worker=1 --image_size=256 --samples=256
  torchrun --standalone --nproc-per-node 8 examples/dehaze_tfrecord.py --gpu_num_per_worker=1 --image_size=256

There is no dataset on this machine: the script first writes synthetic (hazy, clear) pairs with the atmospheric
scattering model I = J t + A (1 - t) into TFRecord files under --data_dir (default: a temporary directory).
"""
import os
import sys
import tempfile

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import distributed_tensorflow_framework_b200 as dtf  # noqa: E402
from distributed_tensorflow_framework_b200 import annotations, flags, models  # noqa: E402,F401
from distributed_tensorflow_framework_b200.io import tfrecord as tfr  # noqa: E402
from distributed_tensorflow_framework_b200.plugins import input as Input  # noqa: E402
from distributed_tensorflow_framework_b200.plugins.learningrate import LearningRate  # noqa: E402
from distributed_tensorflow_framework_b200.runtime import container  # noqa: E402

flags.FLAGS.define("gpu_num_per_worker", int, 0, "towers per worker process (0 = CPU)")
flags.FLAGS.define("model_dir", str, "", "checkpoint directory (default: a temp dir)")
flags.FLAGS.define("records_dir", str, "", "where the TFRecord files live / are written (default: a temp dir)")
flags.FLAGS.define("image_size", int, 32, "height = width of the synthetic images")
flags.FLAGS.define("samples", int, 64, "training pairs")
flags.FLAGS.define("eval_samples", int, 16, "evaluation pairs")
flags.FLAGS.define("epochs", int, 2, "epochs")
flags.FLAGS.define("batch", int, 8, "batch size per tower")
flags.parse()
F = flags.FLAGS
F.input_image_height = F.input_image_width = F.image_size
_rank = int(os.environ.get("RANK", "0"))


def _shared_dir(flag_value, env_key, kind):
    """All ranks of one job must agree on the directory: the flag, else what the first process exported, else (under
    a launcher) a name derived from the job's rendezvous id, else a fresh temporary directory."""
    if flag_value or os.environ.get(env_key):
        return flag_value or os.environ[env_key]
    if "WORLD_SIZE" in os.environ:
        job = os.environ.get("TORCHELASTIC_RUN_ID", "") + "_" + os.environ.get("MASTER_PORT", "0")
        path = os.path.join(tempfile.gettempdir(), "dtf_dehaze_%s_%s" % (kind, job))
        os.makedirs(path, exist_ok=True)
        return path
    return tempfile.mkdtemp(prefix="dtf_dehaze_%s_" % kind)


_model_dir = _shared_dir(F.model_dir, "DTF_EXAMPLE_MODEL_DIR", "model")
_records_dir = _shared_dir(F.records_dir, "DTF_EXAMPLE_RECORDS_DIR", "rec")
FEATURES = {"hazed_image_raw": tfr.FixedLenFeature([], "string"),
            "clear_image_raw": tfr.FixedLenFeature([], "string")}


def write_records(path, n, size, seed):
    """n synthetic pairs: smooth random scenes J, transmission t falling off along a random direction, airlight A."""
    rng = np.random.RandomState(seed)
    yy, xx = np.mgrid[0:size, 0:size].astype(np.float32) / max(size - 1, 1)
    with tfr.TFRecordWriter(path) as w:
        for _ in range(n):
            coeff = rng.rand(3, 3).astype(np.float32)
            clear = np.stack([np.clip(c[0] * xx + c[1] * yy + c[2] * np.sin(6.0 * xx * yy), 0, 1) for c in coeff], -1)
            ang = rng.rand() * 2 * np.pi
            depth = np.clip(0.5 + 0.5 * (np.cos(ang) * (xx - 0.5) + np.sin(ang) * (yy - 0.5)) * 2, 0, 1)
            t = np.exp(-(0.5 + rng.rand()) * depth)[..., None]
            hazy = clear * t + (0.7 + 0.3 * rng.rand()) * (1.0 - t)
            w.write(tfr.make_example({"hazed_image_raw": (hazy * 255).astype(np.uint8).tobytes(),
                                      "clear_image_raw": (clear * 255).astype(np.uint8).tobytes()}))


class DehazeRecordLoader(Input.TFRecordDataLoader):
    """The user-side loader: one decode hook for training; evaluation streams the held-out file once."""

    def _decode_raw_data(self, raw_features, height, width, *args, **kwargs):
        hazy = np.frombuffer(raw_features["hazed_image_raw"], dtype=np.uint8).reshape(height, width, 3)
        clear = np.frombuffer(raw_features["clear_image_raw"], dtype=np.uint8).reshape(height, width, 3)
        return [hazy.astype(np.float32) / 255.0, clear.astype(np.float32) / 255.0]

    def load_eval_batch(self, *args, **kwargs):
        import torch
        if getattr(self, "_eval_it", None) is None:
            self._eval_it = tfr.tf_record_iterator(os.path.join(_records_dir, "eval.tfrecord"))
        raws, gts = [], []
        for rec in self._eval_it:
            h, c = self._decode_raw_data(tfr.parse_single_example(rec, FEATURES), F.image_size, F.image_size)
            raws.append(torch.from_numpy(h))
            gts.append(torch.from_numpy(c))
            if len(raws) == self.batch_size:
                break
        if not raws:
            self._eval_it = None
            return None
        return torch.stack(raws), torch.stack(gts)


def announce(args, kwargs):
    dtf.log.info("dehaze example: records in %s, checkpoints in %s", _records_dir, _model_dir)


def report(results, args, kwargs):
    if _rank != 0:
        return
    mse = results["loss"]
    psnr = 10.0 * np.log10(1.0 / max(mse, 1e-12))
    print("EVAL step %d: mse %.5f  psnr %.2f dB over %d batches" % (results["global_step"], mse, psnr,
                                                                   results["batches"]), flush=True)


def experiment(mode):
    """The decorator stack of the reference's entry script, once per mode."""

    @annotations.current_model(model="DehazeFCN")
    @annotations.optimizer(optimizer=dtf.AdamOptimizer(0.001))
    @annotations.loss(loss="MSELoss")
    @annotations.current_mode(mode=mode)
    @annotations.current_input(input="DehazeRecordLoader")
    @annotations.current_feature(features=FEATURES)
    @annotations.get_advice(pre_fn=announce, post_eval_fn=report)
    @annotations.parse_data_dir(parse_data_dir_fn=lambda d: os.path.join(d, "train.tfrecord"))
    @annotations.learning_rate(learning_rate=LearningRate(0.001, os.path.join(_model_dir, "lr.json"),
                                                          decay_factor=0.5))
    @annotations.gpu_num(gpu_num=F.gpu_num_per_worker)
    @annotations.ps_hosts(ps_hosts=F.ps_hosts)
    @annotations.worker_hosts(worker_hosts=F.worker_hosts or "127.0.0.1:29521")
    @annotations.job_name(job_name=F.job_name or "worker")
    @annotations.task_index(task_index=F.task_index or 0)
    @annotations.batch_size(batch_size=F.batch)
    @annotations.sample_number(sample_number=F.samples)
    @annotations.epoch_num(epoch_num=F.epochs)
    @annotations.model_dir(model_dir=_model_dir)
    @annotations.data_dir(data_dir=_records_dir)
    def main(argv=None):
        pass

    return main


if __name__ == "__main__":
    # spawned towers / torchrun ranks must share the directories the first process picked
    os.environ.setdefault("DTF_EXAMPLE_MODEL_DIR", _model_dir)
    os.environ.setdefault("DTF_EXAMPLE_RECORDS_DIR", _records_dir)
    train_file = os.path.join(_records_dir, "train.tfrecord")
    if _rank == 0 and not os.environ.get("DTF_SPAWNED") and not os.path.exists(train_file):
        write_records(train_file + ".tmp", F.samples, F.image_size, seed=1)
        write_records(os.path.join(_records_dir, "eval.tfrecord"), F.eval_samples, F.image_size, seed=2)
        os.replace(train_file + ".tmp", train_file)
    while not os.path.exists(train_file):  # other ranks of a launcher wait for rank 0's files
        import time
        time.sleep(0.1)
    dtf.run_main(experiment("Train"))
    if not os.environ.get("DTF_SPAWNED"):  # under a launcher every rank evaluates its share of the batches
        op = container.build_operator(experiment("Eval"))
        res = op.run()
        assert res["global_step"] > 0 and res["loss"] == res["loss"]
