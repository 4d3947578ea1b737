"""tf.estimator-style use of the framework (/root/reference/distribute_estimator.py:25-35): a `model_fn` that returns an
EstimatorSpec, `train` with periodic TF1-layout checkpoints in model_dir, `evaluate` from the latest checkpoint,
`predict`, and a second `train` call that resumes from the checkpoint.

  python examples/estimator.py                                   # CPU: CIFAR-10 CNN on 32x32 synthetic images
  python examples/estimator.py --model=ResNet18 --lr=0.002       # CPU: ResNet-18
  python examples/estimator.py --model=ResNet50 --image_size=224 --batch=64 --steps=50    # one B200
"""
import os
import sys
import tempfile

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import distributed_tensorflow_framework_b200 as dtf  # noqa: E402
from distributed_tensorflow_framework_b200 import flags, models  # noqa: E402

flags.FLAGS.define("model", str, "Cifar10CNN", "ResNet18 | ResNet50 | Cifar10CNN")
flags.FLAGS.define("model_dir", str, "", "checkpoint directory (default: a temp dir)")
flags.FLAGS.define("image_size", int, 32, "height = width of the synthetic images")
flags.FLAGS.define("classes", int, 10, "number of classes of the synthetic labels")
flags.FLAGS.define("batch", int, 32, "batch size")
flags.FLAGS.define("steps", int, 80, "training steps per train() call")
flags.FLAGS.define("lr", float, 0.05, "momentum-SGD learning rate")
flags.parse()
F = flags.FLAGS


def model_fn(features, labels, mode, params):
    model = getattr(models, params["model"])()
    if hasattr(model, "num_classes"):
        model.num_classes = params["classes"]
    return dtf.EstimatorSpec(net=model, loss=dtf.SoftmaxCrossEntropyLoss(),
                             optimizer=dtf.MomentumOptimizer(params["lr"], 0.9),
                             predictions_fn=lambda logits: logits.argmax(-1))


def input_fn(n_batches, seed):
    """Synthetic batches whose label is visible in the image (class k = brightness level k plus noise)."""
    def make():
        g = torch.Generator().manual_seed(seed)
        for _ in range(n_batches):
            y = torch.randint(0, F.classes, (F.batch,), generator=g)
            x = 0.3 * torch.rand(F.batch, F.image_size, F.image_size, 3, generator=g)
            x = x + 0.7 * (y.float() / max(F.classes - 1, 1)).view(-1, 1, 1, 1)
            yield x, y
    return make


if __name__ == "__main__":
    model_dir = F.model_dir or tempfile.mkdtemp(prefix="dtf_estimator_")
    config = dtf.RunConfig(model_dir=model_dir, save_checkpoints_steps=max(F.steps // 2, 1))
    params = {"model": F.model, "classes": F.classes, "lr": F.lr}
    est = dtf.Estimator(model_fn, config=config, params=params)
    est.train(input_fn(F.steps, seed=0), steps=F.steps)
    print("trained %d steps, last loss %.4f, checkpoints in %s" % (F.steps, est.last_loss, model_dir), flush=True)
    metrics = est.evaluate(input_fn(2, seed=1))
    print("evaluate:", {k: (round(v, 4) if isinstance(v, float) else v) for k, v in metrics.items()}, flush=True)
    preds = next(iter(est.predict(input_fn(1, seed=2))))
    print("predict: %s" % preds.tolist(), flush=True)
    # a fresh Estimator on the same model_dir resumes from the latest checkpoint
    est2 = dtf.Estimator(model_fn, config=config, params=params)
    est2.train(input_fn(F.steps, seed=3), steps=F.steps)
    resumed = est2.evaluate(input_fn(2, seed=1))
    assert resumed["global_step"] == 2 * F.steps, resumed
    print("resumed to step %d" % resumed["global_step"], flush=True)
