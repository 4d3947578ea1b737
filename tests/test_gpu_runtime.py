"""The Train / Eval operators on the CUDA path through the IoC container: kernels + arenas + TF1 checkpoints."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu


def _main(model_dir, mode="Train", epochs=1, samples=256, batch=32, model="Cifar10CNN", loader="Cifar10SyntheticLoader",
          loss="SoftmaxCrossEntropyLoss", opt=None):
    import distributed_tensorflow_framework_b200 as dtf
    from distributed_tensorflow_framework_b200 import models  # noqa: F401
    from distributed_tensorflow_framework_b200.api import annotations as ann

    @ann.current_model(model=model)
    @ann.optimizer(optimizer=opt or dtf.MomentumOptimizer(0.01, 0.9))
    @ann.loss(loss=loss)
    @ann.current_mode(mode=mode)
    @ann.current_input(input=loader)
    @ann.gpu_num(gpu_num=1)
    @ann.ps_hosts(ps_hosts="")
    @ann.worker_hosts(worker_hosts="127.0.0.1:29711")
    @ann.job_name(job_name="worker")
    @ann.task_index(task_index=0)
    @ann.batch_size(batch_size=batch)
    @ann.sample_number(sample_number=samples)
    @ann.epoch_num(epoch_num=epochs)
    @ann.model_dir(model_dir=model_dir)
    @ann.data_dir(data_dir="")
    def main():
        pass

    return main


def test_train_checkpoint_resume_eval_on_cuda(tmp_path):
    import distributed_tensorflow_framework_b200 as dtf
    from distributed_tensorflow_framework_b200.io import tensorbundle as tb
    md = str(tmp_path / "m")
    os.makedirs(md)
    op = dtf.run_main(_main(md, epochs=2))
    assert op.trainer.device.type == "cuda" and op.trainer.engine is not None
    assert op.trainer.store.global_step == 16 and op.final_loss == op.final_loss
    latest = tb.latest_checkpoint(md)
    with tb.BundleReader(latest) as r:
        assert r.get_tensor("conv1/weights").shape == (5, 5, 3, 64)  # logical shape, not the padded store shape
        assert "conv1/weights/Momentum" in r.keys() and int(r.get_tensor("global_step").reshape(-1)[0]) == 16
        w_saved = torch.from_numpy(r.get_tensor("local3/weights"))
    assert torch.allclose(w_saved, op.trainer.store.vars["local3/weights"].master.cpu(), atol=0)
    # resume on CUDA: restored weights land in the arenas (master + bf16 shadows) and training continues
    op2 = dtf.run_main(_main(md, epochs=3))
    assert op2.trainer.store.global_step == 24
    # Eval operator on CUDA restores the same checkpoint
    ev = dtf.run_main(_main(md, mode="Eval", epochs=3))
    assert ev.results["global_step"] == 24 and 0.0 <= ev.results["accuracy"] <= 1.0 and ev.results["loss"] > 0


def test_cpu_checkpoint_loads_into_cuda_arenas_and_matches_forward(tmp_path):
    import distributed_tensorflow_framework_b200 as dtf
    from distributed_tensorflow_framework_b200 import models
    from distributed_tensorflow_framework_b200.runtime.session import Saver
    torch.manual_seed(0)
    x = torch.rand(8, 32, 32, 3)
    cpu = dtf.Trainer(dtf.Net(models.Cifar10CNN()), dtf.SoftmaxCrossEntropyLoss(), dtf.GradientDescentOptimizer(0.1))
    cpu.build(x)
    prefix = Saver(cpu.store).save(str(tmp_path), 7)
    gpu = dtf.Trainer(dtf.Net(models.Cifar10CNN()), dtf.SoftmaxCrossEntropyLoss(), dtf.GradientDescentOptimizer(0.1),
                      device="cuda:0")
    gpu.build(x)
    assert Saver(gpu.store).restore(prefix) == 7
    yc, yg = cpu.forward(x), gpu.forward(x).float().cpu()
    assert (yc - yg).abs().max() < 3e-2 * (yc.abs().max() + 1e-3) + 1e-2
