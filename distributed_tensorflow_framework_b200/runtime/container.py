"""IoC container: turns the annotation stack on `main` into a wired operator and runs it.

Parity: /root/reference/distribute.py:46-132 -- read the annotations, build ClusterSpec + Server, instantiate loss /
loader / model by reflection (`__new__`, no `__init__`), pick the operator class from the mode string, inject ~15
attributes with setattr, call `operator.run()`. Quirk handling: host strings stripped (Q4), ps hides GPUs
(distribute.py:61-62), model_dir must exist (:74-75), loader kind dispatched on the class attribute `type` (:87-97),
ps does not get model / net (:122-125). When launched by torchrun the cluster is derived from the environment.
"""
import os
import sys

from ..api import annotations
from ..api import flags
from ..plugins import input as Input
from ..plugins import loss as Loss
from ..plugins import model as model_mod
from ..plugins import net as net_mod
from . import evalop as Eval
from . import train as Train
from .cluster import ClusterSpec, Server, cluster_from_env

_OPERATOR_MODULES = {"Train": Train, "Eval": Eval}


def default_optimizer():
    from ..engine.optimizers import AdamOptimizer
    return AdamOptimizer(0.001)


def build_operator(main, model_module=None, loss_module=None, input_module=None, operator_modules=None):
    """Everything of distribute.py:main up to (not including) operator.run(). Returns the wired operator."""
    os.environ.setdefault("TF_SYNC_ON_FINISH", "0")  # kept for parity (distribute.py:48-49); unused here
    os.environ.setdefault("TF_ENABLE_WINOGRAD_NONFUSED", "1")
    g = annotations.get_value_from_annotation

    # Step 1: distributed information
    job_name = getattr(main, "job_name", None) or flags.FLAGS.job_name or "worker"
    task_index = getattr(main, "task_index", None)
    if task_index is None:
        task_index = flags.FLAGS.task_index or 0
    if "TORCHELASTIC_RUN_ID" in os.environ or (not hasattr(main, "worker_hosts") and "WORLD_SIZE" in os.environ):
        cluster, job_name, task_index = cluster_from_env()
    else:
        ps_hosts = getattr(main, "ps_hosts", "") or flags.FLAGS.ps_hosts
        worker_hosts = getattr(main, "worker_hosts", "") or flags.FLAGS.worker_hosts or "127.0.0.1:29511"
        cluster = ClusterSpec({"ps": ps_hosts, "worker": worker_hosts})
    if job_name == "ps":
        os.environ["CUDA_VISIBLE_DEVICES"] = "-1"

    # Step 2: operator parameters
    optimizer = getattr(main, "optimizer", None)
    mode = g(main, "mode")
    if mode != "Train" and mode != "Eval":
        raise ValueError("mode must be set in the annotation @current_mode")
    batch_size = g(main, "batch_size")
    epoch_num = g(main, "epoch_num")
    sample_number = g(main, "sample_number")
    data_dir = g(main, "data_dir")
    model_dir = g(main, "model_dir")
    if not os.path.exists(model_dir):
        raise ValueError("Path to save or restore model doesn't exist")
    loss = annotations.get_instance_from_annotation(main, "loss", loss_module or Loss)
    cluster.num_tasks("worker")
    server = Server(cluster, job_name=job_name, task_index=task_index)
    gpu_num = g(main, "gpu_num")

    # Step 3: data loader
    data_loader = annotations.get_instance_from_annotation(main, "input", input_module or Input)
    input_mode = None
    if data_loader.type == "TFRecordDataLoader":
        if not hasattr(main, "features"):
            raise ValueError("Please use @current_feature to create your features for data_loader")
        setattr(data_loader, "features", g(main, "features"))
        input_mode = Input.InputOptions.TF_RECORD
    elif data_loader.type == "PlaceholderDataLoader":
        input_mode = Input.InputOptions.PLACEHOLDER
    elif data_loader.type == "DataPathDataLoader":
        input_mode = Input.InputOptions.DATAPATHLOADER
    assert input_mode is not None, "Must assign a valid input_mode by using @current_input!"
    setattr(data_loader, "batch_size", batch_size)
    setattr(data_loader, "sample_number", sample_number)
    setattr(data_loader, "data_dir", data_dir)
    setattr(data_loader, "gpu_num", gpu_num)

    # Step 4: operator instance (module and class share the mode's name; user modules in __main__ win)
    mods = dict(_OPERATOR_MODULES)
    mods.update(operator_modules or {})
    main_mod = sys.modules.get("__main__")
    operator_module = mods.get(mode) or getattr(main_mod, mode)
    class_obj = getattr(operator_module, mode)
    operator = class_obj.__new__(class_obj)

    # Step 5: injection
    setattr(operator, "task_index", task_index)
    setattr(operator, "job_name", job_name)
    setattr(operator, "optimizer", optimizer if optimizer is not None else default_optimizer())
    setattr(operator, "server", server)
    setattr(operator, "cluster", cluster)
    setattr(operator, "data_loader", data_loader)
    setattr(operator, "input_mode", input_mode)
    setattr(operator, "batch_size", batch_size)
    setattr(operator, "epoch_num", epoch_num)
    setattr(operator, "sample_number", sample_number)
    setattr(operator, "model_dir", model_dir)
    setattr(operator, "data_dir", data_dir)
    if job_name != "ps":
        experiment_model = annotations.get_instance_from_annotation(main, "model", model_module or model_mod)
        experiment_net = net_mod.Net(model=experiment_model)
        setattr(operator, "net", experiment_net)
    setattr(operator, "loss", loss)
    setattr(operator, "gpu_num", gpu_num)
    # advices placed on `main` (in addition to the class-level @get_advice of the reference) are forwarded
    for key in ("pre_fn", "post_fn", "pre_process_fn", "post_process_fn", "init_fn", "parse_data_dir_fn",
                "pre_eval_fn", "post_eval_fn"):
        if getattr(main, key, None) is not None:
            setattr(operator, key, getattr(main, key))
    return operator


def run_main(main, **kwargs):
    operator = build_operator(main, **kwargs)
    operator.run()
    return operator
