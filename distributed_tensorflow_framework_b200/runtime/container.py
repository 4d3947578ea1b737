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
    if "TORCHELASTIC_RUN_ID" in os.environ or os.environ.get("DTF_SPAWNED") or \
            (not hasattr(main, "worker_hosts") and "WORLD_SIZE" in os.environ):
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
    if os.environ.get("DTF_SPAWNED") and os.environ.get("DTF_SPAWN_MODEL_DIR"):
        # a spawned tower re-executes the user's script: all towers must share the directory the parent resolved
        # (the script may have created a fresh temporary one per process)
        model_dir = os.environ["DTF_SPAWN_MODEL_DIR"]
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
    if os.environ.get("DTF_SPAWNED"):
        gpu_num = 1  # this process is ONE of the gpu_num towers spawned by run_main (one rank per GPU)
    setattr(operator, "gpu_num", gpu_num)
    # learning-rate handler (distribute_learningrate.py:12-50): an instance, or a class name looked up like the other
    # plugins and constructed from --train_learning_rate / --learning_rate_json. The optimizer reads its current
    # value every step (on CUDA through the device-resident hyper-parameters), the chief persists it with every
    # checkpoint, and a restarted run picks the saved value up again.
    lr_handler = getattr(main, "learning_rate", None)
    if lr_handler is not None:
        if isinstance(lr_handler, (str, type)):
            from ..plugins import learningrate as lr_mod
            cls = lr_handler if isinstance(lr_handler, type) else \
                (getattr(lr_mod, lr_handler, None) or annotations.registered(lr_handler))
            lr_handler = cls(flags.FLAGS.train_learning_rate, flags.FLAGS.learning_rate_json)
        setattr(operator, "learning_rate", lr_handler)
        opt = operator.optimizer
        if optimizer is None or not (callable(opt._lr) or hasattr(opt._lr, "learning_rate")):
            opt._lr = lr_handler
    if getattr(main, "sync_mode", None):
        setattr(operator, "sync_mode", getattr(main, "sync_mode"))
    # advices placed on `main` (in addition to the class-level @get_advice of the reference) are forwarded
    for key in ("pre_fn", "post_fn", "pre_process_fn", "post_process_fn", "init_fn", "parse_data_dir_fn",
                "pre_eval_fn", "post_eval_fn"):
        if getattr(main, key, None) is not None:
            setattr(operator, key, getattr(main, key))
    return operator


def _spawn_towers(n, main=None):
    """`@gpu_num(N)` without a launcher: the reference drives N GPUs from one process
    (`for i in range(num_devices): with tf.device('/gpu:%d')`, distribute_train.py:116-139). Here every tower is its
    own process on its own GPU (one rank per GPU over the peer-memory collectives): re-run this script N times with
    the torchrun environment variables set and wait for all of them. Returns the worst exit code."""
    import socket
    import subprocess
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    procs = []
    for r in range(n):
        env = dict(os.environ, DTF_SPAWNED="1", RANK=str(r), LOCAL_RANK=str(r), WORLD_SIZE=str(n),
                   LOCAL_WORLD_SIZE=str(n), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        if main is not None and getattr(main, "model_dir", None):
            env["DTF_SPAWN_MODEL_DIR"] = str(main.model_dir)
        procs.append(subprocess.Popen([sys.executable] + sys.argv, env=env))
    rc = 0
    for p in procs:
        rc = max(rc, abs(p.wait()))
    return rc


def _should_spawn(main):
    if os.environ.get("DTF_SPAWNED") or "TORCHELASTIC_RUN_ID" in os.environ or "WORLD_SIZE" in os.environ:
        return 0
    n = int(getattr(main, "gpu_num", 0) or 0)
    job = getattr(main, "job_name", None) or flags.FLAGS.job_name or "worker"
    workers = getattr(main, "worker_hosts", "") or flags.FLAGS.worker_hosts or ""
    if n <= 1 or job == "ps" or getattr(main, "mode", None) != "Train" or len([h for h in str(workers).split(",") if h.strip()]) > 1:
        return 0
    try:
        import torch
        return n if torch.cuda.is_available() and torch.cuda.device_count() >= n else 0
    except Exception:  # pragma: no cover
        return 0


def run_main(main, **kwargs):
    n = _should_spawn(main)
    if n:
        rc = _spawn_towers(n, main)
        if rc:
            raise RuntimeError("a tower process failed (exit code %d)" % rc)
        return None
    operator = build_operator(main, **kwargs)
    operator.run()
    return operator
