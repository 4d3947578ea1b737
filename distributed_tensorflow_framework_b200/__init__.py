"""distributed_tensorflow_framework_b200: a B200-native framework with the capabilities and plugin API of
Seanforfun/Distributed-Tensorflow-Framework (annotation / reflection driven registration of model, net, loss, input,
learning-rate and eval handlers; ps / worker cluster flags; sync, backup-worker and async replicas; train / eval
operators with experiment and estimator wrappers; TF1 Saver checkpoints), built on PyTorch + hand-written sm_100a
CUDA kernels + NVLink peer-memory collectives. See DESIGN.md.

Module map (reference file -> here):
  distribute_annotations -> api.annotations      distribute_flags   -> api.flags
  distribute_constants   -> api.constants        distribute_log     -> api.log
  distribute_model/net/loss/input/learningrate -> plugins.*
  distribute_tools       -> engine.layers        distribute_tower   -> engine.tower
  distribute_train       -> runtime.train        distribute_eval    -> runtime.evalop
  distribute_experiment  -> runtime.experiment   distribute_estimator -> runtime.estimator
  distribute_utils       -> runtime.hooks / runtime.placement / runtime.estimator.RunConfig
  distribute.py (entry)  -> runtime.container + examples/distribute.py
"""
from .api import annotations, constants, flags, log  # noqa: F401
from .api.flags import FLAGS  # noqa: F401
from .plugins import (Model, Net, Loss, SoftmaxCrossEntropyLoss, MSELoss, LearningRate, Dataloader,  # noqa: F401
                      TFRecordDataLoader, DataPathDataLoader, PlaceholderDataLoader, SyntheticDataLoader,
                      InputOptions)
from .engine import layers as tools  # noqa: F401
from .engine import layers  # noqa: F401
from .engine.optimizers import AdamOptimizer, GradientDescentOptimizer, MomentumOptimizer  # noqa: F401
from .engine.tower import Tower  # noqa: F401
from .engine.trainer import Trainer  # noqa: F401
from .engine.variables import VariableStore, variable_scope  # noqa: F401
from .runtime.cluster import ClusterSpec, Server, DoneQueue  # noqa: F401
from .runtime.container import build_operator, run_main  # noqa: F401
from .runtime.estimator import DistributeEstimator, Estimator, EstimatorSpec, RunConfig  # noqa: F401
from .runtime.evalop import Eval  # noqa: F401
from .runtime.experiment import DistributeExperiment  # noqa: F401
from .runtime.hooks import (ExamplesPerSecondHook, StopAtStepHook, StepCounterHook, CheckpointSaverHook,  # noqa: F401
                            SummarySaverHook, SessionRunHook)
from .runtime.placement import local_device_setter, replica_device_setter, owner_map  # noqa: F401
from .runtime.session import MonitoredTrainingSession, Scaffold, Saver, WorkerLostError  # noqa: F401
from .runtime.sync_replicas import (SyncReplicasOptimizer, AsyncReplicasOptimizer,  # noqa: F401
                                    InProcessParameterServer, ConditionalAccumulator)
from .runtime.train import Train  # noqa: F401

__version__ = "0.1.0"
