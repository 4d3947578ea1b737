"""Variable placement policies.

Parity: `local_device_setter` (/root/reference/distribute_utils.py:118-145) -- a single-host "parameter server":
variable ops are placed round-robin on `ps_device_type:k`, everything else on `worker_device` -- and
tf.train.replica_device_setter as used at /root/reference/distribute_train.py:109-110 (variables round-robin over the
ps tasks in creation order). Here a "ps device" is the owner shard of a GPU rank; the chooser returned below maps a
variable (or a (op_type, name) pair) to a device string, and `owner_map` gives the parameter -> owner-rank table
that the K-RS-OPT / async kernels use.
"""
from typing import Iterable, List

PS_OPS = ["Variable", "VariableV2", "VarHandleOp"]


class _RoundRobinStrategy(object):
    def __init__(self, num_tasks):
        self._num_tasks = max(int(num_tasks), 1)
        self._next = 0

    def __call__(self, op=None):
        task = self._next
        self._next = (self._next + 1) % self._num_tasks
        return task


class _GreedyLoadBalancingStrategy(object):
    """Place each variable on the least-loaded shard (by bytes)."""

    def __init__(self, num_tasks, load_fn=None):
        self._loads = [0] * max(int(num_tasks), 1)
        self._load_fn = load_fn or (lambda op: getattr(op, "numel", 1))

    def __call__(self, op=None):
        task = min(range(len(self._loads)), key=lambda i: self._loads[i])
        self._loads[task] += self._load_fn(op)
        return task


def local_device_setter(num_devices=1, ps_device_type="cpu", worker_device="/cpu:0", ps_ops=None, ps_strategy=None):
    if ps_ops is None:
        ps_ops = list(PS_OPS)
    if ps_strategy is None:
        ps_strategy = _RoundRobinStrategy(num_devices)
    if not callable(ps_strategy):
        raise TypeError("ps_strategy must be callable")

    def _local_device_chooser(op):
        op_type = getattr(op, "op_type", None)
        if op_type is None:
            if isinstance(op, tuple):
                op_type = op[0]
            elif hasattr(op, "trainable") or hasattr(op, "store_shape"):
                op_type = "VariableV2"
            else:
                op_type = str(op)
        if op_type in ps_ops:
            return "/%s:%d" % (ps_device_type, ps_strategy(op))
        return worker_device or ""

    return _local_device_chooser


def replica_device_setter(ps_tasks, worker_device="/job:worker/task:0", ps_device="/job:ps", ps_strategy=None):
    strategy = ps_strategy or _RoundRobinStrategy(ps_tasks)

    def chooser(op):
        op_type = op[0] if isinstance(op, tuple) else ("VariableV2" if hasattr(op, "store_shape") else str(op))
        if op_type in PS_OPS and ps_tasks > 0:
            return "%s/task:%d" % (ps_device, strategy(op))
        return worker_device

    return chooser


def owner_map(variables: Iterable, num_owners: int, policy="round_robin") -> List[int]:
    """parameter -> owner rank. 'round_robin' mirrors replica_device_setter (creation order); 'bytes' balances."""
    variables = list(variables)
    if policy == "round_robin":
        s = _RoundRobinStrategy(num_owners)
    else:
        s = _GreedyLoadBalancingStrategy(num_owners)
    return [s(v) for v in variables]
