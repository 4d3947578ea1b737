"""Cluster description and process bootstrap (the framework's replacement for tf.train.ClusterSpec / tf.train.Server).

Parity: /root/reference/distribute.py:55-81 builds ClusterSpec({"ps": [...], "worker": [...]}) from comma-separated
host strings and starts an in-process gRPC server. Here the "server" is the CPU control plane only: a TCPStore hosted
by ps task 0 (or worker 0 when there is no ps) used for rendezvous, the done queue and the exchange of CUDA IPC
handles; tensors never travel through it. Variables are not placed on ps processes: each GPU rank owns a shard
(runtime/placement.py), so a `ps` process merely keeps the reference's life-cycle (it blocks on the done queue).
Host strings are whitespace-stripped.
"""
import datetime
import os
import time
from typing import Dict, List

from ..api import log as logger


def _split_hosts(s):
    if s is None:
        return []
    if isinstance(s, (list, tuple)):
        return [h.strip().replace(" ", "") for h in s if h and h.strip()]
    return [h.strip().replace(" ", "") for h in str(s).split(",") if h.strip()]


class ClusterSpec(object):
    def __init__(self, jobs: Dict[str, List[str]]):
        self._jobs = {k: _split_hosts(v) for k, v in jobs.items()}

    def num_tasks(self, job_name):
        return len(self._jobs.get(job_name, []))

    def job_tasks(self, job_name):
        return list(self._jobs.get(job_name, []))

    def task_address(self, job_name, task_index):
        return self._jobs[job_name][task_index]

    def as_dict(self):
        return {k: list(v) for k, v in self._jobs.items()}

    @property
    def jobs(self):
        return list(self._jobs)

    def __repr__(self):
        return "ClusterSpec(%r)" % self._jobs


def cluster_from_env():
    """torchrun launch (RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT): N workers, no ps."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    addr = os.environ.get("MASTER_ADDR", "127.0.0.1")
    port = int(os.environ.get("MASTER_PORT", "29500"))
    workers = ["%s:%d" % (addr, port + i) for i in range(world)]
    return ClusterSpec({"ps": [], "worker": workers}), "worker", int(os.environ.get("RANK", "0"))


class Server(object):
    """Control-plane endpoint of one task. `target` names the rendezvous store; `join()` blocks like
    tf.train.Server.join(). For workers, `init_process_group()` brings up torch.distributed (gloo on CPU, nccl on
    CUDA) over the same store."""

    def __init__(self, cluster: ClusterSpec, job_name: str, task_index: int, start=True, timeout_s=300):
        self.cluster, self.job_name, self.task_index = cluster, job_name, int(task_index or 0)
        self.timeout_s = timeout_s
        self.store = None
        self._pg = None
        ps = cluster.job_tasks("ps")
        workers = cluster.job_tasks("worker")
        host = ps[0] if ps else (workers[0] if workers else "127.0.0.1:29500")
        self._host, port = host.rsplit(":", 1)
        self._port = int(port)
        self.is_store_master = (job_name == "ps" and self.task_index == 0) if ps else \
            (job_name == "worker" and self.task_index == 0)
        self.single = (cluster.num_tasks("worker") <= 1 and not ps)
        if start and not self.single and "TORCHELASTIC_RUN_ID" not in os.environ:
            self._start_store()

    @property
    def target(self):
        return "dtf://%s:%d" % (self._host, self._port)

    def _start_store(self):
        from torch.distributed import TCPStore
        n_clients = self.cluster.num_tasks("worker") + self.cluster.num_tasks("ps")
        self.store = TCPStore(self._host, self._port, world_size=None if self.is_store_master else n_clients,
                              is_master=self.is_store_master, timeout=datetime.timedelta(seconds=self.timeout_s),
                              wait_for_workers=False)

    def init_process_group(self, backend=None):
        """Bring up torch.distributed among the WORKERS (ps tasks do not participate)."""
        import torch
        import torch.distributed as dist
        n = self.cluster.num_tasks("worker")
        if n <= 1:
            return None
        if dist.is_initialized():
            return dist.group.WORLD
        backend = backend or ("nccl" if torch.cuda.is_available() else "gloo")
        if "TORCHELASTIC_RUN_ID" in os.environ or self.store is None:
            restart = os.environ.get("TORCHELASTIC_RESTART_COUNT", "0")
            if restart != "0" and "MASTER_ADDR" in os.environ and "MASTER_PORT" in os.environ:
                # a group restarted by the launcher after a lost worker: rendezvous under a fresh key prefix, otherwise
                # a rank may read the previous attempt's (dead) NCCL bootstrap address from the launcher's store
                from torch.distributed import PrefixStore, TCPStore
                base = TCPStore(os.environ["MASTER_ADDR"], int(os.environ["MASTER_PORT"]), world_size=None,
                                is_master=False, timeout=datetime.timedelta(seconds=self.timeout_s))
                dist.init_process_group(backend=backend, store=PrefixStore("dtf_attempt_%s" % restart, base),
                                        rank=int(os.environ["RANK"]), world_size=int(os.environ["WORLD_SIZE"]),
                                        timeout=datetime.timedelta(seconds=self.timeout_s))
            else:
                dist.init_process_group(backend=backend)
        else:
            from torch.distributed import PrefixStore
            dist.init_process_group(backend=backend, store=PrefixStore("pg", self.store), rank=self.task_index,
                                    world_size=n, timeout=datetime.timedelta(seconds=self.timeout_s))
        return dist.group.WORLD

    def control_store(self):
        """Key-value store of the CPU control plane: the Server's own TCPStore, or -- under torchrun -- the store
        torch.distributed was initialised with. None in a single-process run."""
        if self.store is not None:
            return self.store
        try:
            import torch.distributed as dist
            if dist.is_initialized():
                from torch.distributed import distributed_c10d as c10d
                return c10d._get_default_store()
        except Exception:  # pragma: no cover
            pass
        return None

    def agree(self, tag, step, chief_value, is_chief):
        """All workers adopt the chief's boolean for (tag, step): the chief publishes it, the others block on the key."""
        st = self.control_store()
        if st is None:
            return bool(chief_value)
        key = "agree/%s/%d" % (tag, int(step))
        if is_chief:
            st.set(key, b"1" if chief_value else b"0")
            return bool(chief_value)
        return st.get(key) == b"1"

    def join(self):
        while True:
            time.sleep(3600)


class DoneQueue(object):
    """The shared FIFOQueue "done_queue0" (/root/reference/distribute_train.py:43-46): workers enqueue one token when
    they stop, ps tasks dequeue `num_workers` tokens and exit. Backed by the control-plane store; the ps wait has a
    timeout and workers signal on ANY stop, not only at step == total_step."""

    KEY = "done_queue0"

    def __init__(self, server: Server, num_workers: int):
        self.server, self.num_workers = server, num_workers
        self._local = 0

    def enqueue(self):
        if self.server.store is not None:
            self.server.store.add(self.KEY, 1)
        else:
            self._local += 1

    def size(self):
        if self.server.store is not None:
            return int(self.server.store.add(self.KEY, 0))
        return self._local

    def wait_all(self, timeout_s=None, poll_s=0.2):
        """ps side: block until every worker has signalled. Returns False on timeout."""
        t0 = time.time()
        while self.size() < self.num_workers:
            if timeout_s is not None and time.time() - t0 > timeout_s:
                logger.warn("done queue: timed out with %d of %d workers", self.size(), self.num_workers)
                return False
            time.sleep(poll_s)
        return True
