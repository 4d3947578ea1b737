"""Train operator: parameter-server / worker roles, tower loop, synchronous (or async) replicas, monitored session.

Parity: /root/reference/distribute_train.py:22-221 -- `Train.train(pre_fn, post_fn, init_fn, pre_process_fn,
post_process_fn, parse_data_dir_fn, *args, **kwargs)` and `Train.run()`. Attributes are injected by the entry script
(/root/reference/distribute.py:110-127): task_index, job_name, optimizer, server, cluster, data_loader, input_mode,
batch_size, epoch_num, sample_number, model_dir, data_dir, net, loss, gpu_num.

Behavioural fixes: `(raw, gt) = pre_process_fn(raw, gt, ...)` (Q7); the log line reports real
seconds per batch (Q8); one shared global step (Q9); the placeholder queue is built once and every tower is fed
(Q10); the ps exits on a timeout and workers signal the done queue on any stop (Q19).
"""
import os
import time

import torch

from ..api import annotations
from ..api import flags
from ..api import log as logger
from ..engine.trainer import Trainer
from ..io.events import EventFileWriter
from ..plugins import input as Input
from ..utils import fault, nvtx
from . import hooks as H
from .cluster import DoneQueue
from .session import MonitoredTrainingSession, Saver, Scaffold
from .sync_replicas import AsyncReplicasOptimizer, SyncReplicasOptimizer


def _sync_mode(annotated=None):
    """sync | backup | async: the --sync_mode flag wins, then the @sync_mode annotation, then --sync_replicas."""
    m = (flags.FLAGS.sync_mode or annotated or "").lower()
    if m:
        return m
    if not flags.FLAGS.sync_replicas:
        return "async"
    return "sync"


def _record_files(data_dir):
    """TFRecord inputs named by @data_dir / parse_data_dir_fn: a file (the reference passes one path to
    string_input_producer, distribute_train.py:119-128), a list of files, a glob pattern, or a directory whose
    *.tfrecord / *.tfrecords / *.record files are taken in sorted order."""
    import glob
    if not isinstance(data_dir, str):
        return [str(p) for p in data_dir]
    if any(ch in data_dir for ch in "*?["):
        files = sorted(glob.glob(data_dir))
    elif os.path.isdir(data_dir):
        files = sorted(os.path.join(data_dir, f) for f in os.listdir(data_dir)
                       if f.endswith((".tfrecord", ".tfrecords", ".record")) or ".tfrecord-" in f)
    else:
        return [data_dir]
    if not files:
        raise ValueError("no TFRecord files found under %r" % data_dir)
    return files


@annotations.get_advice(pre_fn=None)
class Train(object):
    """Attributes injected by the IoC container: data_loader, input_mode, batch_size, epoch_num, sample_number,
    model_dir, data_dir, net, optimizer, loss, gpu_num, cluster, server, job_name, task_index; optional advices
    pre_fn, post_fn, pre_process_fn, post_process_fn, init_fn, parse_data_dir_fn."""

    @staticmethod
    def _create_done_queue(server, num_workers):
        return DoneQueue(server, num_workers)

    def _pick_device(self):
        if self.job_name == "ps" or self.gpu_num == 0 or not torch.cuda.is_available():
            return torch.device("cpu"), 1 if self.gpu_num == 0 else max(self.gpu_num, 1)
        local = int(os.environ.get("LOCAL_RANK", self.task_index % max(torch.cuda.device_count(), 1)))
        torch.cuda.set_device(local)
        return torch.device("cuda", local), self.gpu_num

    def _next_host_batch(self, name_queue, placeholder_queue):
        if self.input_mode == Input.InputOptions.PLACEHOLDER:
            return self.data_loader.load_placeholder_data(placeholder_queue)
        return self.data_loader.load_train_batch(name_queue)

    def train(self, pre_fn=None, post_fn=None, init_fn=None, pre_process_fn=None, post_process_fn=None,
              parse_data_dir_fn=None, *args, **kwargs):
        is_chief = self.task_index == 0
        replicas_to_aggregate = flags.FLAGS.replicas_to_aggregate
        total_step = self.epoch_num * (self.sample_number // self.batch_size)
        num_workers = max(self.cluster.num_tasks("worker"), 1)
        kill_ps_queue = Train._create_done_queue(self.server, num_workers)
        data_dir = self.data_dir if parse_data_dir_fn is None else parse_data_dir_fn(self.data_dir)

        # ---------------------------------------------------------------- parameter server role
        if self.job_name == "ps":
            # Parameters live on the owner shards of the worker GPUs; a ps process only keeps the reference's
            # life cycle: block until every worker has signalled the done queue (with a timeout, Q19).
            kill_ps_queue.wait_all(timeout_s=getattr(self, "ps_timeout_s", 24 * 3600))
            return

        # ---------------------------------------------------------------- worker role
        device, num_devices = self._pick_device()
        group = self.server.init_process_group() if num_workers > 1 else None
        pre_train_result = None if pre_fn is None else pre_fn(args, kwargs)

        mode = _sync_mode(getattr(self, "sync_mode", None))
        if replicas_to_aggregate is None:
            replicas_to_aggregate = num_workers
        if mode == "backup" and replicas_to_aggregate >= num_workers:
            replicas_to_aggregate = max(num_workers - 1, 1)  # "backup" without a count: one backup worker
        if mode == "sync" and replicas_to_aggregate < num_workers:
            mode = "backup"
        algo = flags.FLAGS.comm_algo
        if mode == "async" and device.type == "cuda" and num_workers > 1:
            algo = "async"
        elif device.type == "cuda" and num_workers > 1 and self.cluster.num_tasks("ps") > 0 and algo == "auto":
            # a cluster that declares ps tasks asks for parameter-server semantics: owner-sharded masters, push =
            # reduce-scatter + owner update, pull = all-gather (distribute_train.py:109-110)
            algo = "twoshot"
        base_opt = self.optimizer
        use_graph = getattr(self, "use_cuda_graph", None)
        if use_graph is None:
            use_graph = flags.FLAGS.use_cuda_graph
        trainer = Trainer(self.net, self.loss, base_opt, device=device, rank=self.task_index, world=num_workers,
                          group=group, gpu_num=num_devices, algo=algo, placement=flags.FLAGS.placement,
                          use_cuda_graph=use_graph, pre_process_fn=pre_process_fn,
                          post_process_fn=post_process_fn, pre_train_result=pre_train_result,
                          seed=getattr(self, "seed", 0),
                          replicas_to_aggregate=replicas_to_aggregate if mode == "backup" else None,
                          max_staleness=flags.FLAGS.max_staleness if mode == "async" else None)
        self.trainer = trainer

        # input pipeline (one logical pipeline per tower, like distribute_train.py:119-128)
        name_queue = placeholder_queue = None
        if self.input_mode == Input.InputOptions.TF_RECORD:
            name_queue = Input.FilenameQueue(_record_files(data_dir), shuffle=True)
        elif self.input_mode == Input.InputOptions.DATAPATHLOADER:
            name_queue = self.data_loader.create_name_queue(data_dir)
        elif self.input_mode == Input.InputOptions.PLACEHOLDER:
            self.data_loader.load_train_batch()  # placeholder specs (informational)
            placeholder_queue = self.data_loader.load_queue_for_placeholder(data_dir)  # built once (Q10)

        first = [self._next_host_batch(name_queue, placeholder_queue) for _ in range(num_devices)]
        trainer.build(first[0][0], first[0][1])
        store = trainer.store

        if mode == "async":
            optimizer = AsyncReplicasOptimizer(base_opt, total_num_replicas=num_workers, engine=trainer.engine)
        else:
            optimizer = SyncReplicasOptimizer(base_opt, use_locking=False, replicas_to_aggregate=replicas_to_aggregate,
                                              total_num_replicas=num_workers, name="sync_replicas",
                                              engine=trainer.engine, group=group, server=self.server)
        trainer.sync_optimizer = optimizer
        sync_replicas_hook = optimizer.make_session_run_hook(is_chief)

        if is_chief:
            logger.info("Worker %d: Initializing session..." % self.task_index)
        else:
            logger.info("Worker %d: Waiting for session to be initialized..." % self.task_index)

        writer = EventFileWriter(self.model_dir) if (is_chief and self.model_dir) else None
        engine = trainer.engine
        collective_save = engine is not None and engine.needs_gather()
        # owner-sharded masters: one data shard per owner rank (what the reference's ps sharding produces) and a
        # gather that every rank enters at the agreed step; replicated masters: the chief saves alone
        sharded = getattr(self, "sharded_checkpoint", None)
        if sharded is None:
            sharded = collective_save
        saver = Saver(store, base_opt, num_shards=num_workers if sharded else 1,
                      gather_fn=engine.gather_masters if collective_save else None,
                      lr_handler=getattr(self, "learning_rate", None) if is_chief else None)

        def agree(step, chief_decision):
            return self.server.agree("ckpt", step, chief_decision, is_chief)

        def on_recover(err):
            if engine is not None and engine.comm is not None:
                torch.cuda.synchronize(device)
                engine.comm.clear_error()
                torch.cuda.synchronize(device)

        def barrier():
            if group is not None:
                import torch.distributed as dist
                dist.barrier(group=group)

        session_hooks = [H.StopAtStepHook(last_step=total_step)]
        if sync_replicas_hook is not None:
            session_hooks.append(sync_replicas_hook)
        if trainer.engine is not None and trainer.engine.comm is not None and num_workers > 1:
            session_hooks.append(H.HeartbeatHook(trainer.engine.comm, every_n_steps=50, timeout_s=120.0))
        if flags.FLAGS.metrics_jsonl:
            session_hooks.append(H.MetricsJsonlHook(flags.FLAGS.metrics_jsonl, self.batch_size * num_devices * num_workers,
                                                    every_n_steps=10, extra={"rank": self.task_index}))
        session_hooks += list(getattr(self, "extra_hooks", []))
        if flags.FLAGS.log_device_placement:  # honoured here; the reference hard-codes False (distribute_train.py:176-178)
            for v in store.all_variables():
                logger.info("placement: %s -> owner rank %d (offset %d, %d elements)", v.name, v.owner, v.offset,
                            v.numel)
        pending = {"batches": first}
        state = {"loss": None}

        def train_step():
            batches = pending.pop("batches", None)
            if batches is None:
                batches = [self._next_host_batch(name_queue, placeholder_queue) for _ in range(num_devices)]
            with nvtx.range("train_step"):
                loss = trainer.step(batches)
            state["loss"] = loss
            if store.global_step < total_step:
                # while the device runs this step: load the next host batch and start its H2D copy on the copy stream
                nxt = [self._next_host_batch(name_queue, placeholder_queue) for _ in range(num_devices)]
                trainer.prefetch(nxt)
                pending["batches"] = nxt
            fault.maybe_kill(self.task_index, store.global_step)
            return {"loss": loss, "summaries": trainer.last_summaries}

        with MonitoredTrainingSession(master=self.server.target, is_chief=is_chief, checkpoint_dir=self.model_dir,
                                      scaffold=Scaffold(init_op=None, init_fn=init_fn), hooks=session_hooks,
                                      save_checkpoint_secs=flags.FLAGS.save_checkpoint_secs,
                                      config={"allow_soft_placement": True, "log_device_placement": False},
                                      stop_grace_period_secs=60, log_step_count_steps=100, store=store,
                                      optimizer=base_opt, saver=saver, summary_writer=writer,
                                      barrier_fn=barrier, collective_save=collective_save, agree_fn=agree,
                                      save_checkpoint_steps=flags.FLAGS.save_checkpoint_steps or None,
                                      on_recover=on_recover) as sess:
            logger.info("Worker %d: Session initialization complete." % self.task_index)
            self.session = sess
            while not sess.should_stop():
                if sess.global_step >= total_step:
                    break
                start = time.time()
                values = sess.run(train_step)
                step = sess.global_step
                if step % 10 == 0:
                    loss_value = float(values["loss"])  # device->host sync only on logging steps
                    duration = time.time() - start
                    num_examples_per_step = self.batch_size * max(self.gpu_num, 1)
                    examples_per_sec = num_examples_per_step / max(duration, 1e-9)
                    sec_per_batch = duration / max(self.gpu_num, 1)
                    format_str = ("step %d, loss = %.8f (%.1f examples/sec; %.3f sec/batch)")
                    logger.info(format_str % (step, loss_value, examples_per_sec, sec_per_batch))
                if step >= total_step:
                    break
            self.final_loss = float(state["loss"]) if state["loss"] is not None else None
            sess.close()
        # the session's end hooks have run (the chief has written the final checkpoint): no worker returns from
        # train() before that, so a script that goes on to evaluate finds the checkpoint on every rank
        barrier()
        kill_ps_queue.enqueue()
        logger.info("kill_ps_enqueue_op done....")
        if writer is not None:
            writer.close()

        if post_fn is not None:
            post_fn(args, kwargs)

    def run(self):
        g = lambda n: getattr(self, n, None)  # noqa: E731
        self.train(pre_fn=g("pre_fn"), post_fn=g("post_fn"), pre_process_fn=g("pre_process_fn"),
                   post_process_fn=g("post_process_fn"), init_fn=g("init_fn"),
                   parse_data_dir_fn=g("parse_data_dir_fn"))
