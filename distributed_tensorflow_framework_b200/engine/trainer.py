"""Trainer: one replica's training step (towers -> loss -> gradients -> fused collective + optimizer).

This is the per-step body that the reference builds once as a TF graph at /root/reference/distribute_train.py:106-158
(tower loop, Tower.process, average_gradients, SyncReplicasOptimizer.apply_gradients) and runs with
`sess.run([train_op, global_step, loss])` at :185. It is the public step API of the framework: the Train operator
(runtime/train.py) and bench.py both drive it.

CUDA path: the whole step (zero-grad, forward, loss, backward with bucketed fused collectives on a side stream,
optimizer, shadow refresh) can be captured ONCE into a CUDA graph and replayed; inputs are copied from pinned host
memory into static device buffers before every replay, and step-dependent optimizer scalars live in device memory.
"""

import torch

from ..api import constants
from ..api import log as logger
from . import layers, variables as V
from .tower import Tower


class Trainer(object):
    def __init__(self, net, loss, optimizer, device="cpu", rank=0, world=1, group=None, gpu_num=1, algo="auto",
                 placement="bytes", use_cuda_graph=False, seed=0, sync_optimizer=None, pre_process_fn=None,
                 post_process_fn=None, pre_train_result=None, bucket_bytes=4 << 20, overlap=True):
        self.net, self.loss, self.optimizer = net, loss, optimizer
        self.device = torch.device(device)
        self.rank, self.world, self.group = rank, world, group
        self.num_towers = max(int(gpu_num), 1)
        self.algo, self.placement = algo, placement
        self.use_cuda_graph = use_cuda_graph and self.device.type == "cuda"
        self.store = V.VariableStore(self.device, seed=seed)
        self.sync_optimizer = sync_optimizer
        self.pre_process_fn, self.post_process_fn = pre_process_fn, post_process_fn
        self.pre_train_result = pre_train_result
        self.bucket_bytes, self.overlap = bucket_bytes, overlap
        self.engine = None
        self.built = False
        self._graph = None
        self._static_in = None
        self._loss_buf = None
        self._graph_launches = 0
        self.last_summaries = {}
        self.last_logits = None

    # ------------------------------------------------------------------ build (variable discovery + arenas)
    def _to_device(self, t):
        if isinstance(t, torch.Tensor) and t.device != self.device:
            return t.to(self.device, non_blocking=True)
        return t

    def build(self, sample_raw, sample_gt=None):
        """Run one forward pass without gradients to create every variable (the reference's graph-construction
        phase), then pack them into arenas and attach the gradient engine."""
        if self.built:
            return
        with V.using_store(self.store), torch.no_grad():
            raw = self._to_device(sample_raw)
            self.store.losses = []
            self.net.process(raw)
        if self.device.type == "cuda":
            from ..parallel.comm import GradientEngine
            with torch.cuda.device(self.device):
                self.engine = GradientEngine(self.store, self.optimizer, self.rank, self.world, self.algo, self.group,
                                             bucket_bytes=self.bucket_bytes, placement=self.placement,
                                             overlap=self.overlap)
            self.algo = self.engine.algo
        else:
            self.store.finalize()
        self.built = True

    # ------------------------------------------------------------------ one step
    def _forward_backward(self, batches):
        """batches: list of (raw, gt) per tower, already on the device. Returns the mean tower loss (tensor)."""
        tower_grads, tower_losses = [], []
        with V.using_store(self.store):
            for i, (raw, gt) in enumerate(batches):
                scope = "%s_%d" % (constants.TOWER_NAME, i)
                if self.pre_process_fn is not None:
                    raw, gt = self.pre_process_fn(raw, gt, (), {})
                tower = Tower(self.net, scope, tower_grads, raw, gt, self.loss, self.optimizer, self.store)
                if self.pre_train_result is not None:
                    summaries, loss, logits = tower.process(self.post_process_fn, self.pre_train_result)
                else:
                    summaries, loss, logits = tower.process(self.post_process_fn)
                tower_losses.append(loss.detach().float())
                self.last_summaries, self.last_logits = summaries, logits
        loss = tower_losses[0] if len(tower_losses) == 1 else torch.stack(tower_losses).mean()
        return loss, tower_grads

    def _step_body(self, batches):
        store = self.store
        step = store.global_step + 1
        if self.device.type == "cuda":
            self.engine.begin_step(step)
            loss, _ = self._forward_backward(batches)
            # in-process towers accumulate into the same arena: their mean is folded into the collective's scale
            if self.num_towers > 1:
                store.grad_arena.mul_(1.0 / self.num_towers)
            self.engine.finish_step()
            return loss
        loss, tower_grads = self._forward_backward(batches)
        grads = Tower.average_gradients(tower_grads)
        opt = self.sync_optimizer or self.optimizer
        opt.apply_gradients(grads, global_step=store.global_step, store=store)
        return loss

    def step_device(self, batches):
        """One training step on device-resident batches. Returns the loss as a device tensor (no sync)."""
        if not self.built:
            self.build(batches[0][0], batches[0][1])
        if not self.use_cuda_graph:
            return self._step_body(batches)
        return self._graph_step(batches)

    def step(self, host_batches):
        """Public per-step API: host (pinned) batches -> async H2D -> step -> loss tensor on the device.
        `host_batches` is one (raw, gt) pair or a list with one pair per tower."""
        if isinstance(host_batches, tuple):
            host_batches = [host_batches]
        dev = [(self._to_device(r), self._to_device(g)) for r, g in host_batches]
        return self.step_device(dev)

    # ------------------------------------------------------------------ CUDA graph replay
    def _graph_step(self, batches):
        if self._graph is None:
            self._capture(batches)
        for (sr, sg), (r, g) in zip(self._static_in, batches):
            if r.data_ptr() != sr.data_ptr():
                sr.copy_(r, non_blocking=True)
            if g.data_ptr() != sg.data_ptr():
                sg.copy_(g, non_blocking=True)
        self._update_hyper()
        self._graph.replay()
        self.store.global_step += 1
        self.engine._step = self.store.global_step
        return self._loss_buf

    def _update_hyper(self):
        pass  # reserved: device-resident lr / bias-correction scalars for schedules under graph replay

    def _capture(self, batches):
        if self.optimizer.kind == "adam":
            logger.warn("CUDA graph replay freezes Adam's bias correction at the capture step; using eager steps")
            self.use_cuda_graph = False
            return
        self._static_in = [(r.clone(), g.clone()) for r, g in batches]
        # Autograd ties every parameter leaf to the stream that was current when the leaf was first used and joins
        # that stream at the end of backward; inside a capture this must be the capturing stream itself, otherwise
        # CUDA reports cudaErrorStreamCaptureIsolation. So: fresh leaves, warm-up and capture all on ONE side stream.
        self.store.rebind_shadow_leaves()
        s = self._cap_stream = torch.cuda.Stream(device=self.device)
        s.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(s):
            for _ in range(2):
                self._step_body(self._static_in)
        torch.cuda.current_stream(self.device).wait_stream(s)
        torch.cuda.synchronize(self.device)
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier(group=self.group)
        self._graph = torch.cuda.CUDAGraph()
        from ..ops import _lib
        n0 = _lib.launch_count
        with torch.cuda.graph(self._graph, stream=s, capture_error_mode="thread_local"):
            self._loss_buf = self._step_body(self._static_in)
        self._graph_launches = _lib.launch_count - n0  # kernels of ours replayed by every graph launch
        # capture does not execute: undo the step increments of the capture pass
        self.store.global_step -= 1

    # ------------------------------------------------------------------ evaluation forward
    def forward(self, raw):
        with V.using_store(self.store), torch.no_grad():
            layers.set_training(False)
            try:
                return self.net.process(self._to_device(raw))
            finally:
                layers.set_training(True)
