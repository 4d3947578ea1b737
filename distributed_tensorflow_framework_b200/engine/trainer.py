"""Trainer: one replica's training step (towers -> loss -> gradients -> fused collective + optimizer).

This is the per-step body that the reference builds once as a TF graph at /root/reference/distribute_train.py:106-158
(tower loop, Tower.process, average_gradients, SyncReplicasOptimizer.apply_gradients) and runs with
`sess.run([train_op, global_step, loss])` at :185. It is the public step API of the framework: the Train operator
(runtime/train.py), the Estimator (runtime/estimator.py) and bench.py all drive it.

CUDA path: the whole step (zero-grad, forward, loss, backward with bucketed fused collectives on a side stream,
optimizer, shadow refresh) is captured ONCE into a CUDA graph and replayed. Step-dependent optimizer scalars
(learning rate, Adam's bias-corrected rate, the step counter) live in device memory (`GradientEngine.hyper`), so a
replayed graph follows a learning-rate schedule and Adam can be captured. Inputs travel host -> device on a copy
stream into a pair of staging buffers (`prefetch`), one step ahead of the compute stream -- the analogue of the
reference's queue runners feeding `shuffle_batch` (/root/reference/distribute_input.py:54-69).
"""

import torch

from ..api import constants
from . import layers, variables as V
from .tower import Tower


class Trainer(object):
    def __init__(self, net, loss, optimizer, device="cpu", rank=0, world=1, group=None, gpu_num=1, algo="auto",
                 placement="bytes", use_cuda_graph=False, seed=0, sync_optimizer=None, pre_process_fn=None,
                 post_process_fn=None, pre_train_result=None, bucket_bytes=4 << 20, overlap=True,
                 replicas_to_aggregate=None, max_staleness=None):
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
        self.replicas_to_aggregate, self.max_staleness = replicas_to_aggregate, max_staleness
        self.engine = None
        self.built = False
        self._graph = None
        self._static_in = None
        self._loss_buf = None
        self._graph_launches = 0
        self.last_summaries = {}
        self.last_logits = None
        # input staging (CUDA): two device copies of the batch, filled by the copy stream one step ahead
        self._copy_stream = None
        self._staging = [None, None]
        self._staged_src = [None, None]  # host batch objects whose data sits in the slot (identity)
        self._staged_ready = [None, None]  # event: H2D into the slot has finished (copy stream)
        self._slot_free = [None, None]  # event: the step that read the slot has finished (compute stream)
        self._next_slot = 0

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
            raw = self._to_device(self._host_pairs((sample_raw, sample_gt))[0][0][0])
            self.store.losses = []
            self.net.process(raw)
        if self.device.type == "cuda":
            from ..parallel.comm import GradientEngine
            with torch.cuda.device(self.device):
                self.engine = GradientEngine(self.store, self.optimizer, self.rank, self.world, self.algo, self.group,
                                             bucket_bytes=self.bucket_bytes, placement=self.placement,
                                             overlap=self.overlap, replicas_to_aggregate=self.replicas_to_aggregate,
                                             max_staleness=self.max_staleness, num_towers=self.num_towers)
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

    def _step_body(self, batches, capturing=False):
        store = self.store
        step = store.global_step + 1
        if self.device.type == "cuda":
            # in-process towers accumulate into one gradient arena; a bucket's collective starts when the LAST tower
            # has produced it, and the 1/num_towers factor is part of the collective's scale (engine.num_towers)
            self.engine.begin_step(step, capturing=capturing)
            loss, _ = self._forward_backward(batches)
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

    # ------------------------------------------------------------------ host batches: staged H2D on a copy stream
    @staticmethod
    def _host_pairs(host_batches):
        """Normalise what a loader hands over: one (raw, gt) pair -- tuple or list -- or a list with one pair per
        tower; numpy arrays / nested lists become tensors (float64 -> float32, the default dtype of numpy data).
        Returns (pairs of tensors, the original objects as the identity key of the staging cache)."""
        import numpy as np
        if isinstance(host_batches, (tuple, list)) and len(host_batches) == 2 and \
                not isinstance(host_batches[0], (tuple, list)):
            host_batches = [host_batches]
        keys = [(r, g) for r, g in host_batches]

        def conv(t):
            if t is None or isinstance(t, torch.Tensor):
                return t
            t = torch.as_tensor(np.ascontiguousarray(t))
            return t.float() if t.dtype == torch.float64 else t
        return [(conv(r), conv(g)) for r, g in keys], keys

    def _stage(self, host_batches, keys=None):
        """Enqueue the H2D copy of `host_batches` (list of (raw, gt)) into the next staging slot on the copy stream.
        Returns the slot index."""
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=self.device)
        k = self._next_slot
        self._next_slot ^= 1
        cs = self._copy_stream
        if self._slot_free[k] is not None:
            cs.wait_event(self._slot_free[k])  # the step that last read this slot must have finished
        bufs = self._staging[k]
        ok = bufs is not None and len(bufs) == len(host_batches) and all(
            b[0].shape == r.shape and b[0].dtype == r.dtype and b[1].shape == g.shape and b[1].dtype == g.dtype
            for b, (r, g) in zip(bufs, host_batches))
        if not ok:
            bufs = [(torch.empty(r.shape, dtype=r.dtype, device=self.device),
                     torch.empty(g.shape, dtype=g.dtype, device=self.device)) for r, g in host_batches]
            self._staging[k] = bufs
            cs.wait_stream(torch.cuda.current_stream(self.device))  # fresh memory: order after its previous users
        with torch.cuda.stream(cs):
            for (dr, dg), (r, g) in zip(bufs, host_batches):
                dr.copy_(r, non_blocking=True)
                dg.copy_(g, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(cs)
        self._staged_ready[k] = ev
        self._staged_src[k] = keys if keys is not None else host_batches
        return k

    @staticmethod
    def _same_batches(a, b):
        return a is not None and len(a) == len(b) and all(x[0] is y[0] and x[1] is y[1] for x, y in zip(a, b))

    def prefetch(self, host_batches):
        """Start the H2D copy of the NEXT step's (pinned) host batch now, on the copy stream, while the current step
        is still running; the next `step()` call with the same batch objects finds the data already on the device."""
        if self.device.type != "cuda":
            return
        host_batches, keys = self._host_pairs(host_batches)
        if any(r.device.type == "cuda" for r, _ in host_batches):
            return
        for k in (0, 1):
            if self._same_batches(self._staged_src[k], keys):
                return
        self._stage(host_batches, keys)

    def step(self, host_batches):
        """Public per-step API: host (pinned) batches -> H2D on the copy stream -> step -> loss tensor on the device.
        `host_batches` is one (raw, gt) pair or a list with one pair per tower."""
        host_batches, keys = self._host_pairs(host_batches)
        if self.device.type != "cuda" or any(r.device.type == "cuda" for r, _ in host_batches):
            return self.step_device([(self._to_device(r), self._to_device(g)) for r, g in host_batches])
        k = None
        for j in (0, 1):
            if self._same_batches(self._staged_src[j], keys):
                k = j
        if k is None:
            k = self._stage(host_batches, keys)
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(self._staged_ready[k])
        self._staged_src[k] = None  # consumed: the same host object may carry new data next time
        loss = self.step_device(self._staging[k])
        ev = torch.cuda.Event()
        ev.record(cur)
        self._slot_free[k] = ev
        return loss

    # ------------------------------------------------------------------ CUDA graph replay
    def _graph_step(self, batches):
        if self._graph is None:
            self._capture(batches)
        if len(batches) != len(self._static_in) or any(
                r.shape != sr.shape or g.shape != sg.shape or r.dtype != sr.dtype or g.dtype != sg.dtype
                for (sr, sg), (r, g) in zip(self._static_in, batches)):
            # a batch of another shape (the smaller last batch of an epoch, say) cannot be replayed: this one step
            # runs eagerly; the device-side step counter / learning rate are shared, so replays simply continue after
            return self._step_body(batches)
        for (sr, sg), (r, g) in zip(self._static_in, batches):
            if r.data_ptr() != sr.data_ptr():
                sr.copy_(r, non_blocking=True)
            if g.data_ptr() != sg.data_ptr():
                sg.copy_(g, non_blocking=True)
        self._update_hyper()
        self._graph.replay()
        self.store.global_step += 1
        self.engine._step = self.store.global_step
        self.engine.note_replay()
        return self._loss_buf

    def _update_hyper(self):
        """Before a replay: push the learning rate to the device if the schedule (a LearningRate handler or a callable
        given to the optimizer) moved it, and the step counter if it was set from outside (checkpoint restore)."""
        self.engine.sync_hyper(self.store.global_step)

    def _state_tensors(self):
        st = self.store
        ts = [st.master_arena, st.shadow_arena, st.shadow_t_arena, st.slot_m, st.slot_v]
        ts += [v.master for v in st.all_variables() if not v.trainable]  # batch-norm running statistics
        return [t for t in ts if t is not None]

    def _quiesce(self):
        """Every rank's device work (including peer stores into this rank's heap) has finished."""
        torch.cuda.synchronize(self.device)
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier(group=self.group)

    def _capture(self, batches):
        self._static_in = [(r.clone(), g.clone()) for r, g in batches]
        # Autograd ties every parameter leaf to the stream that was current when the leaf was first used and joins
        # that stream at the end of backward; inside a capture this must be the capturing stream itself, otherwise
        # CUDA reports cudaErrorStreamCaptureIsolation. So: fresh leaves, warm-up and capture all on ONE side stream.
        self.store.rebind_shadow_leaves()
        s = self._cap_stream = torch.cuda.Stream(device=self.device)
        s.wait_stream(torch.cuda.current_stream(self.device))
        # The warm-up steps are real steps on the first batch; their effect is rolled back afterwards (parameters,
        # optimizer slots, running statistics, step counters), so that the first replay is step global_step + 1 and
        # the first batch is applied exactly once.
        step0 = self.store.global_step
        with torch.cuda.stream(s):
            saved = [t.clone() for t in self._state_tensors()]
            for _ in range(2):
                self._step_body(self._static_in)
        self._quiesce()
        with torch.cuda.stream(s):
            for t, c in zip(self._state_tensors(), saved):
                t.copy_(c)
        self.store.global_step = step0
        self.engine._step = step0
        torch.cuda.current_stream(self.device).wait_stream(s)
        self._quiesce()
        del saved
        self.engine.sync_hyper(step0)
        self._quiesce()
        self._graph = torch.cuda.CUDAGraph()
        from ..ops import _lib
        n0 = _lib.launch_count
        with torch.cuda.graph(self._graph, stream=s, capture_error_mode="thread_local"):
            self._loss_buf = self._step_body(self._static_in, capturing=True)
        self._graph_launches = _lib.launch_count - n0  # kernels of ours replayed by every graph launch
        # capture does not execute: undo the bookkeeping of the capture pass
        self.store.global_step = step0
        self.engine._step = step0
        self.engine._hyper_step = step0

    # ------------------------------------------------------------------ evaluation forward
    def forward(self, raw):
        with V.using_store(self.store), torch.no_grad():
            layers.set_training(False)
            try:
                return self.net.process(self._to_device(raw))
            finally:
                layers.set_training(True)
