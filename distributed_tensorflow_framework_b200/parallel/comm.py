"""Python side of the peer-memory communication layer (csrc/comm.cu).

* `SymmetricHeap`  -- one cudaMalloc'd heap per rank with identical layout everywhere; buffers are torch views at
                      symmetric offsets, so a kernel can address a peer's copy as peer_base[r] + offset.
* `Communicator`   -- exchanges CUDA IPC handles over the torch.distributed control plane (bootstrap only) and wraps
                      the fused collectives: K-AR (allreduce+optimizer), K-RS-OPT (+all-gather by peer stores),
                      async push/apply, device barrier, heartbeat.
* `GradientEngine` -- binds a VariableStore to a communicator: arenas live in the heap, buckets of the gradient arena
                      are reduced as soon as backward has produced them (on a side stream), fused with x1/N and the
                      optimizer update.

Replaces: cross-tower gather+mean (/root/reference/distribute_tower.py:78-114), SyncReplicasOptimizer push/apply/pull
through the TF gRPC runtime (/root/reference/distribute_train.py:109-110,146-160) and the shared FIFOQueue barriers
(/root/reference/distribute_train.py:43-46,86-90).
"""
import ctypes
import os
from typing import List, Optional

import torch

from ..engine import variables as V
from ..ops import _lib

_ALIGN = 256


class _RawCuda(object):
    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class SymmetricHeap(object):
    """backend 'ipc': cudaMalloc + CUDA IPC handles (default). backend 'nvls': torch symmetric memory (VMM
    allocation bound to an NVLS multicast object); gives the same peer pointers plus `multicast_ptr`, which lets
    K-RS-OPT reduce and broadcast inside the NVSwitch (multimem.ld_reduce / multimem.st)."""

    def __init__(self, nbytes, device, backend="ipc", group=None):
        self.device = torch.device(device)
        self.nbytes = int(nbytes + _ALIGN - 1) // _ALIGN * _ALIGN
        self.backend = backend
        self.peer_ptrs = None
        self.multicast_ptr = 0
        if backend == "nvls":
            import torch.distributed as dist
            import torch.distributed._symmetric_memory as symm
            group = group or dist.group.WORLD
            self.nbytes = (self.nbytes + (2 << 20) - 1) // (2 << 20) * (2 << 20)
            self.bytes = symm.empty(self.nbytes, dtype=torch.uint8, device=self.device)
            self._hdl = symm.rendezvous(self.bytes, group.group_name)
            self.bytes.zero_()
            torch.cuda.synchronize(self.device)
            dist.barrier(group=group)
            self.ptr = self.bytes.data_ptr()
            self.peer_ptrs = [int(p) for p in self._hdl.buffer_ptrs]
            self.multicast_ptr = int(self._hdl.multicast_ptr or 0)
            self.handle = None
        else:
            ptr = ctypes.c_void_p()
            self.handle = ctypes.create_string_buffer(64)
            with torch.cuda.device(self.device):
                _lib.call("dtf_comm_heap_alloc", self.nbytes, ctypes.byref(ptr), self.handle)
            self.ptr = ptr.value
            self._raw = _RawCuda(self.ptr, self.nbytes)
            self.bytes = torch.as_tensor(self._raw, device=self.device)
        self._top = _lib.lib().dtf_comm_flag_page_bytes()  # the flag page sits at offset 0
        self.allocs = {}

    def alloc(self, tag, numel, dtype):
        """Bump allocation (deterministic => symmetric across ranks). Returns a zeroed flat tensor view."""
        esize = torch.empty((), dtype=dtype).element_size()
        off = (self._top + _ALIGN - 1) // _ALIGN * _ALIGN
        nbytes = numel * esize
        if off + nbytes > self.nbytes:
            raise MemoryError("symmetric heap exhausted: need %d more bytes" % (off + nbytes - self.nbytes))
        self._top = off + nbytes
        self.allocs[tag] = (off, numel, dtype)
        return self.bytes[off:off + nbytes].view(dtype)

    def offset_of(self, tag):
        return self.allocs[tag][0]

    def free(self):
        if self.ptr and self.backend == "ipc":
            _lib.call("dtf_comm_heap_free", ctypes.c_void_p(self.ptr))
            self.ptr = None


class Communicator(object):
    def __init__(self, heap: SymmetricHeap, rank: int, world: int, group=None, local_heaps: Optional[List] = None):
        self.heap, self.rank, self.world = heap, rank, world
        self.handle = ctypes.c_void_p()
        self.stream = torch.cuda.Stream(device=heap.device)
        self.blocks = int(os.environ.get("DTF_COMM_BLOCKS", "64"))  # CTAs per collective kernel
        if heap.peer_ptrs is not None:  # peers already mapped (torch symmetric memory / NVLS backend)
            arr = (ctypes.c_void_p * world)(*heap.peer_ptrs)
            _lib.call("dtf_comm_create_local", rank, world, arr, heap.nbytes, ctypes.byref(self.handle))
            if heap.multicast_ptr:
                _lib.call("dtf_comm_set_multicast", self.handle, ctypes.c_void_p(heap.multicast_ptr))
        elif local_heaps is not None:  # single process driving several peer-enabled GPUs (tests)
            arr = (ctypes.c_void_p * world)(*[h.ptr for h in local_heaps])
            _lib.call("dtf_comm_create_local", rank, world, arr, heap.nbytes, ctypes.byref(self.handle))
        else:
            handles = bytes(heap.handle.raw)
            if world > 1:
                import torch.distributed as dist
                gathered = [None] * world
                dist.all_gather_object(gathered, handles, group=group)
            else:
                gathered = [handles]
            blob = b"".join(gathered)
            with torch.cuda.device(heap.device):
                _lib.call("dtf_comm_create", rank, world, ctypes.c_void_p(heap.ptr), heap.nbytes, blob,
                          ctypes.byref(self.handle))
        self._snap = torch.zeros(8, device=heap.device, dtype=torch.int32)

    def _s(self, stream=None):
        return _lib.stream_ptr(stream)

    @staticmethod
    def _opt_args(opt, step, gscale, hyper=None):
        from ..ops.raw import OPTIM_KINDS
        return (OPTIM_KINDS[opt.kind], opt.learning_rate, opt.b1, opt.b2, opt.eps, opt.wd, gscale, step, hyper)

    def allreduce_optim(self, grad_off, start, count, opt, step, master, m, v, shadow, gscale=None, stream=None,
                        hyper=None):
        """K-AR: grads[start:start+count) of every rank -> mean -> optimizer on the local replica."""
        gscale = 1.0 / self.world if gscale is None else gscale
        _lib.call("dtf_comm_allreduce_optim", self.handle, grad_off, start, count,
                  *self._opt_args(opt, step, gscale, hyper), master, m, v, shadow, None, self.blocks, self._s(stream))

    def allreduce_mean(self, grad_off, start, count, out, gscale=None, stream=None):
        gscale = 1.0 / self.world if gscale is None else gscale
        _lib.call("dtf_comm_allreduce_optim", self.handle, grad_off, start, count, 0, 0.0, 0.0, 0.0, 0.0, 0.0, gscale,
                  1, None, None, None, None, None, out, self.blocks, self._s(stream))

    def rs_optim_ag(self, grad_off, bounds, opt, step, master, m, v, shadow_off, master_off=None, f32_range=None,
                    gscale=None, stream=None, nopush=(), hyper=None, backup_r=0):
        """K-RS-OPT: reduce-scatter into the owner shard + owner update + all-gather of bf16 weights by stores.
        `master_off` + `f32_range` (lo, hi): the owner also broadcasts its fp32 masters inside that element range
        (default: everywhere when only master_off is given). `nopush`: up to 8 (lo, hi) element ranges that stay
        owner-local (pulled later by K-AG-GEMM). `backup_r` in (0, world): aggregate the first backup_r arrivals
        only (backup workers); the default scale is then 1/backup_r."""
        backup = 0 < backup_r < self.world
        if gscale is None:
            gscale = 1.0 / (backup_r if backup else self.world)
        b = (ctypes.c_longlong * (self.world + 1))(*bounds)
        flat = [x for r in nopush for x in r]
        npr = (ctypes.c_longlong * max(len(flat), 1))(*flat) if flat else None
        if master_off is None:
            master_off, lo, hi = (1 << 64) - 1, 0, 0
        else:
            lo, hi = f32_range if f32_range is not None else (0, bounds[-1])
        _lib.call("dtf_comm_rs_optim_ag", self.handle, grad_off, b, *self._opt_args(opt, step, gscale, hyper), master,
                  m, v, shadow_off, master_off, lo, hi, len(nopush), npr, backup_r if backup else 0, self.blocks,
                  self._s(stream))

    def peer_ptr(self, r):
        return _lib.lib().dtf_comm_peer_ptr(self.handle, r)

    def allgather_slices(self, buf_off, bounds16, stream=None):
        b = (ctypes.c_longlong * (self.world + 1))(*bounds16)
        _lib.call("dtf_comm_allgather_slices", self.handle, buf_off, b, self.blocks, self._s(stream))

    def async_begin(self, stream=None):
        """Start of an asynchronous step: stamp the owner updates this step's weights correspond to."""
        _lib.call("dtf_comm_async_begin", self.handle, self._s(stream))

    def async_push(self, grad_off, mailbox_off, slice_cap, bounds, stream=None):
        """The version stamp lives in the flag page and is bumped by the kernel (graph replay publishes new ones)."""
        b = (ctypes.c_longlong * (self.world + 1))(*bounds)
        _lib.call("dtf_comm_async_push", self.handle, grad_off, mailbox_off, slice_cap, b, self.blocks,
                  self._s(stream))

    def async_apply(self, mailbox_off, slice_cap, bounds, opt, step, master, m, v, shadow_off, stream=None,
                    master_off=None, f32_range=None, max_staleness=-1, hyper=None):
        b = (ctypes.c_longlong * (self.world + 1))(*bounds)
        if master_off is None:
            master_off, lo, hi = (1 << 64) - 1, 0, 0
        else:
            lo, hi = f32_range if f32_range is not None else (0, bounds[-1])
        _lib.call("dtf_comm_async_apply", self.handle, mailbox_off, slice_cap, b,
                  *self._opt_args(opt, step, 1.0, hyper), master, m, v, shadow_off, master_off, lo, hi,
                  -1 if max_staleness is None else int(max_staleness), self._snap, self.blocks, self._s(stream))
        return self._snap

    def async_stats(self, stream=None):
        """{'dropped': pushes dropped as too stale, 'applied': updates applied by this owner, 'pushed': versions
        pushed by this rank}. Synchronises the stream."""
        out = (ctypes.c_uint * 3)()
        _lib.call("dtf_comm_async_stats", self.handle, out, self._s(stream))
        return {"dropped": int(out[0]), "applied": int(out[1]), "pushed": int(out[2])}

    def set_timeout_ms(self, ms):
        _lib.call("dtf_comm_set_timeout_ms", self.handle, int(ms))

    def error(self):
        """Host watchdog read (no CUDA call): None when healthy, else a dict describing the wait that timed out."""
        out = (ctypes.c_uint * 4)()
        code = _lib.call_raw("dtf_comm_error", self.handle, out)
        if not code:
            return None
        what = {1: "barrier", 2: "gradient arrivals", 3: "collective mismatch (a peer entered a different collective)"
                }.get(int(code), "wait")
        return {"code": int(code), "what": what, "block": int(out[1]), "peer": int(out[2]), "epoch": int(out[3])}

    def check(self):
        """Raise WorkerLostError when a device-side wait has expired (a peer is gone or stuck)."""
        e = self.error()
        if e is not None:
            from ..runtime.session import WorkerLostError
            raise WorkerLostError("rank %d: %s: device-side wait failed (block %d, peer/arrivals %d, epoch %d)"
                                  % (self.rank, e["what"], e["block"], e["peer"], e["epoch"]))

    def clear_error(self, stream=None):
        _lib.call("dtf_comm_clear_error", self.handle, self._s(stream))

    def sleep(self, seconds, stream=None):
        """Fault injection: keep the stream busy (an artificial straggler)."""
        _lib.call("dtf_debug_sleep", int(seconds * 1e9), self._s(stream))

    def barrier(self, stream=None):
        _lib.call("dtf_comm_barrier", self.handle, self._s(stream))

    def heartbeat(self, beat, stream=None):
        seen = torch.zeros(8, device=self.heap.device, dtype=torch.int32)
        _lib.call("dtf_comm_heartbeat", self.handle, beat, seen, self._s(stream))
        return seen

    def destroy(self):
        if self.handle:
            _lib.call("dtf_comm_destroy", self.handle)
            self.handle = None


def plan_buckets(variables, n_params, bucket_bytes):
    """Gradient buckets: contiguous [start, end) ranges of the arena in arena order, each closed at the first
    variable boundary at or past `bucket_bytes` of fp32 gradients (at least 1024 elements); the last one runs to the
    end of the arena. Returns [[start, end, [variables]]]: a partition of [0, n_params) in which every variable
    (with its alignment padding) lies in exactly one bucket."""
    per_bucket = max(bucket_bytes // 4, 1024)
    tv = sorted(variables, key=lambda v: v.offset)
    buckets = []
    cur_start, cur_vars = 0, []
    for v in tv:
        cur_vars.append(v)
        end = v.offset + V.padded_numel(v.numel)
        if end - cur_start >= per_bucket:
            buckets.append([cur_start, end, cur_vars])
            cur_start, cur_vars = end, []
    if cur_vars:
        buckets.append([cur_start, n_params, cur_vars])
    return buckets


class GradientEngine(object):
    """Connects a finalized VariableStore to the fused gradient collectives.

    algo: 'oneshot' (K-AR, replicated fp32 masters, bitwise-identical replicas), 'twoshot' (K-RS-OPT: each owner
    updates its shard and pushes bf16 weights; parameter-server semantics), 'nccl' (baseline: ncclAllReduce +
    separate optimizer kernel), 'local' (single replica: optimizer only), 'async' (K-ASYNC push/apply).
    Buckets are contiguous ranges of the gradient arena in reverse layer order; a bucket's collective is enqueued
    on the communication stream as soon as backward has produced all of its gradients.
    """

    def __init__(self, store, optimizer, rank=0, world=1, algo="auto", group=None,
                 bucket_bytes=int(os.environ.get("DTF_BUCKET_MB", "4")) << 20,
                 placement="bytes", heap_extra_bytes=0, overlap=True, replicas_to_aggregate=None, max_staleness=None,
                 num_towers=1):
        self.store, self.opt, self.rank, self.world, self.group = store, optimizer, rank, world, group
        dev = store.device
        self.num_towers = max(int(num_towers), 1)
        self.max_staleness = max_staleness
        # backup workers: aggregate the first R arrivals per owner (SyncReplicasOptimizer, distribute_train.py:146-155)
        r = world if replicas_to_aggregate is None else int(replicas_to_aggregate)
        self.backup_r = r if 0 < r < world else 0
        n_est = sum(V.padded_numel(v.numel) for v in store.trainable_variables())
        if self.backup_r:
            if algo not in ("auto", "twoshot"):
                raise ValueError("replicas_to_aggregate < num_workers needs the owner-sharded collective "
                                 "(comm_algo auto or twoshot), not %r" % algo)
            algo = "twoshot"  # peer-load kernel: the in-switch (NVLS) reduction cannot leave ranks out
        if algo == "auto":
            # small model on a pair of GPUs: one-shot K-AR (one kernel, replicated masters). Otherwise the two-shot
            # K-RS-OPT, over NVLS (in-switch reduce / broadcast) when a multicast mapping can be had.
            if world == 1:
                algo = "local"
            elif world <= 2 and n_est * 4 <= (16 << 20):
                algo = "oneshot"
            else:
                import os
                algo = "twoshot_nvls" if os.environ.get("DTF_NVLS", "1") != "0" else "twoshot"
        self.algo = algo
        self.overlap = overlap and algo in ("oneshot", "nccl", "twoshot", "twoshot_pull", "twoshot_nvls",
                                            "twoshot_nvls_pull")
        self.comm = None
        self.heap = None
        # "twoshot_nvls_pull": reduce-scatter / push over NVLS multimem, large FC matrices pulled by K-AG-GEMM
        self.pull = algo in ("twoshot_pull", "twoshot_nvls_pull")
        if algo == "twoshot_nvls_pull":
            algo = "twoshot_nvls"
        self.nvls = algo == "twoshot_nvls"
        if self.nvls:
            algo = self.algo = "twoshot"
        if algo == "twoshot_pull":
            algo = self.algo = "twoshot"
        use_heap = algo in ("oneshot", "twoshot", "async")
        if use_heap:
            mailbox = n_est * 4 * 1 if algo == "async" else 0
            need = n_est * (4 + 4 + 2 + 2 + 4 + 4) + mailbox * max(world, 1) + (1 << 20) + heap_extra_bytes
            self.heap = None
            if self.nvls and world > 1:
                self.heap = self._try_nvls_heap(need, dev, group)
            if self.heap is None:
                self.nvls = False
                self.heap = SymmetricHeap(need, dev, backend="ipc")
            store.finalize(allocator=self.heap.alloc, placement=placement, world=world, slots=optimizer.num_slots())
            if algo == "async":
                per = max(store.owner_bounds[i + 1] - store.owner_bounds[i] for i in range(world))
                self.slice_cap = (per + 3) // 4 * 4
                self.mailbox = self.heap.alloc("mailbox", self.slice_cap * world, torch.float32)
            self.comm = Communicator(self.heap, rank, world, group)
            self.nopush = []
            if self.pull and world > 1:
                self._setup_pull()
        else:
            store.finalize(placement=placement, world=world, slots=optimizer.num_slots())
        # fp32 masters of the filters are complete on every rank only when every rank applies the full update
        store.masters_replicated = self.algo in ("local", "oneshot", "nccl") or world == 1
        self.n = store.n_params
        # buckets: contiguous [start, end) in arena order, flushed from the tail (last layers finish first)
        self.buckets = plan_buckets(store.trainable_variables(), self.n, bucket_bytes)
        self._bucket_of = {}
        for bi, (_, _, vs) in enumerate(self.buckets):
            for v in vs:
                self._bucket_of[v.name] = bi
        self._pending = [0] * len(self.buckets)
        self._launched = [False] * len(self.buckets)
        self._step = 0
        self.comm_stream = self.comm.stream if self.comm is not None else torch.cuda.Stream(device=dev)
        if self.overlap:
            store.grad_ready_hook = self._on_grad_ready
        self.kernel_launches = 0
        # device-resident hyper-parameters: [lr, lr_t (Adam), int32 step, -]; the optimizer kernels read them at run
        # time, so a replayed CUDA graph follows a learning-rate schedule and Adam's bias correction
        self.hyper = torch.zeros(4, device=dev, dtype=torch.float32)
        self._hyper_lr = None
        self._hyper_step = None
        self.sync_hyper(store.global_step)

    # ------------------------------------------------------------------ device-resident hyper-parameters
    def sync_hyper(self, global_step):
        """Host -> device: the learning rate when the schedule moved it, the step counter when it was set from
        outside (checkpoint restore). Never called inside a graph capture; cheap when nothing changed."""
        lr = float(self.opt.learning_rate)
        if lr != self._hyper_lr:
            self.hyper[0:1].copy_(torch.tensor([lr], dtype=torch.float32))
            self._hyper_lr = lr
        if self._hyper_step != int(global_step):
            self.hyper[2:3].view(torch.int32).copy_(torch.tensor([int(global_step)], dtype=torch.int32))
            self._hyper_step = int(global_step)

    def _hyper_advance(self):
        """In-stream (capturable): step += 1 and lr_t refreshed on the device."""
        from ..ops.raw import OPTIM_KINDS
        _lib.call("dtf_optim_hyper_step", self.hyper, OPTIM_KINDS[self.opt.kind], self.opt.b1, self.opt.b2,
                  _lib.stream_ptr())

    def _try_nvls_heap(self, need, dev, group):
        """NVLS heap through torch symmetric memory; every rank must succeed, otherwise all fall back to CUDA IPC."""
        import torch.distributed as dist
        heap, ok = None, 1
        try:
            heap = SymmetricHeap(need, dev, backend="nvls", group=group)
            if not heap.multicast_ptr:
                ok = 0
        except Exception as e:  # no multicast support, old driver, fabric not configured ...
            if os.environ.get("DTF_NVLS", "1") == "require":
                raise  # a misconfigured fabric must not silently halve the collective's speed
            from ..api import log as logger
            self.nvls_fallback_reason = str(e)[:200]
            logger.warn("NVLS heap unavailable (%s); using CUDA IPC peer mappings (DTF_NVLS=require makes this an "
                        "error)", self.nvls_fallback_reason)
            ok = 0
        flag = torch.tensor([ok], device=dev, dtype=torch.int32)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        return heap if int(flag.item()) == 1 else None

    def _setup_pull(self):
        """Parameter-server PULL fused into the consumer (K-AG-GEMM): large FC weights are not broadcast after the
        owner update; the forward / dX GEMMs read each tile from its owner's heap. Owner boundaries that fall inside
        such a matrix are snapped to multiples of 128 weight rows so that every k-block / n-tile has one owner."""
        st = self.store
        fcs = [v for v in st.trainable_variables() if v.kind == "fc_w" and v.store_shape[0] % 128 == 0
               and v.store_shape[1] % 8 == 0 and v.numel >= (1 << 16)][:8]
        if not fcs:
            return
        bounds = list(st.owner_bounds)
        for v in fcs:
            fin, fout = v.store_shape
            unit = 128 * fout
            for i in range(1, self.world):
                if v.offset < bounds[i] < v.offset + v.numel:
                    k = int(round((bounds[i] - v.offset) / float(unit)))
                    bounds[i] = v.offset + min(max(k, 0), fin // 128) * unit
        for i in range(1, self.world + 1):
            bounds[i] = max(bounds[i], bounds[i - 1])
        st.owner_bounds = bounds
        sh_off, sht_off = self.heap.offset_of("shadow"), self.heap.offset_of("shadow_t")
        for v in fcs:
            v.ag = dict(world=self.world, bounds=bounds, off=v.offset, shape=tuple(v.store_shape),
                        peers=[self.comm.peer_ptr(r) + sh_off + 2 * v.offset for r in range(self.world)],
                        peers_t=[self.comm.peer_ptr(r) + sht_off + 2 * v.offset for r in range(self.world)])
            self.nopush.append((v.offset, v.offset + v.numel))

    # ------------------------------------------------------------------ per-step protocol
    def begin_step(self, step, capturing=False):
        """Call before the forward pass: zero the gradient arena, advance the device-side step / Adam rate, re-arm the
        buckets (a bucket is complete when every tower has produced every gradient in it)."""
        self._step = step
        if not capturing:
            self.sync_hyper(step - 1)
        self._hyper_advance()
        self._hyper_step = step
        self.store.zero_grads()
        if self.algo == "async" and self.comm is not None:
            self.comm.async_begin()
        for bi, (_, _, vs) in enumerate(self.buckets):
            self._pending[bi] = len(vs) * self.num_towers
            self._launched[bi] = False

    def note_replay(self):
        """A captured step was replayed: mirror the device-side step counter on the host."""
        self._hyper_step = (self._hyper_step or 0) + 1

    def _on_grad_ready(self, var):
        bi = self._bucket_of.get(var.name)
        if bi is None or self._launched[bi]:
            return
        self._pending[bi] -= 1
        if self._pending[bi] <= 0:
            self._launch_bucket(bi)

    def _launch_bucket(self, bi):
        start, end, _ = self.buckets[bi]
        self._launched[bi] = True
        cur = torch.cuda.current_stream(self.store.device)
        self.comm_stream.wait_stream(cur)
        with torch.cuda.stream(self.comm_stream):
            self._reduce_and_update(start, end - start)

    def _reduce_and_update(self, start, count):
        st, opt = self.store, self.opt
        step = self._step
        tw = 1.0 / self.num_towers  # in-process towers summed into one arena: their mean folds into the scale
        if self.algo == "oneshot":
            self.comm.allreduce_optim(self.heap.offset_of("grad"), start, count, opt, step, st.master_arena,
                                      st.slot_m, st.slot_v, st.shadow_arena, gscale=tw / self.world,
                                      stream=self.comm_stream, hyper=self.hyper)
        elif self.algo == "nccl":
            import torch.distributed as dist
            dist.all_reduce(st.grad_arena[start:start + count], group=self.group)
            opt.apply_arena(st, step, tw / self.world, start, count, refresh=False, hyper=self.hyper)
        elif self.algo == "twoshot":
            # the bucket is split among its owners: owner bounds clamped to [start, start + count)
            end = start + count
            bounds = [min(max(b, start), end) for b in st.owner_bounds]
            n_agg = self.backup_r or self.world
            self.comm.rs_optim_ag(self.heap.offset_of("grad"), bounds, opt, step, st.master_arena, st.slot_m,
                                  st.slot_v, self.heap.offset_of("shadow"), master_off=self.heap.offset_of("master"),
                                  f32_range=st.f32_ranges[self.rank], gscale=tw / n_agg, stream=self.comm_stream,
                                  nopush=self.nopush, hyper=self.hyper, backup_r=self.backup_r)
        else:
            raise RuntimeError(self.algo)

    def finish_step(self):
        """Call after backward: flush whatever has not been reduced yet, apply, refresh transposed shadows."""
        st, opt = self.store, self.opt
        step = self._step
        cur = torch.cuda.current_stream(st.device)
        if self.algo == "local":
            opt.apply_arena(st, step, 1.0 / self.num_towers, hyper=self.hyper)
            st.global_step = step
            return
        self.comm_stream.wait_stream(cur)
        with torch.cuda.stream(self.comm_stream):
            if self.algo in ("oneshot", "nccl", "twoshot"):
                if self.overlap:
                    for bi in range(len(self.buckets) - 1, -1, -1):
                        if not self._launched[bi]:
                            s, e, _ = self.buckets[bi]
                            self._launched[bi] = True
                            self._reduce_and_update(s, e - s)
                else:
                    self._reduce_and_update(0, self.n)
            elif self.algo == "async":
                self.comm.async_push(self.heap.offset_of("grad"), self.heap.offset_of("mailbox"), self.slice_cap,
                                     st.owner_bounds, stream=self.comm_stream)
                self.comm.async_apply(self.heap.offset_of("mailbox"), self.slice_cap, st.owner_bounds, opt, step,
                                      st.master_arena, st.slot_m, st.slot_v, self.heap.offset_of("shadow"),
                                      stream=self.comm_stream, master_off=self.heap.offset_of("master"),
                                      f32_range=st.f32_ranges[self.rank], max_staleness=self.max_staleness,
                                      hyper=self.hyper)
            st.refresh_transposed_shadows()
            if self.algo == "twoshot" and self.nopush:
                # peers pull the (transposed) shadows of these matrices from their owners: every rank must have
                # refreshed its own columns before anybody's next forward / backward starts
                self.comm.barrier(stream=self.comm_stream)
        cur.wait_stream(self.comm_stream)
        st.global_step = step

    def needs_gather(self):
        """True when fp32 masters / optimizer slots are complete only on their owner (owner-sharded modes)."""
        return self.algo in ("twoshot", "async") and self.world > 1 and self.comm is not None

    def gather_masters(self, slots=True):
        """Make every rank's fp32 master arena (and optimizer slots) complete in the owner-sharded modes.
        COLLECTIVE: every rank must call it at the same point of its stream (the session's checkpoint hooks agree on
        the step through the control plane); a rank calling it alone would pair its barrier with the peers' next
        gradient collective."""
        if not self.needs_gather():
            return
        b16 = [b // 4 for b in self.store.owner_bounds]  # fp32 elements -> 16-byte units
        tags = ["master"]
        if slots:
            tags += [t for t in ("slot_m", "slot_v") if t in self.heap.allocs]
        for tag in tags:
            self.comm.allgather_slices(self.heap.offset_of(tag), b16)
        torch.cuda.current_stream().synchronize()
        self.comm.check()

    def replicas_equal(self):
        """Collective check (every rank calls it): are the bf16 weights bit-identical on all ranks?"""
        if self.world == 1 or self.group is None:
            return True
        import torch.distributed as dist
        acc = self.store.shadow_arena.view(torch.int16).to(torch.int64)
        idx = torch.arange(acc.numel(), device=acc.device, dtype=torch.int64) % 65521 + 1
        sig = torch.stack([acc.sum(), (acc * idx).sum()])  # two checksums: plain and position-weighted
        lo, hi = sig.clone(), sig.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=self.group)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=self.group)
        return bool(torch.equal(lo, hi))
