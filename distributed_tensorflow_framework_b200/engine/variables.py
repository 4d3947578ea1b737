"""Variable store: TF1-style named variables (`variable_scope` + `get_variable`) on flat device arenas.

The reference creates variables with tf.get_variable under tf.variable_scope(layer_name)
(/root/reference/distribute_tools.py:24-66,70-75,98-103,133-135,198-203); those names ("conv1/weights", ...) are the
checkpoint keys, and towers share variables through scope reuse (/root/reference/distribute_tower.py:45-48).

B200 design: after the first forward pass has discovered every variable, `finalize()` packs all trainable
parameters into a few flat arenas -- fp32 master, fp32 gradient, bf16 shadow (tensor-core operand, TF HWIO / [in,out]
layout) and a second bf16 shadow with every filter tap transposed (operand of the data-gradient / transposed-conv
GEMM) plus optimizer slots -- so that the gradient collective + optimizer is ONE fused kernel per bucket over
contiguous memory and the arenas can live in the NVLink symmetric heap. Variables are views into the arenas.
"""
from __future__ import annotations

import contextlib
import math
import threading
from typing import Callable, Dict, List, Optional

import torch

_tls = threading.local()


def _scope_stack() -> List[str]:
    if not hasattr(_tls, "scopes"):
        _tls.scopes = []
    return _tls.scopes


@contextlib.contextmanager
def variable_scope(name: str, reuse=None):
    """tf.variable_scope(name): prefixes variable names. `reuse` is accepted for API parity; variables are always
    created on first use and reused afterwards (towers and *_eval layers share variables by name)."""
    _scope_stack().append(name)
    try:
        yield "/".join(_scope_stack())
    finally:
        _scope_stack().pop()


def current_scope() -> str:
    return "/".join(_scope_stack())


# ---- initializers (TF semantics)
class Initializer(object):
    kind = "constant"

    def params(self, shape):
        return 0.0, 0.0


class ConstantInitializer(Initializer):
    kind = "constant"

    def __init__(self, value=0.0):
        self.value = float(value)

    def params(self, shape):
        return self.value, 0.0


class TruncatedNormalInitializer(Initializer):
    """tf.truncated_normal_initializer(stddev): N(0, stddev) re-drawn outside 2 sigma (distribute_tools.py:62)."""
    kind = "truncated_normal"

    def __init__(self, stddev=1.0, mean=0.0):
        self.stddev, self.mean = float(stddev), float(mean)

    def params(self, shape):
        return self.mean, self.stddev


class XavierInitializer(Initializer):
    """tf.contrib.layers.xavier_initializer() (uniform): U(-l, l), l = sqrt(6 / (fan_in + fan_out))
    (distribute_tools.py:100,135,200)."""
    kind = "uniform"

    def params(self, shape):
        shape = list(shape)
        if len(shape) < 1:
            fan_in = fan_out = 1.0
        elif len(shape) == 1:
            fan_in = fan_out = float(shape[0])
        elif len(shape) == 2:
            fan_in, fan_out = float(shape[0]), float(shape[1])
        else:
            rf = 1.0
            for d in shape[:-2]:
                rf *= d
            fan_in, fan_out = shape[-2] * rf, shape[-1] * rf
        limit = math.sqrt(6.0 / (fan_in + fan_out))
        return -limit, limit


class Variable(object):
    """A named parameter. `shape` is the logical (checkpoint) shape; `store_shape` may pad the input-channel
    dimension of a filter to a multiple of 8 so that 16-byte im2col chunks never straddle taps."""

    def __init__(self, name, shape, initializer, trainable=True, kind="generic", dtype=torch.float32,
                 store_shape=None, wd=None):
        self.name = name
        self.shape = tuple(int(s) for s in shape)
        self.store_shape = tuple(int(s) for s in (store_shape if store_shape is not None else shape))
        self.initializer = initializer
        self.trainable = trainable
        self.kind = kind  # conv_w | fc_w | deconv_w | bias | generic
        self.dtype = dtype
        self.wd = wd
        self.master: Optional[torch.Tensor] = None  # fp32 (store_shape)
        self.grad: Optional[torch.Tensor] = None  # fp32 view into the gradient arena
        self.shadow: Optional[torch.Tensor] = None  # bf16 leaf (requires_grad) used by autograd Functions
        self.shadow_t: Optional[torch.Tensor] = None  # bf16, taps transposed
        self.offset = -1  # element offset in the arenas
        self.owner = 0
        self.index = -1

    @property
    def numel(self):
        n = 1
        for s in self.store_shape:
            n *= s
        return n

    def logical(self, t: torch.Tensor) -> torch.Tensor:
        """Slice a store_shape tensor down to the logical shape."""
        if self.store_shape == self.shape:
            return t
        idx = tuple(slice(0, s) for s in self.shape)
        return t[idx]

    def value(self) -> torch.Tensor:
        """The variable as a tensor for USER code written in plain torch ops (what `weight()` / `bias()` of the
        reference hand to the user, distribute_tools.py:211-224): the fp32 master on CPU; on CUDA the bf16 shadow
        behind a bridge whose backward adds the incoming gradient into this variable's slice of the gradient arena
        and announces it to the bucketed collectives, like the library's own layers do. Logical shape."""
        if self.shadow is None:
            return self.logical(self.master)
        if not self.trainable:
            return self.logical(self.shadow)
        from . import functional as Fn
        return self.logical(Fn.use_variable(self))

    def __repr__(self):
        return "Variable(%s, %s)" % (self.name, self.shape)


def _host_init(var: Variable, seed: int) -> torch.Tensor:
    g = torch.Generator().manual_seed((seed * 1000003 + var.index * 7919 + 17) & 0x7fffffff)
    a, b = var.initializer.params(var.shape)
    kind = var.initializer.kind
    t = torch.zeros(var.store_shape, dtype=torch.float32)
    lt = var.logical(t)
    if kind == "constant":
        lt.fill_(a)
    elif kind == "uniform":
        lt.copy_(torch.rand(var.shape, generator=g) * (b - a) + a)
    elif kind == "truncated_normal":
        v = torch.empty(var.shape)
        torch.nn.init.trunc_normal_(v, mean=a, std=b, a=a - 2 * b, b=a + 2 * b, generator=g)
        lt.copy_(v)
    else:
        lt.copy_(torch.randn(var.shape, generator=g) * b + a)
    return t


class VariableStore(object):
    """Holds every variable of one model replica on one device."""

    def __init__(self, device="cpu", seed=0, compute_dtype=torch.bfloat16):
        self.device = torch.device(device)
        self.seed = seed
        self.compute_dtype = compute_dtype
        self.vars: Dict[str, Variable] = {}
        self.order: List[Variable] = []
        self.finalized = False
        self.losses: List = []  # per-tower 'losses' collection: (name, tensor)
        self.grad_ready_hook: Optional[Callable[[Variable], None]] = None
        # arenas (CUDA path)
        self.master_arena = self.grad_arena = self.shadow_arena = self.shadow_t_arena = None
        self.slot_m = self.slot_v = None
        self.n_params = 0
        self.owner_bounds = [0, 0]
        self.f32_ranges = [(0, 0)]
        self._tap_table = None
        self.global_step = 0

    # ------------------------------------------------------------------ creation
    def get_variable(self, name, shape, initializer=None, trainable=True, kind="generic", store_shape=None,
                     wd=None) -> Variable:
        full = (current_scope() + "/" + name) if current_scope() else name
        v = self.vars.get(full)
        if v is not None:
            if tuple(shape) != v.shape:
                raise ValueError("variable %s exists with shape %s, requested %s" % (full, v.shape, tuple(shape)))
            return v
        if self.finalized and trainable:
            raise RuntimeError("cannot create trainable variable %s after the store was finalized" % full)
        v = Variable(full, shape, initializer or ConstantInitializer(0.0), trainable, kind, store_shape=store_shape,
                     wd=wd)
        v.index = len(self.order)
        self._materialize(v)
        self.vars[full] = v
        self.order.append(v)
        return v

    def _materialize(self, v: Variable):
        if self.device.type == "cuda":
            from ..ops import raw
            t = torch.zeros(v.store_shape, device=self.device, dtype=torch.float32)
            a, b = v.initializer.params(v.shape)
            if v.store_shape == v.shape:
                raw.init_(t, v.initializer.kind, a, b, seed=self.seed, stream_id=v.index)
            else:
                tmp = torch.empty(v.shape, device=self.device, dtype=torch.float32)
                raw.init_(tmp, v.initializer.kind, a, b, seed=self.seed, stream_id=v.index)
                v.logical(t).copy_(tmp)
            v.master = t
            if v.trainable:
                v.shadow = t.to(self.compute_dtype).requires_grad_(True)
                v.grad = torch.zeros_like(t)
                if v.kind in ("conv_w", "fc_w", "deconv_w"):
                    v.shadow_t = _tap_transpose_torch(v.shadow.detach(), v.store_shape)
        else:
            v.master = _host_init(v, self.seed).requires_grad_(v.trainable)

    def trainable_variables(self) -> List[Variable]:
        return [v for v in self.order if v.trainable]

    def all_variables(self) -> List[Variable]:
        return list(self.order)

    # ------------------------------------------------------------------ arenas
    def finalize(self, allocator: Optional[Callable[[str, int, torch.dtype], torch.Tensor]] = None,
                 placement: str = "bytes", world: int = 1, slots: int = 2):
        """Pack trainable variables into flat arenas. `allocator(tag, numel, dtype)` returns a zeroed flat tensor
        (e.g. carved out of the NVLink symmetric heap); default is a plain torch allocation.

        placement: 'bytes' keeps layer order (owner shards are equal byte ranges); 'round_robin' groups variables by
        owner = index % world, mirroring tf.train.replica_device_setter's round-robin over ps tasks
        (/root/reference/distribute_train.py:109-110), so each owner's shard is still one contiguous range."""
        if self.finalized or self.device.type != "cuda":
            self.finalized = True
            return
        tv = self.trainable_variables()
        # Variables that kernels read in fp32 straight from the master arena (biases, batch-norm gamma / beta, user
        # `weight()` / `bias()` variables) are packed in front of the tensor-core operands (conv / FC / deconv filters,
        # read through the bf16 shadow), so that in the owner-sharded modes the "fp32 region" an owner must broadcast
        # after its update is ONE contiguous range per owner (f32_ranges[owner]).
        small = lambda v: v.kind not in ("conv_w", "fc_w", "deconv_w")  # noqa: E731
        rr = placement == "round_robin" and world > 1
        if rr:
            for i, v in enumerate(tv):
                v.owner = i % world
            tv = sorted(tv, key=lambda v: (v.owner, 0 if small(v) else 1, v.index))
        else:
            tv = sorted(tv, key=lambda v: (0 if small(v) else 1, v.index))
        off = 0
        self.owner_bounds = [0] * (world + 1)
        cur_owner = 0
        f32_lo, f32_hi = {}, {}
        for v in tv:
            if rr:
                while cur_owner < v.owner:
                    cur_owner += 1
                    self.owner_bounds[cur_owner] = off
            v.offset = off
            # every variable starts on a 128-byte line of the bf16 arenas (256 bytes in the fp32 ones): TMA tiles of a
            # filter and the red.add rows of its gradient then never straddle lines (an odd-sized variable in front,
            # e.g. 1000 logits biases, used to shift every later filter by 80 bytes: 1x1-conv GEMMs ran 20 % slower)
            off += padded_numel(v.numel)
            if small(v):
                key = v.owner if rr else 0
                f32_lo.setdefault(key, v.offset)
                f32_hi[key] = off
        n = off
        if rr:
            for o in range(cur_owner + 1, world + 1):
                self.owner_bounds[o] = n
            self.f32_ranges = [(f32_lo.get(o, 0), f32_hi.get(o, 0)) for o in range(world)]
        else:
            per = ((n + world - 1) // world + 7) // 8 * 8
            self.owner_bounds = [min(i * per, n) for i in range(world)] + [n]
            self.f32_ranges = [(f32_lo.get(0, 0), f32_hi.get(0, 0))] * max(world, 1)
        self.n_params = n
        alloc = allocator or (lambda tag, numel, dtype: torch.zeros(numel, device=self.device, dtype=dtype))
        self.master_arena = alloc("master", n, torch.float32)
        self.grad_arena = alloc("grad", n, torch.float32)
        self.shadow_arena = alloc("shadow", n, self.compute_dtype)
        self.shadow_t_arena = alloc("shadow_t", n, self.compute_dtype)
        self.slot_m = alloc("slot_m", n, torch.float32) if slots >= 1 else None
        self.slot_v = alloc("slot_v", n, torch.float32) if slots >= 2 else None
        descs = []
        tile_start = 0
        for v in tv:
            sl = slice(v.offset, v.offset + v.numel)
            self.master_arena[sl].copy_(v.master.reshape(-1))
            v.master = self.master_arena[sl].view(v.store_shape)
            v.grad = self.grad_arena[sl].view(v.store_shape)
            self.shadow_arena[sl].copy_(v.master.reshape(-1))
            v.shadow = self.shadow_arena[sl].view(v.store_shape).detach().requires_grad_(True)
            if v.kind in ("conv_w", "fc_w", "deconv_w"):
                T, Ci, Co = _taps(v.store_shape)
                v.shadow_t = self.shadow_t_arena[sl].view(v.store_shape[:-2] + (Co, Ci))
                descs.append((v.offset, v.offset, T, Ci, Co, tile_start))
                tile_start += T * ((Ci + 31) // 32) * ((Co + 31) // 32)
        if not (placement == "round_robin" and world > 1):
            for v in tv:  # owner = the shard that holds the variable's first element
                v.owner = max(i for i in range(world) if self.owner_bounds[i] <= v.offset)
        self._build_tap_table(descs, tile_start)
        self.refresh_transposed_shadows()
        self.finalized = True

    def _build_tap_table(self, descs, total_tiles):
        import struct
        self._tap_tiles = total_tiles
        self._tap_n = len(descs)
        if not descs:
            self._tap_table = None
            return
        blob = b"".join(struct.pack("<qqiiii", *d) for d in descs)
        self._tap_table = torch.frombuffer(bytearray(blob), dtype=torch.uint8).to(self.device)

    def refresh_transposed_shadows(self):
        """shadow_t[tap] = shadow[tap]^T for every filter (one batched kernel)."""
        if self._tap_table is None:
            return
        from ..ops import _lib
        _lib.call("dtf_tap_transpose", self.shadow_arena, self.shadow_t_arena, self._tap_table, self._tap_n,
                  self._tap_tiles, _lib.stream_ptr())

    def rebind_shadow_leaves(self):
        """Re-create the autograd leaves of the bf16 shadows (same storage, new grad accumulators)."""
        for v in self.trainable_variables():
            if v.shadow is not None:
                v.shadow = v.shadow.detach().requires_grad_(True)

    def zero_grads(self):
        self._bridge_uses = {}  # outstanding Variable.value() applications of this step (engine/functional.py)
        if self.grad_arena is not None:
            from ..ops import _lib
            _lib.call("dtf_fill_u32", self.grad_arena, self.grad_arena.numel(), 0, _lib.stream_ptr())
        else:
            for v in self.trainable_variables():
                if self.device.type == "cuda":
                    v.grad.zero_()
                elif v.master.grad is not None:
                    v.master.grad = None

    def notify_grad_ready(self, v: Variable):
        if self.grad_ready_hook is not None:
            self.grad_ready_hook(v)

    # ------------------------------------------------------------------ state dict (checkpoint keys = names)
    def state_dict(self) -> Dict[str, torch.Tensor]:
        out = {}
        for v in self.order:
            out[v.name] = v.logical(v.master.detach()).cpu().contiguous()
        return out

    def load_state_dict(self, state: Dict[str, torch.Tensor], strict=False):
        missing = []
        with torch.no_grad():
            for v in self.order:
                if v.name not in state:
                    missing.append(v.name)
                    continue
                src = torch.as_tensor(state[v.name])
                if [d for d in src.shape if d != 1] != [d for d in v.shape if d != 1]:
                    # (TF: "Assign requires shapes of both tensors to match")
                    raise ValueError("variable %r: the checkpoint holds shape %s, the model declares %s"
                                     % (v.name, tuple(src.shape), tuple(v.shape)))
                src = src.to(torch.float32).reshape(v.shape)
                v.logical(v.master).copy_(src.to(v.master.device))
                if v.shadow is not None:
                    v.shadow.data.copy_(v.master)
        if self.device.type == "cuda":
            if self.finalized:
                self.refresh_transposed_shadows()
            else:
                for v in self.order:
                    if v.shadow_t is not None:
                        v.shadow_t = _tap_transpose_torch(v.shadow.detach(), v.store_shape)
        if strict and missing:
            raise KeyError("missing variables in checkpoint: %s" % missing)
        return missing


ARENA_ALIGN = 64  # elements


def padded_numel(n):
    return (n + ARENA_ALIGN - 1) // ARENA_ALIGN * ARENA_ALIGN


def _taps(shape):
    T = 1
    for d in shape[:-2]:
        T *= d
    return T, shape[-2], shape[-1]


def _tap_transpose_torch(t, shape):
    T, Ci, Co = _taps(shape)
    return t.reshape(T, Ci, Co).transpose(1, 2).contiguous().view(tuple(shape[:-2]) + (Co, Ci))


# ---- default store (the reference has one default graph); swappable for tests / multiple replicas
_default_store: Optional[VariableStore] = None


def default_store() -> VariableStore:
    global _default_store
    if _default_store is None:
        _default_store = VariableStore("cpu")
    return _default_store


def set_default_store(store: Optional[VariableStore]):
    global _default_store
    _default_store = store
    return store


@contextlib.contextmanager
def using_store(store: VariableStore):
    global _default_store
    prev = _default_store
    _default_store = store
    try:
        yield store
    finally:
        _default_store = prev
