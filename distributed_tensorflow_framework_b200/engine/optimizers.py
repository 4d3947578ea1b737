"""Optimizers with the tf.train.Optimizer calling convention used by the reference
(`compute_gradients(loss)` at /root/reference/distribute_tower.py:26-27, `apply_gradients(grads, global_step)` at
/root/reference/distribute_train.py:158; default AdamOptimizer(0.001) at /root/reference/distribute.py:26,112).

CUDA path: one fused kernel over the flat arenas (fp32 master + slots, bf16 shadow written in the same pass); when a
communicator is attached the same update is fused into the gradient collective (parallel/comm.py). CPU path: plain
torch math (oracle)."""
import math
from typing import List, Optional, Tuple

import torch

from . import variables as V


def _lr_value(lr):
    if callable(lr):
        return float(lr())
    if hasattr(lr, "learning_rate"):
        return float(lr.learning_rate)
    return float(lr)


class Optimizer(object):
    kind = "sgd"
    slot_names: Tuple[str, ...] = ()

    def __init__(self, learning_rate, name=None):
        self._lr = learning_rate
        self.name = name or type(self).__name__
        self._slots = {}  # CPU path: var.name -> [tensors]
        self.b1, self.b2, self.eps, self.wd = 0.9, 0.999, 1e-8, 0.0

    @property
    def learning_rate(self):
        return _lr_value(self._lr)

    def num_slots(self):
        return len(self.slot_names)

    # ------------------------------------------------------------------ gradients
    def compute_gradients(self, loss, var_list: Optional[List[V.Variable]] = None, store=None, retain_graph=False):
        """Backward pass. Returns [(grad, var)] in variable-creation order (like TF's trainable_variables order)."""
        store = store or V.default_store()
        vs = var_list if var_list is not None else store.trainable_variables()
        if store.device.type == "cuda":
            loss.backward(retain_graph=retain_graph)
            return [(v.grad, v) for v in vs]
        for v in vs:
            v.master.grad = None
        loss.backward(retain_graph=retain_graph)
        out = []
        for v in vs:
            g = v.master.grad if v.master.grad is not None else torch.zeros_like(v.master)
            out.append((g.detach().clone(), v))
        return out

    # ------------------------------------------------------------------ update
    def apply_gradients(self, grads_and_vars, global_step=None, gscale=1.0, store=None):
        store = store or V.default_store()
        step = (store.global_step + 1) if global_step is None else int(global_step) + 1
        if store.device.type == "cuda" and store.finalized and store.grad_arena is not None:
            self.apply_arena(store, step, gscale)
        else:
            for g, v in grads_and_vars:
                self._apply_one(v, g, step, gscale)
        store.global_step = step
        return step

    def apply_arena(self, store, step, gscale=1.0, start=0, count=None, refresh=True, hyper=None):
        """`hyper`: optional device tensor [lr, lr_t, step, -] read by the kernel instead of the by-value lr / step
        (learning-rate schedules and Adam under CUDA-graph replay)."""
        from ..ops import raw
        n = store.n_params if count is None else count
        sl = slice(start, start + n)
        raw.optim_step(self.kind, store.master_arena[sl], store.grad_arena[sl],
                       store.slot_m[sl] if store.slot_m is not None else None,
                       store.slot_v[sl] if store.slot_v is not None else None, store.shadow_arena[sl],
                       self.learning_rate, self.b1, self.b2, self.eps, self.wd, gscale, step, hyper=hyper)
        if refresh:
            store.refresh_transposed_shadows()

    def _apply_one(self, v, g, step, gscale):
        lr = self.learning_rate
        with torch.no_grad():
            p = v.master
            g = g.to(p.device, torch.float32).reshape(p.shape) * gscale
            if self.wd:
                g = g + self.wd * p
            slots = self._slots.setdefault(v.name, [torch.zeros_like(p) for _ in self.slot_names])
            if self.kind == "sgd":
                p.sub_(lr * g)
            elif self.kind == "momentum":
                slots[0].mul_(self.b1).add_(g)
                p.sub_(lr * slots[0])
            else:
                m, vv = slots
                m.mul_(self.b1).add_((1 - self.b1) * g)
                vv.mul_(self.b2).add_((1 - self.b2) * g * g)
                lr_t = lr * math.sqrt(1 - self.b2 ** step) / (1 - self.b1 ** step)
                p.sub_(lr_t * m / (vv.sqrt() + self.eps))
            if v.shadow is not None:
                v.shadow.data.copy_(p)
                if v.shadow_t is not None:
                    v.shadow_t.copy_(V._tap_transpose_torch(v.shadow.detach(), v.store_shape))

    # ------------------------------------------------------------------ slots for checkpoints (TF names)
    def slot_tensors(self, store):
        """name -> cpu tensor, using TF's slot naming ("<var>/Adam", "<var>/Adam_1", "<var>/Momentum")."""
        out = {}
        for v in store.trainable_variables():
            for i, sname in enumerate(self.slot_names):
                if store.device.type == "cuda" and store.finalized and store.slot_m is not None:
                    arena = store.slot_m if i == 0 else store.slot_v
                    t = arena[v.offset:v.offset + v.numel].view(v.store_shape)
                else:
                    sl = self._slots.get(v.name)
                    t = sl[i] if sl is not None else torch.zeros_like(v.master)
                out[v.name + "/" + sname] = v.logical(t.detach()).cpu().contiguous()
        if self.kind == "adam":
            out["beta1_power"] = torch.tensor(self.b1 ** max(store.global_step, 0) * self.b1, dtype=torch.float32)
            out["beta2_power"] = torch.tensor(self.b2 ** max(store.global_step, 0) * self.b2, dtype=torch.float32)
        return out

    def load_slot_tensors(self, store, state):
        for v in store.trainable_variables():
            for i, sname in enumerate(self.slot_names):
                key = v.name + "/" + sname
                if key not in state:
                    continue
                src = torch.as_tensor(state[key]).to(torch.float32).reshape(v.shape)
                if store.device.type == "cuda" and store.finalized and store.slot_m is not None:
                    arena = store.slot_m if i == 0 else store.slot_v
                    v.logical(arena[v.offset:v.offset + v.numel].view(v.store_shape)).copy_(src.to(arena.device))
                else:
                    sl = self._slots.setdefault(v.name, [torch.zeros_like(v.master) for _ in self.slot_names])
                    v.logical(sl[i]).copy_(src)


class GradientDescentOptimizer(Optimizer):
    kind = "sgd"

    def __init__(self, learning_rate, name="GradientDescent"):
        super().__init__(learning_rate, name)


class MomentumOptimizer(Optimizer):
    kind = "momentum"
    slot_names = ("Momentum",)

    def __init__(self, learning_rate, momentum=0.9, name="Momentum"):
        super().__init__(learning_rate, name)
        self.b1 = momentum


class AdamOptimizer(Optimizer):
    kind = "adam"
    slot_names = ("Adam", "Adam_1")

    def __init__(self, learning_rate=0.001, beta1=0.9, beta2=0.999, epsilon=1e-8, name="Adam"):
        super().__init__(learning_rate, name)
        self.b1, self.b2, self.eps = beta1, beta2, epsilon
