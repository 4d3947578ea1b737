"""Tower / replica engine: per-device forward + loss + gradients, cross-tower gradient mean.

Parity: /root/reference/distribute_tower.py:16-148 (`Tower.__init__`, `get_gradient`, `tower_loss`, `process`,
`loss_to_scope`, `average_gradients`, `tower_fn`). Fixes: the unused add_n over ALL towers' losses in
`loss_to_scope` is dropped. On the B200 path "the towers" are the ranks (one process per GPU) and
`average_gradients` is replaced by the fused peer-memory allreduce (parallel/comm.py); the literal list-of-lists
version below is the API-compatible reference implementation (CPU oracle, in-process towers, B-LITERAL baseline).
"""
import re

import torch

from ..api import constants
from . import variables as V


class Tower(object):
    def __init__(self, net, scope, tower_grades, raw_data, ground_truth, loss, optimizer, store=None):
        self.net = net
        self.scope = scope
        self.tower_grades = tower_grades
        self.raw_data = raw_data
        self.ground_truth = ground_truth
        self.loss = loss
        self.optimizer = optimizer
        self.store = store
        self.summaries = {}

    def get_gradient(self, loss):
        return self.optimizer.compute_gradients(loss, store=self._store())

    def _store(self):
        return self.store or V.default_store()

    def tower_loss(self, post_process_fn=None, *args, **kwargs):
        return self.__tower_loss(post_process_fn, *args, **kwargs)

    def process(self, post_process_fn=None, *args, **kwargs):
        """Forward this tower, compute its loss and gradients; returns (summaries, loss, logits).
        Variables are shared across towers by name (tower 0 creates them, later towers reuse them)."""
        loss, logits = self.tower_loss(post_process_fn, *args, **kwargs)
        grads = self.get_gradient(loss)
        self.tower_grades.append(grads)
        return self.summaries, loss, logits

    def __loss(self, result):
        return self.loss.loss(result, self.ground_truth)

    def loss_to_scope(self, result):
        loss = self.__loss(result)
        self._store().losses.append(("loss", loss))
        return loss

    @staticmethod
    def average_gradients(tower_grads):
        """[[(g, v), ...] per tower] -> [(mean_g, v)]: the synchronisation point across towers
        (distribute_tower.py:78-114: expand_dims + concat + reduce_mean, first tower's variable returned)."""
        average_grads = []
        for grad_and_vars in zip(*tower_grads):
            grads = [g.unsqueeze(0) for g, _ in grad_and_vars]
            grad = torch.cat(grads, dim=0).mean(dim=0)
            average_grads.append((grad, grad_and_vars[0][1]))
        return average_grads

    def __tower_loss(self, post_process_fn=None, *args, **kwargs):
        store = self._store()
        store.losses = []
        logits = self.net.process(self.raw_data, *args, **kwargs)
        logits = logits if post_process_fn is None else post_process_fn(logits)
        self.loss_to_scope(logits)
        losses = list(store.losses)
        total_loss = losses[0][1]
        for _, l in losses[1:]:
            total_loss = total_loss + l.to(total_loss.dtype)
        # scalar summaries per loss term and for the total, tower prefix stripped (distribute_tower.py:137-143)
        self.summaries = {}
        for name, l in losses + [("total_loss", total_loss)]:
            clean = re.sub("%s_[0-9]*/" % constants.TOWER_NAME, "", name)
            self.summaries[clean] = l.detach()
        return total_loss, logits

    @staticmethod
    def tower_fn(tower):
        return tower.process()
