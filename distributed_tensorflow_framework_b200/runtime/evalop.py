"""Eval operator. The reference's is a stub (/root/reference/distribute_eval.py:8-23: `eval(...)` is `pass`); this
is the real implementation with the same constructor, method names and advice cut-points.

Flow: restore the latest checkpoint from model_dir (TF1 Saver layout) -> pre_eval_fn -> for every batch of
`data_loader.load_eval_batch()`: pre_process_fn -> net.process (inference statistics for batch norm) ->
post_process_fn -> loss and, for classifiers, top-1 accuracy -> post_eval_fn(results). Results are also written as
TensorBoard scalars into model_dir/eval.
"""
import os

import torch

from ..api import log as logger
from ..engine import layers, variables as V
from ..io import tensorbundle as tb
from ..io.events import EventFileWriter
from .session import Saver


class Eval(object):
    def __init__(self, data_loader=None, input_mode=None):
        self.data_loader = data_loader
        self.input_mode = input_mode

    def eval(self, pre_eval_fn=None, post_eval_fn=None, pre_process_fn=None, post_process_fn=None, *args, **kwargs):
        use_cuda = getattr(self, "gpu_num", 0) > 0 and torch.cuda.is_available()
        device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0))) if use_cuda else torch.device("cpu")
        store = getattr(self, "store", None) or V.VariableStore(device)
        pre_result = pre_eval_fn(args, kwargs) if pre_eval_fn is not None else None
        total_loss, n_batches, hits, seen = 0.0, 0, 0, 0
        restored_step = None
        # launched on several ranks (torchrun): every rank restores the checkpoint and evaluates every world-th batch
        # of the stream; the sums are combined at the end, so the result is that of one pass over the data
        rank, world = 0, 1
        try:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized():
                rank, world = dist.get_rank(), dist.get_world_size()
        except Exception:  # pragma: no cover
            pass
        seq = -1
        layers.set_training(False)
        try:
            with V.using_store(store), torch.no_grad():
                while True:
                    batch = self.data_loader.load_eval_batch()
                    if batch is None:
                        break
                    seq += 1
                    if world > 1 and seq % world != rank and restored_step is not None:
                        continue
                    raw, gt = batch
                    raw = raw.to(device, non_blocking=True)
                    gt = gt.to(device, non_blocking=True) if isinstance(gt, torch.Tensor) else gt
                    if pre_process_fn is not None:
                        raw, gt = pre_process_fn(raw, gt, args, kwargs)
                    logits = self.net.process(raw)
                    if restored_step is None:
                        # variables exist after the first forward: now restore them (init_or_restore semantics)
                        restored_step = self._restore(store)
                        logits = self.net.process(raw)
                    if world > 1 and seq % world != rank:
                        continue  # (the first batch is run by every rank only to create and restore the variables)
                    if post_process_fn is not None:
                        logits = post_process_fn(logits)
                    if getattr(self, "loss", None) is not None and gt is not None:
                        total_loss += float(self.loss.loss(logits, gt))
                    if gt is not None and gt.dtype in (torch.int64, torch.int32) and logits.dim() == 2:
                        if logits.is_cuda:
                            from ..ops import raw as rawops
                            hits += int(rawops.top1_hits(logits, gt.reshape(-1).long()).item())
                        else:
                            hits += int((logits.argmax(1) == gt.reshape(-1)).sum())
                        seen += logits.shape[0]
                    n_batches += 1
        finally:
            layers.set_training(True)
        if world > 1:
            import torch.distributed as dist
            t = torch.tensor([total_loss, float(n_batches), float(hits), float(seen)], dtype=torch.float64,
                             device=device if device.type == "cuda" else "cpu")
            dist.all_reduce(t)
            total_loss, n_batches, hits, seen = float(t[0]), int(t[1]), int(t[2]), int(t[3])
        results = {"global_step": restored_step or 0, "batches": n_batches,
                   "loss": total_loss / max(n_batches, 1)}
        if seen:
            results["accuracy"] = hits / float(seen)
        logger.info("eval @ step %d: %s", results["global_step"],
                    ", ".join("%s = %.6g" % (k, v) for k, v in results.items() if k != "global_step"))
        model_dir = getattr(self, "model_dir", None)
        if model_dir and os.path.isdir(model_dir):
            w = EventFileWriter(os.path.join(model_dir, "eval"))
            for k, v in results.items():
                if k not in ("global_step", "batches"):
                    w.add_scalar(k, v, results["global_step"])
            w.close()
        self.results = results
        if post_eval_fn is not None:
            post_eval_fn(results, args, kwargs) if _accepts(post_eval_fn, 3) else post_eval_fn(args, kwargs)
        return results

    def _restore(self, store):
        model_dir = getattr(self, "model_dir", None)
        if not model_dir:
            return 0
        prefix = tb.latest_checkpoint(model_dir)
        if prefix is None:
            logger.warn("no checkpoint in %s: evaluating freshly initialised variables", model_dir)
            return 0
        return Saver(store).restore(prefix)

    def run(self):
        g = lambda n: getattr(self, n, None)  # noqa: E731
        return self.eval(pre_eval_fn=g("pre_eval_fn") or g("pre_fn"), post_eval_fn=g("post_eval_fn") or g("post_fn"),
                         pre_process_fn=g("pre_process_fn"), post_process_fn=g("post_process_fn"))


def _accepts(fn, n):
    import inspect
    try:
        params = inspect.signature(fn).parameters.values()
    except (TypeError, ValueError):
        return False
    pos = [p for p in params if p.kind in (p.POSITIONAL_ONLY, p.POSITIONAL_OR_KEYWORD)]
    return len(pos) >= n or any(p.kind == p.VAR_POSITIONAL for p in params)
