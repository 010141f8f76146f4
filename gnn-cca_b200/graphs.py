"""CUDA-graph replay of the training step for batches of a repeating shape.

A training step on EPFL-sized batches is ~45 short kernels (forward, loss, backward): launch-bound.  ``GraphedTrainStep``
captures forward + loss (``loss_stats``, train.py:80-97) + backward of ONE batch shape into a CUDA graph and replays it
for every later batch with the same (nodes, edges, frames) signature; inputs are copied into the graph's static
buffers, every replay leaves the parameters' ``.grad`` pointing at the graph's freshly written gradient tensors, so
any ``torch.optim`` optimiser steps on them as usual.  Batches of another shape get their own graph (up to
``max_graphs``), after that they run eagerly.  The kernels are the same ones the eager path launches; nothing is traced or compiled.

    step = gnncca.GraphedTrainStep(mpn_model)                       # criteria of main_training.py:258-268 via kwargs
    for data_batch in loader:                                       # batches from gnncca.build_graph_batch
        loss = step(data_batch)                                     # replaces train.py:454-458 + loss.backward() (:493)
        optimizer.step()                                            # no optimizer.zero_grad(): .grad is overwritten, not accumulated
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch

from . import _lib
from .graph import Batch
from .loss import loss_stats

_STATIC = ("x", "edge_index", "edge_attr", "edge_labels", "ptr", "edge_ptr")


class _Captured:
    def __init__(self, graph, batch, loss, stats, launches):
        self.graph, self.batch, self.loss, self.stats, self.launches = graph, batch, loss, stats, launches


class GraphedTrainStep:
    def __init__(self, model, pos_weight: float = 1.0, focal: bool = False, focus: float = 5.0, balance: float = 0.9,
                 max_graphs: int = 8, warmup: int = 3):
        self.model, self.max_graphs, self.warmup = model, max_graphs, warmup
        self.loss_args = dict(pos_weight=pos_weight, focal=focal, focus=focus, balance=balance)
        self.graphs: Dict[Tuple, _Captured] = {}
        self.replays = 0
        self.eager = 0
        self.launches_replayed = 0      # kernels of libgnncca.so launched through graph replays (each replay = its capture's count)

    # ------------------------------------------------------------------------------------------------ eager step
    def _forward_backward(self, batch):
        out = self.model(batch)["classified_edges"]
        st = loss_stats(out, batch.edge_labels, **self.loss_args)
        st.loss.backward()
        return st

    def _eager(self, batch):
        for p in self.model.parameters():
            if p.grad is not None:
                p.grad.zero_()
        self.eager += 1
        st = self._forward_backward(batch)
        self.last_stats = st
        return st.loss.detach()

    @staticmethod
    def _signature(batch) -> Optional[Tuple]:
        if not getattr(batch, "gnncca_trusted", False) or getattr(batch, "edge_ptr", None) is None or getattr(batch, "ptr", None) is None:
            return None          # only batches of build_graph_batch are replayed: nothing about them has to be read back
        return (tuple(batch.x.shape), tuple(batch.edge_index.shape), tuple(batch.edge_attr.shape), int(batch.ptr.numel()),
                tuple(getattr(batch, "gnncca_caps", None) or ()), batch.x.device.index)

    # ------------------------------------------------------------------------------------------------ capture
    def _capture(self, batch, sig) -> _Captured:
        model = self.model
        if not model.training:
            raise RuntimeError("GraphedTrainStep is for model.train() (use model(data) for inference)")
        static = Batch(**{k: (v.clone() if torch.is_tensor(v) else v) for k, v in batch.__dict__.items()})
        # BatchNorm buffers (classifier of the BNcls configurations): the rehearsals below must not count as training steps
        bn_state = [(b, b.detach().clone()) for m in model.modules() if isinstance(m, torch.nn.modules.batchnorm._BatchNorm)
                    for b in m.buffers(recurse=False)]
        # warm-up on a side stream (allocator pools, lazily configured kernels)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(self.warmup):
                for p in model.parameters():
                    if p.grad is not None:
                        p.grad.zero_()
                self._forward_backward(static)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        # the CSR of the batch (and its transpose for the backward) must be rebuilt by every replay: edge_index changes
        # in place under the graph, so no plan cached by the rehearsals may be reused
        static.__dict__.pop("gnncca_plan", None)
        g = torch.cuda.CUDAGraph()
        lib = _lib.load()
        n0 = int(lib.gnncca_launch_count())
        with torch.cuda.graph(g):
            # .grad = None: autograd then ADOPTS the gradient tensors the backward produced (memory of the graph's own pool, so
            # the same addresses at every replay) instead of zero-filling and adding into older ones: 2 launches less per parameter
            for p in model.parameters():
                p.grad = None
            st = self._forward_backward(static)
        for b, saved in bn_state:
            b.copy_(saved)
        cap = _Captured(g, static, st.loss.detach(), st, int(lib.gnncca_launch_count()) - n0)
        cap.grads = [p.grad for p in model.parameters()]
        self.graphs[sig] = cap
        return cap

    # ------------------------------------------------------------------------------------------------ call
    def __call__(self, batch) -> torch.Tensor:
        """Forward + loss + backward of `batch`; returns the loss (a device tensor, overwritten by the next call of the
        same shape).  ``self.last_stats`` holds the ``LossStats`` of the step (precisions, per-class losses)."""
        sig = self._signature(batch)
        if sig is None:
            return self._eager(batch)
        cap = self.graphs.get(sig)
        if cap is None:
            if len(self.graphs) >= self.max_graphs:
                return self._eager(batch)
            cap = self._capture(batch, sig)
        for k in _STATIC:
            src, dst = getattr(batch, k), getattr(cap.batch, k)
            if src.data_ptr() != dst.data_ptr():
                dst.copy_(src, non_blocking=True)
        cap.graph.replay()
        for p, g in zip(self.model.parameters(), cap.grads):       # this shape's gradient buffers (each graph owns a set)
            p.grad = g
        self.replays += 1
        self.launches_replayed += cap.launches
        cap.stats._host = None          # counts of the previous replay
        self.last_stats = cap.stats
        return cap.loss

    def static_batch(self, like) -> Optional[Batch]:
        """The graph's own input buffers for batches shaped like `like` (fill them in place to skip the copies)."""
        cap = self.graphs.get(self._signature(like))
        return None if cap is None else cap.batch
