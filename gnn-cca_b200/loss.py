"""Loss and statistics of the classified steps: the reference's ``compute_loss_acc`` (train.py:51-133) on the device.

``compute_loss_acc(outputs, batch, criterion, criterion_no_reduction, mode)`` keeps the reference's signature, return
tuple and criteria (main_training.py:258-268: ``BCE``, ``BCE_weighted`` with a scalar ``pos_weight``, ``Focal`` =
libs/utils.py:264-293).  The reference runs ~10 eager ops per classified step and >= 8 synchronising ``.cpu()`` copies
per batch; here one kernel reads every step's logits once (``gnncca_bce_stats_fwd``), a one-block kernel folds the
per-block sums in a fixed order, the backward seed is one kernel (``gnncca_bce_stats_bwd``), and the precisions come
back in ONE small device->host copy.  ``loss_stats`` is the non-synchronising form (everything stays on the device).
There is no CPU or eager fallback.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import torch

from . import _lib

PER = 5
MAX_STEPS = 8


def _ptr_array(tensors: Sequence[torch.Tensor]):
    return (C.c_void_p * len(tensors))(*[t.data_ptr() for t in tensors])


def _flat(t: torch.Tensor) -> torch.Tensor:
    t = t.detach().reshape(-1)
    if t.dtype != torch.float32:
        t = t.float()
    return t if t.is_contiguous() else t.contiguous()


class _BceStats(torch.autograd.Function):
    """stats[5k + 4] (float64) of k logit vectors; differentiable in its first k entries (the per-step mean losses)."""

    @staticmethod
    def forward(ctx, labels, pos_weight, focal, focus, balance, probs, *logits):
        lib = _lib.load()
        k = len(logits)
        flat = [_flat(x) for x in logits]
        E = flat[0].numel()
        dev = flat[0].device
        if not flat[0].is_cuda:
            raise RuntimeError("gnncca loss kernels need CUDA tensors (there is no CPU path)")
        for x in flat:
            if x.numel() != E or x.device != dev:
                raise ValueError("every classified step must hold one logit per edge on the same device")
        if labels.numel() != E:
            raise ValueError("edge_labels must hold one label per edge")
        blocks = lib.gnncca_loss_blocks(E)
        width = PER * k + 4
        scratch = torch.empty((blocks + 1) * width, dtype=torch.float64, device=dev)
        out = scratch[blocks * width:]
        # the summed loss of the BCE criteria comes out of the same launch as a float32 scalar, and its gradient goes straight
        # into the backward kernel: no reduction / cast / expand kernels between the statistics and the seed of the MPN's backward
        loss_sum = torch.empty((), dtype=torch.float32, device=dev)
        stream = torch.cuda.current_stream(dev).cuda_stream
        _lib.check(lib.gnncca_bce_stats_fwd(_ptr_array(flat), k, labels.data_ptr(), E, pos_weight, int(focal), focus, balance,
                                            _lib.ptr(probs), scratch.data_ptr(), out.data_ptr(), loss_sum.data_ptr(), stream), "gnncca_bce_stats_fwd")
        ctx.flat, ctx.labels, ctx.pos_weight, ctx.shapes = flat, labels, pos_weight, [x.shape for x in logits]
        ctx.set_materialize_grads(False)
        return out, loss_sum

    @staticmethod
    def backward(ctx, g_out, g_sum):
        lib = _lib.load()
        flat, labels = ctx.flat, ctx.labels
        k, E = len(flat), flat[0].numel()
        dev = flat[0].device
        if g_out is None and g_sum is None:
            return (None,) * (6 + k)
        up = None
        if g_out is not None:
            up = g_out.detach()
            if up.dtype != torch.float64 or not up.is_contiguous():
                up = up.double().contiguous()
        us = None
        if g_sum is not None:
            us = g_sum.detach()
            if us.dtype != torch.float32:
                us = us.float()
        grads = torch.empty((k, E), dtype=torch.float32, device=dev)
        rows = [grads[s] for s in range(k)]
        stream = torch.cuda.current_stream(dev).cuda_stream
        _lib.check(lib.gnncca_bce_stats_bwd(_ptr_array(flat), _ptr_array(rows), k, labels.data_ptr(), E, ctx.pos_weight,
                                            _lib.ptr(up), _lib.ptr(us), stream), "gnncca_bce_stats_bwd")
        return (None,) * 6 + tuple(rows[s].view(ctx.shapes[s]) for s in range(k))


class LossStats:
    """Device-resident result of ``loss_stats`` (no synchronisation until a ``precision*``/``counts`` property is read)."""

    def __init__(self, stats: torch.Tensor, k: int, probs: Optional[torch.Tensor], focal, focus, balance, loss_sum=None):
        self.stats, self.k, self.probs = stats, k, probs
        if focal:   # FocalLoss_binary(reduction='mean'): the focal transform of the MEAN cross entropy (utils.py:281-291)
            m = stats[:k]
            m = balance * (1.0 - torch.exp(-m)) ** focus * m
            self.loss = m.sum().float()
        else:
            self.loss = loss_sum        # sum of the k step means, from the statistics launch itself
        self._host = None

    # per-class losses and mean probabilities are read by compute_loss_acc / logging only: nothing is launched for them in a
    # training step that just needs ``loss``
    @property
    def loss_class1(self):
        return self.stats.detach()[self.k:2 * self.k].sum().float()

    @property
    def loss_class0(self):
        return self.stats.detach()[2 * self.k:3 * self.k].sum().float()

    @property
    def mean_prob_class1(self):
        return self.stats.detach()[3 * self.k:4 * self.k]

    @property
    def mean_prob_class0(self):
        return self.stats.detach()[4 * self.k:5 * self.k]

    def counts(self):
        """(n1, n0, correct1, correct0) as Python ints: the one device->host copy."""
        if self._host is None:
            self._host = [int(v) for v in self.stats.detach()[PER * self.k:].tolist()]
        return self._host

    @staticmethod
    def _pct(ok: int, n: int) -> float:
        return 0 if ok == 0 else ok / n * 100.0      # train.py:108-111

    @property
    def precision_class1(self):
        n1, _, c1, _ = self.counts()
        return self._pct(c1, n1)

    @property
    def precision_class0(self):
        _, n0, _, c0 = self.counts()
        return self._pct(c0, n0)

    @property
    def precision_all(self):
        n1, n0, c1, c0 = self.counts()
        return self._pct(c1 + c0, n1 + n0)


def _criterion_args(criterion, criterion_no_reduction, mode):
    """(pos_weight, focal, focus, balance) of the reference's three criteria; anything else is refused."""
    if mode != "train":
        return 1.0, False, 0.0, 1.0      # validation: plain F.binary_cross_entropy_with_logits (train.py:89-94)
    for c in (criterion, criterion_no_reduction):
        if c is None or isinstance(c, str):
            raise ValueError("mode='train' needs the criteria of main_training.py:258-268")
    if hasattr(criterion, "focusing_param"):
        if not hasattr(criterion_no_reduction, "focusing_param"):
            raise NotImplementedError("Focal criterion with a non-focal per-sample criterion")
        if (criterion.focusing_param, criterion.balance_param) != (criterion_no_reduction.focusing_param, criterion_no_reduction.balance_param):
            raise NotImplementedError("the two focal criteria must share their parameters")
        for c in (criterion, criterion_no_reduction):
            if getattr(c.cross_entropy, "pos_weight", None) is not None or getattr(c.cross_entropy, "weight", None) is not None:
                raise NotImplementedError("focal loss over a weighted cross entropy")
        if criterion.cross_entropy.reduction != "mean" or criterion_no_reduction.cross_entropy.reduction != "none":
            raise NotImplementedError("criterion must reduce by 'mean', criterion_no_reduction by 'none'")
        return 1.0, True, float(criterion.focusing_param), float(criterion.balance_param)
    if isinstance(criterion, torch.nn.BCEWithLogitsLoss) and isinstance(criterion_no_reduction, torch.nn.BCEWithLogitsLoss):
        if criterion.reduction != "mean" or criterion_no_reduction.reduction != "none":
            raise NotImplementedError("criterion must reduce by 'mean', criterion_no_reduction by 'none'")
        if criterion.weight is not None or criterion_no_reduction.weight is not None:
            raise NotImplementedError("per-sample `weight` is not used by the reference's criteria")
        pws = []
        for c in (criterion, criterion_no_reduction):
            pw = c.pos_weight
            if pw is not None and pw.numel() != 1:
                raise NotImplementedError("pos_weight must be a scalar (main_training.py:264)")
            pws.append(1.0 if pw is None else float(pw))
        if pws[0] != pws[1]:
            raise NotImplementedError("the two criteria must share pos_weight")
        return pws[0], False, 0.0, 1.0
    raise NotImplementedError("criterion %s: only BCEWithLogitsLoss and FocalLoss_binary are implemented" % type(criterion).__name__)


def loss_stats(logits: Sequence[torch.Tensor], labels: torch.Tensor, pos_weight: float = 1.0, focal: bool = False,
               focus: float = 5.0, balance: float = 0.9, want_probs: bool = False) -> LossStats:
    """Summed loss over the classified steps + class statistics, on the device, without a host synchronisation."""
    k = len(logits)
    if k < 1 or k > MAX_STEPS:
        raise NotImplementedError("1..%d classified steps" % MAX_STEPS)
    lab = _flat(labels)
    E = lab.numel()
    if E == 0:
        raise ValueError("compute_loss_acc on a batch without edges")
    probs = torch.empty((k, E), dtype=torch.float32, device=lab.device) if want_probs else None
    stats, loss_sum = _BceStats.apply(lab, float(pos_weight), bool(focal), float(focus), float(balance), probs, *logits)
    return LossStats(stats, k, probs, focal, focus, balance, loss_sum)


def compute_loss_acc(outputs, batch, criterion, criterion_no_reduction, mode):
    """Drop-in for train.py:51 ``compute_loss_acc``: same arguments, same 7-tuple
    ``(loss, precision_class1, precision_class0, precision_all, loss_class1, loss_class0, list_pred_prob)``."""
    pw, focal, focus, balance = _criterion_args(criterion, criterion_no_reduction, mode)
    logits: List[torch.Tensor] = list(outputs["classified_edges"])
    st = loss_stats(logits, batch.edge_labels, pw, focal, focus, balance, want_probs=True)
    probs = [st.probs[s] for s in range(st.k)]
    return st.loss, [st.precision_class1], [st.precision_class0], [st.precision_all], st.loss_class1, st.loss_class0, probs
