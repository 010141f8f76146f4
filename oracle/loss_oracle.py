"""CPU restatement of the reference's loss and statistics (TEST INFRASTRUCTURE ONLY).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU legs may import this module; the product path
(``gnn-cca_b200/loss.py`` -> ``libgnncca.so``) never does.

Restates ``compute_loss_acc`` (/root/reference/train.py:51-133, return at :208), the three criteria of
/root/reference/main_training.py:258-268 and ``FocalLoss_binary`` (/root/reference/libs/utils.py:264-293) with plain
torch CPU ops in float64-free form (float32, as the reference computes).  Pinned by ``tests/golden/loss_*.npz``, which
``oracle/gen_golden.py`` writes by executing the reference's own, unmodified ``compute_loss_acc``/``FocalLoss_binary``
source (tests/test_oracle_golden.py).
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F


def bce_with_logits(x: torch.Tensor, y: torch.Tensor, pos_weight=None) -> torch.Tensor:
    """Per-sample torch.nn.BCEWithLogitsLoss: (1 - y) x + (1 + (pw - 1) y) softplus(-x)."""
    pw = None if pos_weight is None else torch.as_tensor(pos_weight, dtype=x.dtype)
    return F.binary_cross_entropy_with_logits(x, y, reduction="none", pos_weight=pw)


def focal(ce: torch.Tensor, focusing_param=5, balance_param=0.9) -> torch.Tensor:
    """utils.py:281-291 applied to a cross entropy `ce` (per sample for reduction 'none', the mean for 'mean')."""
    logpt = -ce
    pt = torch.exp(logpt)
    return balance_param * (-((1 - pt) ** focusing_param) * logpt)


def compute_loss_acc(logits, labels, loss_name="BCE", pos_weight=None, mode="train", focusing_param=5, balance_param=0.9):
    """train.py:51-133 on CPU tensors.  `logits`: list of [E] or [E,1] tensors (outputs['classified_edges']), `labels` [E].

    Returns a dict: loss, loss_class1, loss_class0 (0-dim tensors), precision_class1/0/all (Python numbers, in %),
    probs (list of [E] sigmoids), mean_prob_class1/0 (lists; train.py:461-470's per-class mean probabilities)."""
    labels = labels.reshape(-1).float()
    loss = torch.zeros(())
    l1 = torch.zeros(())
    l0 = torch.zeros(())
    probs, mp1, mp0 = [], [], []
    for preds in logits:                                                 # train.py:79 (all steps)
        preds = preds.reshape(-1).float()
        if mode == "train" and loss_name == "Focal":                     # main_training.py:259-261
            per_sample = focal(bce_with_logits(preds, labels), focusing_param, balance_param)
            loss = loss + focal(bce_with_logits(preds, labels).mean(), focusing_param, balance_param)
        elif mode == "train" and loss_name == "BCE_weighted":            # main_training.py:263-265
            per_sample = bce_with_logits(preds, labels, pos_weight)
            loss = loss + per_sample.mean()
        else:                                                            # 'BCE' (main_training.py:267-268) / validate (train.py:89-94)
            per_sample = bce_with_logits(preds, labels)
            loss = loss + per_sample.mean()
        l1 = l1 + per_sample[labels == 1].mean()                         # train.py:86-87 / 91-92 (NaN for an empty class)
        l0 = l0 + per_sample[labels == 0].mean()
        p = torch.sigmoid(preds.detach())                                # train.py:99-102
        probs.append(p)
        mp1.append(p[labels == 1].mean())
        mp0.append(p[labels == 0].mean())
    pred = (probs[-1] >= 0.5).to(torch.int64).numpy()                    # train.py:106-110: last step only
    lab = labels.numpy()

    def pct(mask):
        ok = int(np.sum(pred[mask] == lab[mask]))
        return 0 if ok == 0 else ok / int(mask.sum()) * 100.0            # train.py:113-117

    return dict(loss=loss, loss_class1=l1, loss_class0=l0, precision_class1=pct(lab == 1), precision_class0=pct(lab == 0),
                precision_all=pct(np.ones_like(lab, dtype=bool)), probs=probs, mean_prob_class1=mp1, mean_prob_class0=mp0)
