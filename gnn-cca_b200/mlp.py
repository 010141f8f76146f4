"""Drop-in for the reference's ``models/mlp.py`` (``MLP``, mlp.py:4-28).

Same constructor, same ``fc_layers.<idx>`` parameter names (the index counts Linear, BatchNorm1d, ReLU and
Dropout slots exactly like the reference's ``nn.Sequential``, so ``dropout_p: 0`` still shifts the
numbering), same default initialisation order (so a seeded construction yields the same weights).  The
torch modules are parameter containers only: ``forward`` runs the sm_100a kernels of libgnncca.
"""
from __future__ import annotations

from typing import List, NamedTuple, Optional, Sequence

import torch
from torch import nn


class LayerSpec(NamedTuple):
    linear: int               # index of the Linear inside fc_layers
    bn: Optional[int]         # index of the BatchNorm1d, if any
    relu: bool
    dropout_p: Optional[float]
    in_dim: int
    out_dim: int


def plan_layers(input_dim: int, fc_dims: Sequence[int], dropout_p, use_batchnorm) -> List[LayerSpec]:
    """Slot numbering of mlp.py:12-23: a width-1 layer gets neither BN, ReLU nor Dropout."""
    specs, slot = [], 0
    for dim in fc_dims:
        lin = slot; slot += 1
        bn = None
        hidden = dim != 1
        if use_batchnorm and hidden:
            bn = slot; slot += 1
        if hidden:
            slot += 1  # ReLU
        dp = dropout_p if (dropout_p is not None and hidden) else None
        if dp is not None:
            slot += 1
        specs.append(LayerSpec(lin, bn, hidden, dp, input_dim, dim))
        input_dim = dim
    return specs


class MLP(nn.Module):
    def __init__(self, input_dim, fc_dims, dropout_p=0.4, use_batchnorm=False):
        super().__init__()
        assert isinstance(fc_dims, (list, tuple)), \
            'fc_dims must be either a list or a tuple, but got {}'.format(type(fc_dims))
        self.layer_specs = plan_layers(input_dim, fc_dims, dropout_p, use_batchnorm)
        mods: List[nn.Module] = []
        for sp in self.layer_specs:
            mods.append(nn.Linear(sp.in_dim, sp.out_dim))
            if sp.bn is not None:
                mods.append(nn.BatchNorm1d(sp.out_dim))
            if sp.relu:
                mods.append(nn.ReLU(inplace=True))
            if sp.dropout_p is not None:
                mods.append(nn.Dropout(p=sp.dropout_p))
        self.fc_layers = nn.Sequential(*mods)
        self.input_dim = input_dim

    # ------------------------------------------------------------------ helpers used by the fused pipeline
    def linear(self, i: int) -> nn.Linear:
        return self.fc_layers[self.layer_specs[i].linear]

    def batchnorm(self, i: int) -> Optional[nn.BatchNorm1d]:
        b = self.layer_specs[i].bn
        return None if b is None else self.fc_layers[b]

    def check_runnable(self, what: str) -> None:
        """Configurations libgnncca has no kernel for fail loudly (there is no eager fallback)."""
        for sp in self.layer_specs:
            if self.training and sp.dropout_p not in (None, 0, 0.0):
                raise NotImplementedError("%s: dropout_p=%r in training mode has no CUDA kernel in gnn-cca_b200" % (what, sp.dropout_p))
            if self.training and sp.bn is not None:
                raise NotImplementedError("%s: BatchNorm1d in training mode is only implemented for the classifier" % what)

    def forward(self, input):
        from . import engine
        return engine.mlp_forward(self, input)
