"""Drop-in for the reference's ``models/mpn.py``: same class names, constructor arguments, parameter tree
(``state_dict`` keys and shapes), modes and ``forward(data)`` contract -- but the
arithmetic runs in hand-written sm_100a kernels (libgnncca) as one fused pipeline per message-passing step.

There is no CPU path and no eager fallback: CPU tensors, or configurations the kernels do not cover, raise.
"""
from __future__ import annotations

import os
from typing import Optional

import torch
from torch import nn

from .mlp import MLP


class MetaLayer(nn.Module):
    """One message-passing step: edge update, then node update on the updated edges (mpn.py:10-57)."""

    def __init__(self, edge_model=None, node_model=None):
        super().__init__()
        self.edge_model = edge_model
        self.node_model = node_model
        self.reset_parameters()

    def reset_parameters(self):
        for item in (self.node_model, self.edge_model):
            if hasattr(item, 'reset_parameters'):
                item.reset_parameters()

    def forward(self, x, edge_index, edge_attr):
        """One message-passing step on caller-supplied tensors (mpn.py:32-54): ``x`` [N, 32 | 64] (64 = cat(initial, current),
        mpn.py:285), ``edge_attr`` [E, 6 | 12], ``edge_index`` [2, E] int64 in any order.  Returns (x', edge_attr') like the
        reference.  Runs the same fused kernels as ``MOTMPNet.forward`` (gnncca_node_project + gnncca_mp_step_fwd);
        inference only -- training goes through ``MOTMPNet.forward``."""
        from . import engine
        return engine.meta_layer_forward(self, x, edge_index, edge_attr)

    def __repr__(self):
        return '{}(edge_model={}, node_model={})'.format(self.__class__.__name__, self.edge_model, self.node_model)


class EdgeModel(nn.Module):
    """Parameter holder of the edge update ``edge_mlp(cat[x_src, x_dst, e])`` (mpn.py:59-69)."""

    def __init__(self, edge_mlp):
        super().__init__()
        self.edge_mlp = edge_mlp

    def forward(self, source, target, edge_attr):
        """``edge_mlp(cat[source, target, edge_attr])`` (mpn.py:67-69) on caller-supplied per-edge rows, through the dense FP32
        kernels (inference only; inside ``MOTMPNet.forward`` this is phase A of the fused step kernel and nothing of width 70
        exists)."""
        from . import engine
        return engine.mlp_forward(self.edge_mlp, torch.cat([source, target, edge_attr], dim=1))


class NodeModel(nn.Module):
    """Parameter holder of the node update ``agg_row(node_mlp(cat[x_src, e']))`` (mpn.py:71-101)."""

    def __init__(self, node_mlp, node_agg_fn):
        super().__init__()
        self.node_mlp = node_mlp
        self.node_agg_fn = node_agg_fn

    def forward(self, x, edge_index, edge_attr):
        """``agg_row(node_mlp(cat[x[row], edge_attr]))`` (mpn.py:97-101) on caller-supplied tensors: the MLP through the dense FP32
        kernels, the aggregation through gnncca_segment_agg on the source-sorted rows (deterministic; inference only)."""
        from . import engine
        return engine.node_model_forward(self, x, edge_index, edge_attr)


class MLPGraphIndependent(nn.Module):
    """Independent node / edge MLPs used as encoder and as classifier (mpn.py:103-142)."""

    def __init__(self, edge_in_dim=None, node_in_dim=None, edge_out_dim=None, node_out_dim=None,
                 node_fc_dims=None, edge_fc_dims=None, dropout_p=None, use_batchnorm=None):
        super().__init__()
        if node_in_dim is not None:
            self.node_mlp = MLP(input_dim=node_in_dim, fc_dims=list(node_fc_dims) + [node_out_dim],
                                dropout_p=dropout_p, use_batchnorm=use_batchnorm)
        else:
            self.node_mlp = None
        if edge_in_dim is not None:
            self.edge_mlp = MLP(input_dim=edge_in_dim, fc_dims=list(edge_fc_dims) + [edge_out_dim],
                                dropout_p=dropout_p, use_batchnorm=use_batchnorm)
        else:
            self.edge_mlp = None

    def forward(self, edge_feats=None, nodes_feats=None):
        out_nodes = self.node_mlp(nodes_feats) if (self.node_mlp is not None and nodes_feats is not None) else nodes_feats
        out_edges = self.edge_mlp(edge_feats) if (self.edge_mlp is not None and edge_feats is not None) else edge_feats
        return out_edges, out_nodes


class MOTMPNet(nn.Module):
    """Encoder -> L x (edge update, node update) -> per-step edge classifier (mpn.py:144-299).

    Extra, optional keyword (not in the reference): ``math`` = 'tf32' (default, or $GNNCCA_MATH: the node encoder's
    2048-wide contraction runs on tcgen05 tensor cores with TF32 operands, everything else in FP32) or 'fp32' (FFMA
    everywhere; the tight-tolerance path).  Training always runs in FP32.
    """

    def __init__(self, model_params, bb_encoder=None, arch=None, math: Optional[str] = None):
        super().__init__()
        self.node_cnn = bb_encoder          # stored and never used, like the reference (mpn.py:163)
        self.model_params = model_params
        self.math = math if math is not None else os.environ.get("GNNCCA_MATH", "tf32")
        if self.math not in ("tf32", "fp32"):
            raise ValueError("math must be 'tf32' or 'fp32'")

        edges_params = model_params['encoder_feats_dict']['edges']
        nodes_params = model_params['encoder_feats_dict']['nodes'][arch]
        edges_params.update(nodes_params)   # in-place merge: a side effect callers can observe (mpn.py:169)
        encoder_feats_dict = edges_params
        classifier_feats_dict = model_params['classifier_feats_dict']

        self.encoder = MLPGraphIndependent(**encoder_feats_dict)
        self.classifier = MLPGraphIndependent(**classifier_feats_dict)
        self.MPNet = self._build_core_MPNet(model_params=model_params, encoder_feats_dict=encoder_feats_dict)
        self.num_enc_steps = model_params['num_enc_steps']
        self.num_class_steps = model_params['num_class_steps']

    def _build_core_MPNet(self, model_params, encoder_feats_dict):
        node_agg_fn = model_params['node_agg_fn']
        assert node_agg_fn.lower() in ('mean', 'max', 'sum'), "node_agg_fn can only be 'max', 'mean' or 'sum'."
        if node_agg_fn not in ('mean', 'max', 'sum'):
            # the reference leaves a mixed-case string uncalled-for here and crashes at the first forward
            raise ValueError("node_agg_fn must be spelled 'mean', 'max' or 'sum' (lower case)")
        self.node_agg_name = node_agg_fn

        self.reattach_initial_nodes = model_params['reattach_initial_nodes']
        self.reattach_initial_edges = model_params['reattach_initial_edges']
        edge_factor = 2 if self.reattach_initial_edges else 1
        node_factor = 2 if self.reattach_initial_nodes else 1
        f_node, f_edge = encoder_feats_dict['node_out_dim'], encoder_feats_dict['edge_out_dim']
        edge_model_in_dim = node_factor * 2 * f_node + edge_factor * f_edge
        node_model_in_dim = node_factor * f_node + f_edge

        em, nm = model_params['edge_model_feats_dict'], model_params['node_model_feats_dict']
        edge_mlp = MLP(input_dim=edge_model_in_dim, fc_dims=em['fc_dims'], dropout_p=em['dropout_p'],
                       use_batchnorm=em['use_batchnorm'])
        node_mlp = MLP(input_dim=node_model_in_dim, fc_dims=nm['fc_dims'], dropout_p=nm['dropout_p'],
                       use_batchnorm=nm['use_batchnorm'])
        # the reference builds (and discards) one more Linear here (mpn.py:241-242); draw the same random
        # numbers so that a seeded program continues with the same RNG state after construction
        nn.Linear(2 * f_node, f_node)
        return MetaLayer(edge_model=EdgeModel(edge_mlp=edge_mlp),
                         node_model=NodeModel(node_mlp=node_mlp, node_agg_fn=node_agg_fn))

    def forward(self, data):
        from . import autograd_fn, engine
        edge_index, edge_attr = data.edge_index, data.edge_attr
        needs_grad = torch.is_grad_enabled() and any(p.requires_grad for _, p in self.named_parameters_fast())
        train_path = needs_grad or (self.training and self._has_batchnorm())
        # A batch from build_graph_batch carries the RAW node embeddings + their column norms (graph.Batch.x): inference feeds the raw
        # matrix to the encoder and scales the weight columns of its first Linear instead of materialising the normalised copy
        raw_fn = getattr(data, "raw_x", None)
        raw = raw_fn() if (callable(raw_fn) and not train_path) else None
        if raw is not None:
            x = raw[0]
            enc = self.encoder.node_mlp
            enc.__dict__["_gnncca_w0"] = self._scaled_first_weight(data, raw[1])
            try:
                return self._forward(data, x, edge_index, edge_attr, needs_grad, train_path)
            finally:
                enc.__dict__.pop("_gnncca_w0", None)
        return self._forward(data, data.x, edge_index, edge_attr, needs_grad, train_path)

    def _scaled_first_weight(self, data, norm):
        """W1 / norm (column-wise) of the node encoder's first Linear, cached on the batch while neither changes."""
        from . import _lib
        from .plan import _stream
        w = self.encoder.node_mlp.linear(0).weight
        key = (w.data_ptr(), w._version, norm.data_ptr(), norm._version)
        hit = data.__dict__.get("_gnncca_w1s") if hasattr(data, "__dict__") else None
        if hit is not None and hit[0] == key:
            return hit[1]
        out = torch.empty_like(w)
        _lib.check(_lib.load().gnncca_scale_columns(w.data_ptr(), out.data_ptr(), w.size(0), w.size(1), norm.data_ptr(), _stream()), "scale_columns")
        try:
            data.__dict__["_gnncca_w1s"] = (key, out)
        except AttributeError:
            pass
        return out

    def _forward(self, data, x, edge_index, edge_attr, needs_grad, train_path):
        from . import autograd_fn, engine
        graph_ptr = getattr(data, "ptr", None)             # PyG Batch.ptr / our Batch.ptr: frame boundaries, when known
        caps = getattr(data, "gnncca_caps", None)
        trusted = bool(getattr(data, "gnncca_trusted", False))
        plan = getattr(data, "gnncca_plan", None)          # CSR verified by an earlier forward on this very object
        ptr_key = (graph_ptr.data_ptr(), graph_ptr._version) if torch.is_tensor(graph_ptr) else None
        if plan is not None and (not engine.CACHE_PLANS or plan.key() != (edge_index.data_ptr(), tuple(edge_index.shape),
                                                                          edge_index._version, int(x.size(0)))
                                 or getattr(plan, "ptr_key", None) != ptr_key):
            plan = None       # edge_index or ptr were replaced / edited in place since the plan was verified
        allow_fused = True
        if plan is not None and plan.last_path is not None:      # verified before: nothing left to check
            trusted, allow_fused = True, plan.last_path == "fused"
            caps = getattr(plan, "caps", caps) if allow_fused else caps
        if (not needs_grad and not self.training and caps is not None and getattr(data, "gnncca_trusted", False)
                and getattr(data, "edge_ptr", None) is not None and graph_ptr is not None):
            # batches from build_graph_batch in eval mode: one library call, two launches, nothing read back
            fp = engine.fast_path(self)
            outs = fp.run(self, x, edge_index, edge_attr, graph_ptr, data.edge_ptr, caps) if fp is not None else None
            if outs is not None:
                self.last_plan = engine.FAST_PLAN
                return {'classified_edges': outs}
        if train_path:
            frames = None
            if trusted and getattr(data, "edge_ptr", None) is not None and graph_ptr is not None and caps is not None:
                frames = (graph_ptr, data.edge_ptr, caps)      # closed frames with known offsets: the backward runs as one launch
            outs, plan = autograd_fn.forward_training(self, x, edge_index, edge_attr, plan, trusted=trusted, return_plan=True, frames=frames)
            if engine.CACHE_PLANS and plan is not None:
                try:                    # the CSR (and its transpose for the backward) is reused while edge_index is unchanged
                    plan.ptr_key = ptr_key
                    data.gnncca_plan = plan
                except AttributeError:
                    pass
        else:
            edge_ptr = getattr(data, "edge_ptr", None) if getattr(data, "gnncca_trusted", False) else None
            outs, plan = engine.forward_inference(self, x, edge_index, edge_attr, plan, graph_ptr=graph_ptr, caps=caps,
                                                  return_plan=True, trusted=trusted, allow_fused=allow_fused, edge_ptr=edge_ptr,
                                                  cache_on=data)
            self.last_plan = plan       # introspection only (which device pipeline ran, the CSR)
            if engine.CACHE_PLANS and not getattr(data, "gnncca_trusted", False) and plan is not engine.BLOCKED_PLAN:
                try:                    # a verified CSR is reused by later forwards on the same, unmodified Batch
                    plan.ptr_key = ptr_key
                    data.gnncca_plan = plan
                    if caps is None and getattr(plan, "caps", None) is not None and plan.last_path == "fused":
                        data.gnncca_caps = plan.caps
                except AttributeError:
                    pass
        return {'classified_edges': outs}

    def forward_host(self, x, edge_index, edge_attr, out=None, graph_ptr=None, caps=None):
        """The same forward for HOST tensors (pinned memory recommended): chunked host->device copies overlap the
        node encoder, per-step logits are copied back into ``out`` (list of host tensors [E,1])."""
        from . import engine
        return engine.forward_host(self, x, edge_index, edge_attr, out=out, graph_ptr=graph_ptr, caps=caps)

    def _has_batchnorm(self) -> bool:
        hb = self.__dict__.get("_gnncca_has_bn")
        if hb is None:
            hb = self.__dict__["_gnncca_has_bn"] = any(isinstance(m, nn.BatchNorm1d) for m in self.modules())
        return hb

    def named_parameters_fast(self):
        """``list(self.named_parameters())`` without walking the module tree on every call: the (owner dict, key) of
        each parameter slot is recorded once (the tree is fixed after construction); re-assigned parameters are seen."""
        slots = self.__dict__.get("_gnncca_param_slots")
        if slots is None:
            slots = []
            for mname, m in self.named_modules():
                for k, t in m._parameters.items():
                    if t is not None:
                        slots.append(((mname + "." if mname else "") + k, m._parameters, k))
            self.__dict__["_gnncca_param_slots"] = slots
        return [(n, d[k]) for n, d, k in slots]
