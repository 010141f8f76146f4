"""CPU restatement of the reference message-passing network (TEST INFRASTRUCTURE ONLY).

Follows ``/root/reference/models/mlp.py`` and ``/root/reference/models/mpn.py`` function by function,
written as plain functional torch-CPU fp32 code over a ``state_dict`` so that it can travel to the
GPU box (where ``/root/reference`` does not exist).  It is pinned against outputs of the reference
itself (``tests/golden/mpn_*.npz``, produced by ``oracle/gen_golden.py``) and, when
``/root/reference`` is present, compared live with the imported reference module
(``tests/test_oracle_vs_reference.py``).

The third-party pieces the reference calls but does not vendor are restated here from their
published semantics: ``torch_scatter`` 2.0.8 ``scatter_add / scatter_mean / scatter_max``
(``env_gnn.yml:97``; call sites ``models/mpn.py:196,199,202``).
"""
from __future__ import annotations

import copy
from typing import Dict, List, Optional, Sequence, Tuple

import torch

Tensor = torch.Tensor


# --------------------------------------------------------------------------------------
# torch_scatter 2.0.8 (third-party, absent): restated
# --------------------------------------------------------------------------------------
def scatter_add(src: Tensor, index: Tensor, dim: int = 0, dim_size: Optional[int] = None) -> Tensor:
    """out[i] = sum of src rows whose index == i; rows never hit stay 0 (torch_scatter.scatter_add)."""
    assert dim == 0
    out = torch.zeros((dim_size,) + tuple(src.shape[1:]), dtype=src.dtype)
    return out.index_add_(0, index, src)


def scatter_mean(src: Tensor, index: Tensor, dim: int = 0, dim_size: Optional[int] = None) -> Tensor:
    """sum / max(count, 1) (torch_scatter.scatter_mean: count is clamped to >= 1 before dividing)."""
    assert dim == 0
    out = scatter_add(src, index, 0, dim_size)
    cnt = torch.zeros(dim_size, dtype=src.dtype).index_add_(0, index, torch.ones(index.numel(), dtype=src.dtype))
    cnt = cnt.clamp_(min=1)
    return out / cnt.view(-1, *([1] * (src.dim() - 1)))


def scatter_max(src: Tensor, index: Tensor, dim: int = 0, dim_size: Optional[int] = None) -> Tuple[Tensor, Tensor]:
    """(values, argmax); segments with no entry return value 0 and arg == src.size(dim)."""
    assert dim == 0
    idx = index.view(-1, *([1] * (src.dim() - 1))).expand_as(src)
    lowest = torch.full((dim_size,) + tuple(src.shape[1:]), float("-inf"), dtype=src.dtype)
    out = lowest.scatter_reduce(0, idx, src, reduce="amax", include_self=True)
    hit = torch.zeros(dim_size, dtype=torch.bool)
    hit[index] = True
    hit = hit.view(-1, *([1] * (src.dim() - 1))).expand_as(out)
    out = torch.where(hit, out, torch.zeros_like(out))
    # argmax: first edge attaining the max (only needed for completeness)
    arg = torch.full(out.shape, src.size(0), dtype=torch.long)
    eq = src == out[index]
    pos = torch.arange(src.size(0)).view(-1, *([1] * (src.dim() - 1))).expand_as(src)
    cand = torch.where(eq, pos, torch.full_like(pos, src.size(0)))
    arg = arg.scatter_reduce(0, idx, cand, reduce="amin", include_self=True)
    return out, arg


# --------------------------------------------------------------------------------------
# models/mlp.py:4-28
# --------------------------------------------------------------------------------------
def mlp_layout(input_dim: int, fc_dims: Sequence[int], dropout_p=0.4, use_batchnorm=False) -> List[tuple]:
    """The nn.Sequential the reference builds, as a list of (index, kind, in, out).

    mlp.py:12-23: per dim: Linear; BatchNorm1d iff use_batchnorm and dim != 1; ReLU iff dim != 1;
    Dropout iff dropout_p is not None and dim != 1.  The running index is the ``fc_layers.<idx>`` key.
    """
    assert isinstance(fc_dims, (list, tuple)), "fc_dims must be either a list or a tuple, but got {}".format(type(fc_dims))
    layers, idx = [], 0
    for dim in fc_dims:
        layers.append((idx, "linear", input_dim, dim)); idx += 1
        if use_batchnorm and dim != 1:
            layers.append((idx, "bn", dim, dim)); idx += 1
        if dim != 1:
            layers.append((idx, "relu", dim, dim)); idx += 1
        if dropout_p is not None and dim != 1:
            layers.append((idx, "dropout", dropout_p, None)); idx += 1
        input_dim = dim
    return layers


def mlp_forward(sd: Dict[str, Tensor], prefix: str, layout: List[tuple], inp: Tensor, training: bool = False,
                bn_eps: float = 1e-5) -> Tensor:
    """mlp.py:27-28 ``self.fc_layers(input)`` over the layout above."""
    out = inp
    for idx, kind, a, _b in layout:
        key = "{}.fc_layers.{}".format(prefix, idx)
        if kind == "linear":
            out = torch.nn.functional.linear(out, sd[key + ".weight"], sd[key + ".bias"])  # nn.Linear
        elif kind == "bn":
            # nn.BatchNorm1d (torch): batch statistics (biased var) in training, running stats in eval;
            # eps 1e-5.  Running buffers are cloned: the oracle never mutates the caller's state_dict.
            out = torch.nn.functional.batch_norm(
                out, sd[key + ".running_mean"].clone(), sd[key + ".running_var"].clone(),
                sd[key + ".weight"], sd[key + ".bias"], training=training, momentum=0.1, eps=bn_eps)
        elif kind == "relu":
            out = torch.relu(out)
        elif kind == "dropout":
            if training and a not in (0, 0.0):
                raise NotImplementedError("oracle: dropout_p > 0 in training mode is stochastic")
    return out


# --------------------------------------------------------------------------------------
# models/mpn.py
# --------------------------------------------------------------------------------------
class MPNSpec:
    """Everything ``MOTMPNet.__init__`` derives from ``model_params`` (mpn.py:154-247), no tensors."""

    def __init__(self, model_params: dict, arch: Optional[str]):
        model_params = copy.deepcopy(model_params)  # the reference mutates in place (mpn.py:167-170)
        edges_params = model_params["encoder_feats_dict"]["edges"]
        nodes_params = model_params["encoder_feats_dict"]["nodes"][arch]
        edges_params.update(nodes_params)  # mpn.py:169
        enc = edges_params
        cls = model_params["classifier_feats_dict"]
        self.enc, self.cls = enc, cls

        def gi(d, **kw):  # MLPGraphIndependent.__init__ (mpn.py:112-126)
            full = dict(edge_in_dim=None, node_in_dim=None, edge_out_dim=None, node_out_dim=None,
                        node_fc_dims=None, edge_fc_dims=None, dropout_p=None, use_batchnorm=None)
            full.update(d)
            node = edge = None
            if full["node_in_dim"] is not None:
                node = mlp_layout(full["node_in_dim"], list(full["node_fc_dims"]) + [full["node_out_dim"]],
                                  full["dropout_p"], full["use_batchnorm"])
            if full["edge_in_dim"] is not None:
                edge = mlp_layout(full["edge_in_dim"], list(full["edge_fc_dims"]) + [full["edge_out_dim"]],
                                  full["dropout_p"], full["use_batchnorm"])
            return node, edge

        self.enc_node, self.enc_edge = gi(enc)
        self.cls_node, self.cls_edge = gi(cls)

        agg = model_params["node_agg_fn"]
        assert agg.lower() in ("mean", "max", "sum"), "node_agg_fn can only be 'max', 'mean' or 'sum'."  # mpn.py:193
        self.agg = agg
        self.reattach_nodes = model_params["reattach_initial_nodes"]
        self.reattach_edges = model_params["reattach_initial_edges"]
        ef = 2 if self.reattach_edges else 1
        nf = 2 if self.reattach_nodes else 1
        edge_in = nf * 2 * enc["node_out_dim"] + ef * enc["edge_out_dim"]  # mpn.py:213
        node_in = nf * enc["node_out_dim"] + enc["edge_out_dim"]  # mpn.py:215
        em, nm = model_params["edge_model_feats_dict"], model_params["node_model_feats_dict"]
        self.edge_model = mlp_layout(edge_in, em["fc_dims"], em["dropout_p"], em["use_batchnorm"])
        self.node_model = mlp_layout(node_in, nm["fc_dims"], nm["dropout_p"], nm["use_batchnorm"])
        self.num_enc_steps = model_params["num_enc_steps"]
        self.num_class_steps = model_params["num_class_steps"]


def init_state_dict(spec: MPNSpec, seed: int = 0) -> Dict[str, Tensor]:
    """Random-init weights in the order ``MOTMPNet.__init__`` creates its modules (mpn.py:173-174 encoder then
    classifier, each node MLP before edge MLP, mpn.py:116-126; then edge_mlp, node_mlp and the dead
    ``node_mlp_old`` Linear of mpn.py:217-242), with torch's default ``nn.Linear`` / ``nn.BatchNorm1d`` init."""
    torch.manual_seed(seed)
    sd: Dict[str, Tensor] = {}

    def build(prefix, layout):
        if layout is None:
            return
        for idx, kind, a, b in layout:
            key = "{}.fc_layers.{}".format(prefix, idx)
            if kind == "linear":
                lin = torch.nn.Linear(a, b)
                sd[key + ".weight"], sd[key + ".bias"] = lin.weight.detach(), lin.bias.detach()
            elif kind == "bn":
                bn = torch.nn.BatchNorm1d(a)
                for name, t in bn.state_dict().items():
                    sd[key + "." + name] = t.detach()

    build("encoder.node_mlp", spec.enc_node)
    build("encoder.edge_mlp", spec.enc_edge)
    build("classifier.node_mlp", spec.cls_node)
    build("classifier.edge_mlp", spec.cls_edge)
    build("MPNet.edge_model.edge_mlp", spec.edge_model)
    build("MPNet.node_model.node_mlp", spec.node_model)
    return sd


def motmpnet_forward(spec: MPNSpec, sd: Dict[str, Tensor], x: Tensor, edge_index: Tensor, edge_attr: Tensor,
                     training: bool = False, return_latents: bool = False):
    """``MOTMPNet.forward`` (mpn.py:250-299) on CPU tensors; returns {'classified_edges': [...]}.

    With ``return_latents`` also returns the per-step (node, edge) latent features, used to check the
    CUDA step kernels one step at a time.
    """
    row, col = edge_index[0], edge_index[1]
    # encoder: MLPGraphIndependent.forward (mpn.py:128-142)
    e = mlp_forward(sd, "encoder.edge_mlp", spec.enc_edge, edge_attr, training) if spec.enc_edge is not None else edge_attr
    h = mlp_forward(sd, "encoder.node_mlp", spec.enc_node, x, training) if spec.enc_node is not None else x
    e0, h0 = e, h
    first_class_step = spec.num_enc_steps - spec.num_class_steps + 1  # mpn.py:277
    outs, latents = [], [(h, e)]

    def classify(ee):
        if spec.cls_edge is None:
            return ee
        return mlp_forward(sd, "classifier.edge_mlp", spec.cls_edge, ee, training)

    for step in range(1, spec.num_enc_steps + 1):
        if spec.reattach_edges:
            e = torch.cat((e0, e), dim=1)  # mpn.py:283
        if spec.reattach_nodes:
            h = torch.cat((h0, h), dim=1)  # mpn.py:285
        # MetaLayer.forward (mpn.py:44-54): edge model sees OLD nodes, node model sees NEW edges
        e = mlp_forward(sd, "MPNet.edge_model.edge_mlp", spec.edge_model, torch.cat([h[row], h[col], e], dim=1), training)
        m = mlp_forward(sd, "MPNet.node_model.node_mlp", spec.node_model, torch.cat([h[row], e], dim=1), training)
        n = h.size(0)
        if spec.agg == "mean":
            h = scatter_mean(m, row, 0, n)
        elif spec.agg == "max":
            h = scatter_max(m, row, 0, n)[0]
        else:
            h = scatter_add(m, row, 0, n)
        latents.append((h, e))
        if step >= first_class_step:
            outs.append(classify(e))
    if spec.num_enc_steps == 0:
        outs.append(classify(e))
    out = {"classified_edges": outs}
    return (out, latents) if return_latents else out
