"""Host-side orchestration of the fused pipeline: which kernel runs on which buffer, in which order.
PyTorch is used for device memory (caching allocator) and streams only; all arithmetic is in libgnncca."""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import torch

from . import _lib
from .plan import GraphPlan, _stream

MATH_MODES = {"fp32": _lib.MATH_FP32, "tf32": _lib.MATH_TF32}


def _require_cuda(t: torch.Tensor, name: str) -> None:
    if not t.is_cuda:
        raise RuntimeError("gnn-cca_b200: %s must be a CUDA tensor; this package has no CPU or eager fallback" % name)
    if t.dtype != torch.float32:
        raise TypeError("gnn-cca_b200: %s must be float32, got %s" % (name, t.dtype))


def linear(x: torch.Tensor, lin: torch.nn.Linear, relu: bool, bn: Optional[torch.nn.BatchNorm1d] = None,
           math: int = _lib.MATH_FP32) -> torch.Tensor:
    """models/mlp.py:14-19 for one fc_dims entry: Linear [+ BatchNorm1d(eval)] [+ ReLU]."""
    lib = _lib.load()
    x = x.contiguous()
    rows, in_dim = x.shape
    y = torch.empty((rows, lin.out_features), dtype=torch.float32, device=x.device)
    bnw = bnb = bnm = bnv = None
    eps = 0.0
    if bn is not None:
        bnw, bnb, bnm, bnv, eps = bn.weight, bn.bias, bn.running_mean, bn.running_var, bn.eps
    _lib.check(lib.gnncca_linear(_lib.ptr(x), _lib.ptr(lin.weight), _lib.ptr(lin.bias), _lib.ptr(y), rows, in_dim,
                                 lin.out_features, int(relu), _lib.ptr(bnw), _lib.ptr(bnb), _lib.ptr(bnm), _lib.ptr(bnv),
                                 float(eps), math, _stream()), "linear")
    return y


def mlp_forward(mlp, inp: torch.Tensor, math: int = _lib.MATH_FP32) -> torch.Tensor:
    """``MLP.forward`` (mlp.py:27-28) through the dense kernels; inference only (no autograd graph)."""
    _require_cuda(inp, "input")
    if torch.is_grad_enabled() and any(p.requires_grad for p in mlp.parameters()):
        raise NotImplementedError("stand-alone MLP.forward is inference-only; training runs through MOTMPNet.forward")
    mlp.check_runnable("MLP")
    out = inp
    for i, sp in enumerate(mlp.layer_specs):
        use_tc = math if (sp.in_dim % 32 == 0 and sp.out_dim % 16 == 0 and sp.out_dim <= 256 and sp.bn is None) else _lib.MATH_FP32
        out = linear(out, mlp.linear(i), sp.relu, mlp.batchnorm(i), use_tc)
    return out


def small_mlp_rows(mlp, rows: torch.Tensor) -> torch.Tensor:
    """Per-row MLP chain for narrow inputs (edge encoder 4 -> 6; stand-alone classifier)."""
    lib = _lib.load()
    specs = mlp.layer_specs
    n = len(specs)
    if any(sp.bn is not None for sp in specs):
        raise NotImplementedError("BatchNorm1d inside a per-edge encoder MLP has no kernel in gnn-cca_b200")
    dims = (C.c_int32 * (n + 1))(*([specs[0].in_dim] + [sp.out_dim for sp in specs]))
    Ws = (C.c_void_p * n)(*[mlp.linear(i).weight.data_ptr() for i in range(n)])
    bs = (C.c_void_p * n)(*[mlp.linear(i).bias.data_ptr() for i in range(n)])
    relu = (C.c_int32 * n)(*[int(sp.relu) for sp in specs])
    rows = rows.contiguous()
    y = torch.empty((rows.size(0), specs[-1].out_dim), dtype=torch.float32, device=rows.device)
    _lib.check(lib.gnncca_edge_mlp(_lib.ptr(rows), _lib.ptr(y), rows.size(0), n, dims, Ws, bs, relu, _stream()), "edge_mlp")
    return y


class StepBinding:
    """ctypes views of the step configuration and of the live parameter pointers of a MOTMPNet."""

    def __init__(self, model, math: int):
        enc = model.encoder
        self.f_node = enc.node_mlp.layer_specs[-1].out_dim if enc.node_mlp is not None else None
        self.f_edge = enc.edge_mlp.layer_specs[-1].out_dim if enc.edge_mlp is not None else None
        cls = model.classifier.edge_mlp
        self.cfg = _lib.MpCfg(f_node=self.f_node, f_edge=self.f_edge, cls_hidden=cls.layer_specs[0].out_dim,
                              agg=_lib.AGG[model.node_agg_name], reattach_nodes=int(model.reattach_initial_nodes),
                              reattach_edges=int(model.reattach_initial_edges), math=math, reserved=0)
        em, nm = model.MPNet.edge_model.edge_mlp, model.MPNet.node_model.node_mlp
        bn = cls.batchnorm(0)
        self.w = _lib.MpWeights(
            w_edge=em.linear(0).weight.data_ptr(), b_edge=em.linear(0).bias.data_ptr(),
            w_node=nm.linear(0).weight.data_ptr(), b_node=nm.linear(0).bias.data_ptr(),
            w_c1=cls.linear(0).weight.data_ptr(), b_c1=cls.linear(0).bias.data_ptr(),
            bn_weight=_lib.ptr(bn.weight) if bn is not None else 0, bn_bias=_lib.ptr(bn.bias) if bn is not None else 0,
            bn_mean=_lib.ptr(bn.running_mean) if bn is not None else 0, bn_var=_lib.ptr(bn.running_var) if bn is not None else 0,
            w_c2=cls.linear(1).weight.data_ptr(), b_c2=cls.linear(1).bias.data_ptr(),
            bn_eps=float(bn.eps) if bn is not None else 0.0, pad_=0)


def check_supported(model) -> None:
    """The configuration family the fused step kernels cover; anything else raises (no fallback)."""
    enc, cls = model.encoder, model.classifier
    if enc.node_mlp is None or enc.edge_mlp is None:
        raise NotImplementedError("gnn-cca_b200 needs both encoder MLPs (node_in_dim and edge_in_dim)")
    if cls.edge_mlp is None:
        raise NotImplementedError("gnn-cca_b200 needs a classifier edge MLP")
    em, nm = model.MPNet.edge_model.edge_mlp, model.MPNet.node_model.node_mlp
    f_node, f_edge = enc.node_mlp.layer_specs[-1].out_dim, enc.edge_mlp.layer_specs[-1].out_dim
    if len(em.layer_specs) != 1 or len(nm.layer_specs) != 1:
        raise NotImplementedError("fused step: edge_model / node_model fc_dims must have exactly one layer")
    if em.layer_specs[0].out_dim != f_edge or nm.layer_specs[0].out_dim != f_node:
        raise NotImplementedError("fused step: update MLP widths must equal the encoder's output widths")
    if em.layer_specs[0].bn is not None or nm.layer_specs[0].bn is not None:
        raise NotImplementedError("fused step: BatchNorm inside the edge/node update MLPs is not implemented")
    if not em.layer_specs[0].relu or not nm.layer_specs[0].relu:
        raise NotImplementedError("fused step: width-1 update MLPs (no ReLU) are not implemented")
    cs = cls.edge_mlp.layer_specs
    if len(cs) != 2 or cs[0].in_dim != f_edge or cs[1].out_dim != 1:
        raise NotImplementedError("fused step: classifier must be edge_in_dim=f_edge, one hidden layer, edge_out_dim=1")
    if (f_node, f_edge, cs[0].out_dim) != (32, 6, 4):
        raise NotImplementedError("fused step kernels are built for (node_out_dim, edge_out_dim, classifier hidden) = "
                                  "(32, 6, 4); got (%d, %d, %d)" % (f_node, f_edge, cs[0].out_dim))
    for name, m in (("encoder.node_mlp", enc.node_mlp), ("encoder.edge_mlp", enc.edge_mlp), ("edge_model", em), ("node_model", nm)):
        m.check_runnable(name)
    for sp in cs:
        if model.training and sp.dropout_p not in (None, 0, 0.0):
            raise NotImplementedError("classifier dropout_p > 0 in training mode is not implemented")


def forward_inference(model, x: torch.Tensor, edge_index: torch.Tensor, edge_attr: torch.Tensor,
                      plan: Optional[GraphPlan] = None, return_latents: bool = False):
    """``MOTMPNet.forward`` (models/mpn.py:250-299), no autograd graph.  Returns the list of per-step logits
    [E,1] in the caller's edge order."""
    lib = _lib.load()
    _require_cuda(x, "data.x")
    _require_cuda(edge_attr, "data.edge_attr")
    check_supported(model)
    math = MATH_MODES[model.math]
    N, E = int(x.size(0)), int(edge_index.size(1))
    if edge_attr.size(0) != E:
        raise ValueError("edge_attr has %d rows but edge_index has %d edges" % (edge_attr.size(0), E))
    if plan is None:
        plan = GraphPlan(edge_index, N)          # async: narrow + order check + rowptr
    dev = x.device
    # encoder (MLPGraphIndependent.forward, mpn.py:128-142); the CSR flags are read while it runs
    h = mlp_forward_nograd(model.encoder.node_mlp, x, math)
    plan.resolve()
    e = small_mlp_rows(model.encoder.edge_mlp, plan.to_sorted(edge_attr.contiguous()))
    h0, e0 = h, e
    sb = StepBinding(model, math)
    L, k = model.num_enc_steps, model.num_class_steps
    first_class_step = L - k + 1                                            # mpn.py:277
    outs: List[torch.Tensor] = []
    latents = [(h, e)]
    if L > 0:
        P = torch.empty((N, 8), dtype=torch.float32, device=dev)
        Q = torch.empty((N, 8), dtype=torch.float32, device=dev)
        R = torch.empty((N, sb.f_node), dtype=torch.float32, device=dev)
        ws_bytes = int(lib.gnncca_mp_step_workspace_bytes(C.byref(sb.cfg), N, E))
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        keep = return_latents or model.reattach_initial_edges
        for step in range(1, L + 1):
            _lib.check(lib.gnncca_node_project(C.byref(sb.cfg), C.byref(sb.w), _lib.ptr(h0), _lib.ptr(h), N, _lib.ptr(P),
                                               _lib.ptr(Q), _lib.ptr(R), _stream()), "node_project")
            last = step == L
            classify = step >= first_class_step
            e_out = torch.empty_like(e) if (keep and (step == 1 or return_latents)) else e   # in place unless e0 must survive
            h_out = None if (last and not return_latents) else torch.empty((N, sb.f_node), dtype=torch.float32, device=dev)
            logits = torch.empty((E, 1), dtype=torch.float32, device=dev) if classify else None
            _lib.check(lib.gnncca_mp_step_fwd(C.byref(sb.cfg), C.byref(sb.w), _lib.ptr(plan.rowptr), _lib.ptr(plan.src),
                                              _lib.ptr(plan.dst), N, E, _lib.ptr(P), _lib.ptr(Q), _lib.ptr(R),
                                              _lib.ptr(e0) if model.reattach_initial_edges else 0, _lib.ptr(e), _lib.ptr(e_out),
                                              _lib.ptr(h_out), _lib.ptr(logits), _lib.ptr(ws), ws_bytes, _stream()), "mp_step_fwd")
            e = e_out
            if h_out is not None:
                h = h_out
            latents.append((h, e))
            if classify:
                outs.append(plan.to_caller(logits))
    else:
        outs.append(plan.to_caller(small_mlp_rows_cls(model, e)))            # mpn.py:295-297
    return (outs, latents, plan) if return_latents else outs


def mlp_forward_nograd(mlp, inp, math):
    out = inp
    for i, sp in enumerate(mlp.layer_specs):
        ok_tc = sp.in_dim % 32 == 0 and sp.out_dim % 16 == 0 and sp.out_dim <= 256 and sp.bn is None
        out = linear(out, mlp.linear(i), sp.relu, mlp.batchnorm(i), math if ok_tc else _lib.MATH_FP32)
    return out


def small_mlp_rows_cls(model, e):
    cls = model.classifier.edge_mlp
    if any(sp.bn is not None for sp in cls.layer_specs):
        raise NotImplementedError("num_enc_steps == 0 with a BatchNorm classifier is not implemented")
    return small_mlp_rows(cls, e)
