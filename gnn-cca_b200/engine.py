"""Host-side orchestration of the fused pipeline: which kernel runs on which buffer, in which order.
PyTorch is used for device memory (caching allocator) and streams only; all arithmetic is in libgnncca."""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional

import torch

from . import _lib
from .plan import GraphPlan, TrustedPlan, _stream

MATH_MODES = {"fp32": _lib.MATH_FP32, "tf32": _lib.MATH_TF32}


def _require_cuda(t: torch.Tensor, name: str) -> None:
    if not t.is_cuda:
        raise RuntimeError("gnn-cca_b200: %s must be a CUDA tensor; this package has no CPU or eager fallback" % name)
    if t.dtype != torch.float32:
        raise TypeError("gnn-cca_b200: %s must be float32, got %s" % (name, t.dtype))


def linear(x: torch.Tensor, lin: torch.nn.Linear, relu: bool, bn: Optional[torch.nn.BatchNorm1d] = None,
           math: int = _lib.MATH_FP32, out: Optional[torch.Tensor] = None, weight: Optional[torch.Tensor] = None) -> torch.Tensor:
    """models/mlp.py:14-19 for one fc_dims entry: Linear [+ BatchNorm1d(eval)] [+ ReLU]."""
    lib = _lib.load()
    x = x.contiguous()
    rows, in_dim = x.shape
    y = torch.empty((rows, lin.out_features), dtype=torch.float32, device=x.device) if out is None else out
    assert y.is_contiguous() and y.shape == (rows, lin.out_features)
    bnw = bnb = bnm = bnv = None
    eps = 0.0
    if bn is not None:
        bnw, bnb, bnm, bnv, eps = bn.weight, bn.bias, bn.running_mean, bn.running_var, bn.eps
    _lib.check(lib.gnncca_linear(_lib.ptr(x), _lib.ptr(lin.weight if weight is None else weight), _lib.ptr(lin.bias), _lib.ptr(y), rows, in_dim,
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
                              reattach_edges=int(model.reattach_initial_edges), math=math, edge_phase=0)
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


class StageTimer:
    """CUDA-event marks on the current stream between the stages of a forward (bench.py uses it to time the
    dominant kernel live; costs one event record per stage)."""

    def __init__(self):
        self.marks = []

    def mark(self, name: str) -> None:
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        self.marks.append((name, ev))

    def stages(self):
        """[(stage name, milliseconds)] for every interval between consecutive marks (call after a sync)."""
        return [(self.marks[i][0], self.marks[i - 1][1].elapsed_time(self.marks[i][1])) for i in range(1, len(self.marks))]


CAP_NODE_CLASSES = (32, 64, 128, 256)
CAP_EDGE_CLASSES = (512, 1024, 2048, 4096)


def frame_capacity(model, N: int, E: int, B: int, hint=None):
    """Capacity class (nodes, edges per frame) the fused all-steps kernel is launched for, or None when the frames
    are too large for one CTA.  ``hint`` = exact (max nodes, max edges) when known (our batch builder records them);
    otherwise 1.25 x the batch average, but never below the largest frame this model has met so far (a launch that
    reports an overflow makes the host measure the true maxima, see ``forward_inference``)."""
    if B <= 0:
        return None
    if hint is not None:      # exact maxima: size shared memory to the batch (more frames resident per SM)
        need_n, need_e = max(8, (int(hint[0]) + 7) & ~7), max(32, (int(hint[1]) + 31) & ~31)
        if need_n > CAP_NODE_CLASSES[-1] or need_e > (2048 if model.reattach_initial_edges else CAP_EDGE_CLASSES[-1]):
            return None
        return need_n, need_e
    else:
        seen = getattr(model, "_gnncca_caps_seen", (0, 0))
        need_n = max(int(1.25 * N / B) + 1, seen[0])
        need_e = max(int(1.25 * E / B) + 1, seen[1])
    cn = next((c for c in CAP_NODE_CLASSES if c >= need_n), None)
    ce = next((c for c in CAP_EDGE_CLASSES if c >= need_e), None)
    if cn is None or ce is None:
        return None
    if model.reattach_initial_edges and ce > 2048:
        return None
    return cn, ce


def measure_frames(graph_ptr: torch.Tensor, rowptr: torch.Tensor):
    """(max nodes, max edges) over the frames; one small synchronising read back (slow path only)."""
    gp = graph_ptr.long()
    if gp.numel() < 2:
        return 0, 0
    ok = bool(((gp >= 0) & (gp < rowptr.numel())).all()) and bool((gp[1:] >= gp[:-1]).all())
    if not ok:
        raise ValueError("data.ptr is not a non-decreasing vector of node offsets in [0, num_nodes]")
    ep = rowptr.long()[gp]
    m = torch.stack([(gp[1:] - gp[:-1]).max(), (ep[1:] - ep[:-1]).max()]).cpu()
    return int(m[0]), int(m[1])


def forward_inference(model, x: torch.Tensor, edge_index: torch.Tensor, edge_attr: torch.Tensor,
                      plan: Optional[GraphPlan] = None, return_latents: bool = False, prof: Optional[StageTimer] = None,
                      h_encoded: Optional[torch.Tensor] = None, graph_ptr: Optional[torch.Tensor] = None, caps=None,
                      allow_fused: bool = True, return_plan: bool = False, trusted: bool = False,
                      edge_ptr: Optional[torch.Tensor] = None, _no_block: bool = False, cache_on=None):
    """``MOTMPNet.forward`` (models/mpn.py:250-299), no autograd graph.  Returns the list of per-step logits
    [E,1] in the caller's edge order.

    Two device pipelines: when ``graph_ptr`` (PyG ``Batch.ptr``) is known and the frames are small, ALL steps run
    in one launch with the latents held in shared memory (gnncca_mp_fused_fwd); otherwise one fused launch per
    step (gnncca_mp_step_fwd).  Everything is enqueued optimistically -- the kernels are no-ops when the CSR
    flags say the edges are not source-major -- and the 16-byte flag read-back happens after the last launch,
    so the device never waits for the host; a batch that fails a check is then re-run on the safe path.
    ``h_encoded``: node latents already produced by the node encoder (used by ``forward_host``).
    ``trusted``: the batch comes from ``build_graph_batch`` (source-major, in range, frames closed, exact ``caps``
    by construction) or was verified by an earlier forward: nothing is read back and the call never synchronises.
    ``edge_ptr`` (trusted batches only): int64 [B+1] edge offsets of the frames; with it the fused kernel reads the
    caller's int64 ``edge_index`` directly (RAW form) and no CSR is built at all."""
    lib = _lib.load()
    _require_cuda(edge_attr, "data.edge_attr")
    check_supported(model)
    math = MATH_MODES[model.math]
    N, E = (int(x.size(0)) if h_encoded is None else int(h_encoded.size(0))), int(edge_index.size(1))
    if edge_attr.size(0) != E:
        raise ValueError("edge_attr has %d rows but edge_index has %d edges" % (edge_attr.size(0), E))
    if prof is not None:
        prof.mark("begin")
    L, k = model.num_enc_steps, model.num_class_steps
    first_class_step = L - k + 1                                            # mpn.py:277
    B = int(graph_ptr.numel()) - 1 if graph_ptr is not None else 0
    fused_ok = allow_fused and graph_ptr is not None and L >= 1 and E > 0 and edge_attr.size(1) <= 8 and first_class_step >= 1
    cap = frame_capacity(model, N, E, B, caps) if fused_ok else None
    if (cap is None and plan is None and graph_ptr is not None and not _no_block and not return_latents and h_encoded is None
            and E >= BLOCKED_MIN_EDGES and L >= 1 and first_class_step >= 1 and B > 1 and int(edge_attr.size(1)) == 4):
        # a batch of large frames whose per-edge buffers should not all be resident: blocks of whole frames (see forward_blocked)
        outs = forward_blocked(model, x, edge_index, edge_attr, graph_ptr, edge_ptr, trusted=trusted, cache_on=cache_on, prof=prof)
        return (outs, BLOCKED_PLAN) if return_plan else outs
    raw = plan is None and trusted and cap is not None and edge_ptr is not None and not return_latents
    if raw:
        plan = TrustedPlan(edge_index, N, edge_ptr)
    elif plan is None:
        plan = GraphPlan(edge_index, N)          # async: narrow + order check + rowptr
    optimistic = not plan.resolved and not trusted
    if prof is not None:
        prof.mark("csr")
    dev = edge_attr.device
    # node encoder (MLPGraphIndependent.forward, mpn.py:128-142)
    if h_encoded is None:
        _require_cuda(x, "data.x")
        h = mlp_forward_nograd(model.encoder.node_mlp, x, math)
    else:
        h = h_encoded
    if prof is not None:
        prof.mark("encoder_node")
    sb = StepBinding(model, math)
    ea_sorted = plan.to_sorted(edge_attr.contiguous())

    if cap is not None:
        gp = graph_ptr.to(device=dev, dtype=torch.int64).contiguous()
        n_cls = max(0, L - first_class_step + 1)
        logits = torch.empty((max(n_cls, 1), E), dtype=torch.float32, device=dev)
        h_dump = torch.empty((L + 1, N, sb.f_node), dtype=torch.float32, device=dev) if return_latents else None
        e_dump = torch.empty((L + 1, E, sb.f_edge), dtype=torch.float32, device=dev) if return_latents else None
        enc_e = model.encoder.edge_mlp
        if len(enc_e.layer_specs) != 1 or not enc_e.layer_specs[0].relu or enc_e.layer_specs[0].bn is not None:
            raise NotImplementedError("fused pipeline: the edge encoder must be a single Linear+ReLU (edge_fc_dims=[])")
        _lib.check(lib.gnncca_mp_fused_fwd(C.byref(sb.cfg), C.byref(sb.w), _lib.ptr(enc_e.linear(0).weight), _lib.ptr(enc_e.linear(0).bias),
                                           int(ea_sorted.size(1)), _lib.ptr(gp), B, _lib.ptr(plan.raw_edge_index), _lib.ptr(plan.edge_ptr),
                                           _lib.ptr(plan.rowptr), _lib.ptr(plan.src),
                                           _lib.ptr(plan.dst), N, E, _lib.ptr(h), _lib.ptr(ea_sorted), L, first_class_step, cap[0], cap[1],
                                           _lib.ptr(logits), _lib.ptr(h_dump), _lib.ptr(e_dump), _lib.ptr(plan.flags), _stream()),
                   "mp_fused_fwd")
        if prof is not None:
            prof.mark("mp_fused")
        ok_order = True
        if not trusted:
            plan.request_flags()
            ok_order = plan.source_major()
        if not trusted and (not ok_order or plan.frames_overflowed()):
            if ok_order:        # a frame larger than the class, or an edge leaving its frame
                plan.clear_overflow()
                plan._resolved = True
                true_caps = measure_frames(gp, plan.rowptr)
                seen = getattr(model, "_gnncca_caps_seen", (0, 0))
                model._gnncca_caps_seen = (max(seen[0], true_caps[0]), max(seen[1], true_caps[1]))
                retry = (true_caps[0] > cap[0] or true_caps[1] > cap[1]) and frame_capacity(model, N, E, B, true_caps) is not None
                return forward_inference(model, x, edge_index, edge_attr, plan, return_latents, None, h, graph_ptr,
                                         true_caps if retry else None, retry, return_plan)
            # not source-major
            plan.resolve()
            return forward_inference(model, x, edge_index, edge_attr, plan, return_latents, None, h, graph_ptr, caps, True, return_plan)
        plan._resolved = True
        plan.last_path = "fused"
        plan.caps = caps if caps is not None else cap
        outs = [plan.to_caller(logits[i].view(E, 1)) for i in range(n_cls)]
        if return_latents:
            return outs, [(h_dump[i], e_dump[i]) for i in range(L + 1)], plan
        return (outs, plan) if return_plan else outs

    # ------------------------------------------------------------------ general path: one fused launch per step
    guard = _lib.ptr(plan.flags) if optimistic else 0
    enc_e = model.encoder.edge_mlp
    # edge encoder folded into the step kernel (gnncca_mp_step_fwd_attr): Linear(4, f_edge) + ReLU, latents not requested
    fold = (L > 0 and not return_latents and len(enc_e.layer_specs) == 1 and enc_e.layer_specs[0].relu and enc_e.layer_specs[0].bn is None
            and int(ea_sorted.size(1)) == 4 and ea_sorted.is_contiguous() and ea_sorted.data_ptr() % 16 == 0)
    e = None if fold else small_mlp_rows(enc_e, ea_sorted)
    if prof is not None:
        prof.mark("encoder_edge")
    h0, e0 = h, e
    outs: List[torch.Tensor] = []
    latents = [(h, e)]
    if L > 0:
        P = torch.empty((N, 8), dtype=torch.float32, device=dev)
        Q = torch.empty((N, 8), dtype=torch.float32, device=dev)
        R = torch.empty((N, sb.f_node), dtype=torch.float32, device=dev)
        ws_bytes = int(lib.gnncca_mp_step_workspace_bytes(C.byref(sb.cfg), N, E))
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        keep = return_latents or model.reattach_initial_edges
        le = enc_e.linear(0) if fold else None
        for step in range(1, L + 1):
            if step == 1:      # later steps: the previous step's fold launch has already projected its output (gnncca_mp_step_fwd_next)
                _lib.check(lib.gnncca_node_project(C.byref(sb.cfg), C.byref(sb.w), _lib.ptr(h0), _lib.ptr(h), N, _lib.ptr(P),
                                                   _lib.ptr(Q), _lib.ptr(R), _stream()), "node_project")
            last = step == L
            classify = step >= first_class_step
            h_out = None if (last and not return_latents) else torch.empty((N, sb.f_node), dtype=torch.float32, device=dev)
            logits = torch.empty((E, 1), dtype=torch.float32, device=dev) if classify else None
            dead_e = last and h_out is None and logits is not None       # e^L of an inference forward is never read
            use_attr = fold and (step == 1 or model.reattach_initial_edges)
            if use_attr:
                e_out = None if dead_e else (torch.empty((E, sb.f_edge), dtype=torch.float32, device=dev) if step == 1 else e)
            else:
                e_out = None if dead_e else (torch.empty_like(e) if (keep and (step == 1 or return_latents)) else e)   # in place unless e0 must survive
            if not last:       # the step + the projections of step + 1, in place over P / Q / R
                _lib.check(lib.gnncca_mp_step_fwd_next(C.byref(sb.cfg), C.byref(sb.w), _lib.ptr(plan.rowptr), _lib.ptr(plan.src), _lib.ptr(plan.dst), N, E,
                                                       _lib.ptr(P), _lib.ptr(Q), _lib.ptr(R), _lib.ptr(ea_sorted) if use_attr else 0, 4,
                                                       _lib.ptr(le.weight) if use_attr else 0, _lib.ptr(le.bias) if use_attr else 0,
                                                       _lib.ptr(e0) if (model.reattach_initial_edges and not use_attr) else 0, _lib.ptr(e), _lib.ptr(e_out),
                                                       _lib.ptr(h_out), _lib.ptr(logits), _lib.ptr(h0), _lib.ptr(P), _lib.ptr(Q), _lib.ptr(R), guard,
                                                       _lib.ptr(ws), ws_bytes, _stream()), "mp_step_fwd_next")
            elif use_attr:
                _lib.check(lib.gnncca_mp_step_fwd_attr(C.byref(sb.cfg), C.byref(sb.w), _lib.ptr(plan.rowptr), _lib.ptr(plan.src),
                                                       _lib.ptr(plan.dst), N, E, _lib.ptr(P), _lib.ptr(Q), _lib.ptr(R), _lib.ptr(ea_sorted), 4,
                                                       _lib.ptr(le.weight), _lib.ptr(le.bias), _lib.ptr(e), _lib.ptr(e_out), _lib.ptr(h_out),
                                                       _lib.ptr(logits), guard, _lib.ptr(ws), ws_bytes, _stream()), "mp_step_fwd_attr")
            else:
                _lib.check(lib.gnncca_mp_step_fwd(C.byref(sb.cfg), C.byref(sb.w), _lib.ptr(plan.rowptr), _lib.ptr(plan.src),
                                                  _lib.ptr(plan.dst), N, E, _lib.ptr(P), _lib.ptr(Q), _lib.ptr(R),
                                                  _lib.ptr(e0) if model.reattach_initial_edges else 0, _lib.ptr(e), _lib.ptr(e_out),
                                                  _lib.ptr(h_out), _lib.ptr(logits), guard, _lib.ptr(ws), ws_bytes, _stream()),
                           "mp_step_fwd")
            e = e_out
            if h_out is not None:
                h = h_out
            latents.append((h, e))
            if classify:
                outs.append(plan.to_caller(logits))
            if prof is not None:
                prof.mark("step%d" % step)
    else:
        outs.append(plan.to_caller(small_mlp_rows_cls(model, e)))            # mpn.py:295-297
    if optimistic:
        plan.request_flags()
        if not plan.source_major():
            plan.resolve()   # stable sort by source; outputs go back to the caller's edge order
            return forward_inference(model, x, edge_index, edge_attr, plan, return_latents, None, h0, graph_ptr, caps, allow_fused, return_plan)
    plan._resolved = True
    plan.last_path = "steps"
    if return_latents:
        return outs, latents, plan
    return (outs, plan) if return_plan else outs


# ------------------------------------------------------------------------------------------------------------
# stand-alone MetaLayer / NodeModel forwards (models/mpn.py:32-54, 97-101) on caller-supplied tensors
# ------------------------------------------------------------------------------------------------------------
def _inference_only(module, what: str) -> None:
    if torch.is_grad_enabled() and any(p.requires_grad for p in module.parameters()):
        raise NotImplementedError("stand-alone %s.forward is inference-only (use torch.no_grad()); training runs through MOTMPNet.forward" % what)


def _agg_name(node_model) -> str:
    name = node_model.node_agg_fn
    if not isinstance(name, str) or name not in _lib.AGG:
        raise NotImplementedError("NodeModel.node_agg_fn must be 'sum', 'mean' or 'max' (the reference passes a lambda over torch_scatter; "
                                  "gnn-cca_b200 takes the name, models/mpn.py:192-202)")
    return name


def meta_layer_forward(meta, x: torch.Tensor, edge_index: torch.Tensor, edge_attr: torch.Tensor):
    """``MetaLayer.forward`` (mpn.py:32-54): edge model on the old nodes, node model on the new edges, aggregation by ``row``."""
    lib = _lib.load()
    _require_cuda(x, "x")
    _require_cuda(edge_attr, "edge_attr")
    _inference_only(meta, "MetaLayer")
    em, nm = meta.edge_model.edge_mlp, meta.node_model.node_mlp
    if len(em.layer_specs) != 1 or len(nm.layer_specs) != 1 or not em.layer_specs[0].relu or not nm.layer_specs[0].relu \
            or em.layer_specs[0].bn is not None or nm.layer_specs[0].bn is not None:
        raise NotImplementedError("fused step: edge_model / node_model must be one Linear + ReLU each")
    f_edge, f_node = em.layer_specs[0].out_dim, nm.layer_specs[0].out_dim
    K, EIN = int(x.size(1)), int(edge_attr.size(1))
    if (f_node, f_edge) != (32, 6) or K not in (32, 64) or EIN not in (6, 12):
        raise NotImplementedError("fused step kernels are built for node / edge latent widths 32 / 6 (x: 32 or 64 columns, edge_attr: 6 or 12)")
    if em.layer_specs[0].in_dim != 2 * K + EIN or nm.layer_specs[0].in_dim != K + f_edge:
        raise ValueError("input widths do not match the update MLPs: x %d, edge_attr %d vs edge_mlp in %d, node_mlp in %d"
                         % (K, EIN, em.layer_specs[0].in_dim, nm.layer_specs[0].in_dim))
    N, E = int(x.size(0)), int(edge_index.size(1))
    if int(edge_attr.size(0)) != E:
        raise ValueError("edge_attr has %d rows but edge_index has %d edges" % (edge_attr.size(0), E))
    dev = x.device
    rn, re = K == 64, EIN == 12
    cfg = _lib.MpCfg(f_node=32, f_edge=6, cls_hidden=4, agg=_lib.AGG[_agg_name(meta.node_model)], reattach_nodes=int(rn), reattach_edges=int(re),
                     math=_lib.MATH_FP32, edge_phase=0)
    w = _lib.MpWeights(w_edge=em.linear(0).weight.data_ptr(), b_edge=em.linear(0).bias.data_ptr(), w_node=nm.linear(0).weight.data_ptr(),
                       b_node=nm.linear(0).bias.data_ptr(), w_c1=0, b_c1=0, bn_weight=0, bn_bias=0, bn_mean=0, bn_var=0, w_c2=0, b_c2=0, bn_eps=0.0, pad_=0)
    plan = GraphPlan(edge_index, N).resolve()
    h0 = x[:, :32].contiguous() if rn else None
    h = x[:, 32:].contiguous() if rn else x.contiguous()
    ea = plan.to_sorted(edge_attr.contiguous())
    e0 = ea[:, :6].contiguous() if re else None
    e = ea[:, 6:].contiguous() if re else ea
    P = torch.empty((N, 8), dtype=torch.float32, device=dev)
    Q = torch.empty((N, 8), dtype=torch.float32, device=dev)
    R = torch.empty((N, 32), dtype=torch.float32, device=dev)
    e_out = torch.empty((E, 6), dtype=torch.float32, device=dev)
    h_out = torch.empty((N, 32), dtype=torch.float32, device=dev)
    if N > 0:
        _lib.check(lib.gnncca_node_project(C.byref(cfg), C.byref(w), _lib.ptr(h0), _lib.ptr(h), N, _lib.ptr(P), _lib.ptr(Q), _lib.ptr(R), _stream()),
                   "node_project")
    ws_bytes = int(lib.gnncca_mp_step_workspace_bytes(C.byref(cfg), N, E))
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    _lib.check(lib.gnncca_mp_step_fwd(C.byref(cfg), C.byref(w), _lib.ptr(plan.rowptr), _lib.ptr(plan.src), _lib.ptr(plan.dst), N, E, _lib.ptr(P),
                                      _lib.ptr(Q), _lib.ptr(R), _lib.ptr(e0), _lib.ptr(e), _lib.ptr(e_out), _lib.ptr(h_out), 0, 0, _lib.ptr(ws),
                                      ws_bytes, _stream()), "mp_step_fwd")
    return h_out, plan.to_caller(e_out)


def node_model_forward(node_model, x: torch.Tensor, edge_index: torch.Tensor, edge_attr: torch.Tensor) -> torch.Tensor:
    """``NodeModel.forward`` (mpn.py:97-101)."""
    lib = _lib.load()
    _require_cuda(x, "x")
    _require_cuda(edge_attr, "edge_attr")
    _inference_only(node_model, "NodeModel")
    agg = _lib.AGG[_agg_name(node_model)]
    N = int(x.size(0))
    plan = GraphPlan(edge_index, N).resolve()
    row = plan.src.long()                                     # sorted sources
    msgs = mlp_forward(node_model.node_mlp, torch.cat([x.index_select(0, row), plan.to_sorted(edge_attr.contiguous())], dim=1))
    out = torch.empty((N, msgs.size(1)), dtype=torch.float32, device=x.device)
    _lib.check(lib.gnncca_segment_agg(_lib.ptr(msgs), _lib.ptr(plan.rowptr), N, int(msgs.size(1)), agg, _lib.ptr(out), _stream()), "segment_agg")
    return out


# ------------------------------------------------------------------------------------------------------------
# frame-blocked execution of very large batches (BASELINE config 4: 4096 frames of 8 x 64 detections, 939.5 M edges)
# ------------------------------------------------------------------------------------------------------------
BLOCK_EDGES = int(os.environ.get("GNNCCA_BLOCK_EDGES", str(32 << 20)))     # target edges per block of frames
BLOCKED_MIN_EDGES = int(os.environ.get("GNNCCA_BLOCKED_MIN_EDGES", str(24 << 20)))   # batches above this run blocked


def frame_blocks(graph_ptr_host, edge_ptr_host, target_edges: int):
    """Cut the frames [0, B) into contiguous blocks of about ``target_edges`` edges each (whole frames; sizes as even as
    the frames allow).  Returns [(g0, g1, n0, n1, e0, e1)].  Pure host arithmetic on the two offset vectors."""
    B = len(graph_ptr_host) - 1
    E = int(edge_ptr_host[B]) - int(edge_ptr_host[0])
    if B <= 0:
        return []
    nb = max(1, min(B, -(-E // max(1, target_edges))))
    blocks, g0 = [], 0
    for b in range(1, nb + 1):
        if b == nb:
            g1 = B
        else:
            goal = int(edge_ptr_host[0]) + (E * b) // nb
            lo, hi = g0 + 1, B - (nb - b)         # leave at least one frame for every later block
            while lo < hi:                        # first frame boundary at or beyond the goal
                mid = (lo + hi) // 2
                if int(edge_ptr_host[mid]) < goal:
                    lo = mid + 1
                else:
                    hi = mid
            g1 = lo
        blocks.append((g0, g1, int(graph_ptr_host[g0]), int(graph_ptr_host[g1]), int(edge_ptr_host[g0]), int(edge_ptr_host[g1])))
        g0 = g1
    return blocks


class BlockWorkspace:
    """Per-edge / per-node work buffers of ONE block of frames, allocated once for the largest block and reused by every
    block of every forward on that batch shape (the caching allocator would hand the same memory back anyway; keeping
    them removes ~15 allocator calls per block from the host path)."""

    def __init__(self, lib, sb, max_nodes: int, max_edges: int, n_blocks: int, dev, f_node: int, f_edge: int, reattach_e: bool):
        i32, f32 = torch.int32, torch.float32
        self.max_nodes, self.max_edges, self.n_blocks = max_nodes, max_edges, n_blocks
        self.src = torch.empty(max_edges, dtype=i32, device=dev)
        self.dst = torch.empty(max_edges, dtype=i32, device=dev)
        self.rowptr = torch.empty(max_nodes + 1, dtype=i32, device=dev)
        self.flags = torch.zeros((n_blocks, _lib.CSR_NFLAGS), dtype=i32, device=dev)
        self.h = [torch.empty((max_nodes, f_node), dtype=f32, device=dev) for _ in range(3)]
        self.P = torch.empty((max_nodes, 8), dtype=f32, device=dev)
        self.Q = torch.empty((max_nodes, 8), dtype=f32, device=dev)
        self.R = torch.empty((max_nodes, f_node), dtype=f32, device=dev)
        self.e = torch.empty((max_edges, f_edge), dtype=f32, device=dev)
        self.ws_bytes = int(lib.gnncca_mp_step_workspace_bytes(C.byref(sb.cfg), max_nodes, max_edges))
        self.ws = torch.empty(self.ws_bytes, dtype=torch.uint8, device=dev)

    def fits(self, max_nodes, max_edges, n_blocks):
        return max_nodes <= self.max_nodes and max_edges <= self.max_edges and n_blocks <= self.n_blocks


def forward_blocked(model, x: torch.Tensor, edge_index: torch.Tensor, edge_attr: torch.Tensor, graph_ptr: torch.Tensor,
                    edge_ptr: Optional[torch.Tensor] = None, trusted: bool = False, target_edges: Optional[int] = None,
                    cache_on=None, prof=None):
    """``MOTMPNet.forward`` (models/mpn.py:250-299) for a batch whose per-edge work buffers should not (or cannot: the step
    kernel indexes edges with int32) be resident for all E edges at once.  A PyG ``Batch`` is a disjoint union of frames
    (train.py:450) and nothing in the eval-mode forward mixes frames, so blocks of whole frames run one after the other
    through the per-step pipeline (CSR narrowing -> node encoder -> L x [node_project, mp_step, fold]) on ONE reused
    workspace; each block's logits land in its slice of the batch-wide outputs.  Results are bit-identical to the
    unblocked per-step forward: a block is tiled like the whole batch (``edge_phase`` = position of its first edge inside
    a reduction tile of the batch-wide edge numbering), so every segment is summed in the same order
    (tests/test_gpu_blocked.py).

    ``graph_ptr`` [B+1] node offsets (Batch.ptr); ``edge_ptr`` [B+1] edge offsets when known (``build_graph_batch``),
    otherwise found by a binary search in ``edge_index[0]`` (valid for source-major input, which each block verifies).
    One small read back (the two offset vectors) the first time a batch is seen; untrusted input costs one more 16-byte
    flag read per forward after the last launch."""
    lib = _lib.load()
    _require_cuda(x, "data.x")
    _require_cuda(edge_attr, "data.edge_attr")
    check_supported(model)
    if edge_index.dim() != 2 or edge_index.size(0) != 2 or edge_index.dtype != torch.int64 or not edge_index.is_cuda:
        raise ValueError("edge_index must be a CUDA int64 tensor of shape [2, E]")
    math = MATH_MODES[model.math]
    dev = x.device
    x, edge_index, edge_attr = x.contiguous(), edge_index.contiguous(), edge_attr.contiguous()
    N, E = int(x.size(0)), int(edge_index.size(1))
    L, k = model.num_enc_steps, model.num_class_steps
    first = L - k + 1
    if L < 1 or first < 1:
        raise NotImplementedError("blocked forward needs num_enc_steps >= 1 and num_class_steps <= num_enc_steps")
    enc_e = model.encoder.edge_mlp
    if not (len(enc_e.layer_specs) == 1 and enc_e.layer_specs[0].relu and enc_e.layer_specs[0].bn is None and int(edge_attr.size(1)) == 4):
        raise NotImplementedError("blocked forward: the edge encoder must be the single Linear(4, f_edge) + ReLU")
    target = int(target_edges) if target_edges else BLOCK_EDGES

    cached = getattr(cache_on, "_gnncca_blocks", None) if cache_on is not None else None
    key = (graph_ptr.data_ptr(), graph_ptr._version, edge_index.data_ptr(), edge_index._version, N, E, target)
    if cached is not None and cached[0] == key:
        blocks = cached[1]
    else:
        gp = graph_ptr.to(device=dev, dtype=torch.int64)
        if edge_ptr is None:
            edge_ptr = torch.searchsorted(edge_index[0].contiguous(), gp)      # first edge whose source is >= the frame's first node
        host = torch.stack([gp, edge_ptr.to(device=dev, dtype=torch.int64)]).cpu()
        gph, eph = host[0].tolist(), host[1].tolist()
        if gph[0] != 0 or gph[-1] != N or eph[0] != 0 or eph[-1] != E or any(b < a for a, b in zip(gph, gph[1:])):
            raise ValueError("data.ptr / edge offsets do not cover the batch (ptr must run from 0 to num_nodes, non-decreasing)")
        blocks = frame_blocks(gph, eph, target)
        if cache_on is not None:
            try:
                cache_on._gnncca_blocks = (key, blocks)
            except AttributeError:
                pass
    max_n = max((b[3] - b[2]) for b in blocks)
    max_e = max((b[5] - b[4]) for b in blocks)
    if max_e >= (1 << 31) - 4096:
        raise NotImplementedError("a single frame block of %d edges exceeds the step kernel's int32 edge index" % max_e)
    sb = StepBinding(model, math)
    wsp = model.__dict__.get("_gnncca_block_ws")
    if wsp is None or not wsp.fits(max_n, max_e, len(blocks)) or wsp.ws.device != dev:
        wsp = BlockWorkspace(lib, sb, max_n, max_e, len(blocks), dev, sb.f_node, sb.f_edge, bool(model.reattach_initial_edges))
        model.__dict__["_gnncca_block_ws"] = wsp
    logits = torch.empty((k, E), dtype=torch.float32, device=dev)
    le = enc_e.linear(0)
    row0, col0 = edge_index.data_ptr(), edge_index.data_ptr() + 8 * E
    st = _stream()
    cfg, w = C.byref(sb.cfg), C.byref(sb.w)
    enc_n = model.encoder.node_mlp
    tile = int(lib.gnncca_mp_step_tile())
    if not trusted:
        wsp.flags.zero_()
    for bi, (g0, g1, n0, n1, e0, e1) in enumerate(blocks):
        nn_, ne_ = n1 - n0, e1 - e0
        if ne_ == 0 or nn_ == 0:
            continue                # frames without cross-camera edges: no logits to write
        sb.cfg.edge_phase = e0 % tile
        flags_b = wsp.flags[bi]
        guard = 0 if trusted else flags_b.data_ptr()
        _lib.check(lib.gnncca_csr_from_edge_rows(row0 + 8 * e0, col0 + 8 * e0, n0, nn_, ne_, _lib.ptr(wsp.src), _lib.ptr(wsp.dst),
                                                 _lib.ptr(wsp.rowptr), flags_b.data_ptr(), st), "csr_from_edge_rows")
        h = mlp_forward_nograd(enc_n, x[n0:n1], math, out=wsp.h[0][:nn_])
        h0 = h
        ea = edge_attr[e0:e1]
        hbuf = 1
        _lib.check(lib.gnncca_node_project(cfg, w, h0.data_ptr(), h.data_ptr(), nn_, _lib.ptr(wsp.P), _lib.ptr(wsp.Q), _lib.ptr(wsp.R), st), "node_project")
        for step in range(1, L + 1):
            last = step == L
            h_out = None if last else wsp.h[hbuf][:nn_]
            lg = logits[step - first][e0:e1] if step >= first else None
            dead_e = last and lg is not None
            use_attr = step == 1 or model.reattach_initial_edges
            e_in = None if step == 1 else wsp.e
            if not last:      # the step + (in its fold launch) the projections of step + 1, in place over P / Q / R
                _lib.check(lib.gnncca_mp_step_fwd_next(cfg, w, _lib.ptr(wsp.rowptr), _lib.ptr(wsp.src), _lib.ptr(wsp.dst), nn_, ne_,
                                                       _lib.ptr(wsp.P), _lib.ptr(wsp.Q), _lib.ptr(wsp.R), ea.data_ptr() if use_attr else 0, 4,
                                                       _lib.ptr(le.weight) if use_attr else 0, _lib.ptr(le.bias) if use_attr else 0, 0,
                                                       _lib.ptr(e_in), _lib.ptr(wsp.e), _lib.ptr(h_out), _lib.ptr(lg), h0.data_ptr(),
                                                       _lib.ptr(wsp.P), _lib.ptr(wsp.Q), _lib.ptr(wsp.R), guard, _lib.ptr(wsp.ws), wsp.ws_bytes, st),
                           "mp_step_fwd_next")
            elif use_attr:
                _lib.check(lib.gnncca_mp_step_fwd_attr(cfg, w, _lib.ptr(wsp.rowptr), _lib.ptr(wsp.src), _lib.ptr(wsp.dst), nn_, ne_,
                                                       _lib.ptr(wsp.P), _lib.ptr(wsp.Q), _lib.ptr(wsp.R), ea.data_ptr(), 4, _lib.ptr(le.weight), _lib.ptr(le.bias),
                                                       _lib.ptr(e_in), 0 if dead_e else _lib.ptr(wsp.e), _lib.ptr(h_out), _lib.ptr(lg), guard,
                                                       _lib.ptr(wsp.ws), wsp.ws_bytes, st), "mp_step_fwd_attr")
            else:
                _lib.check(lib.gnncca_mp_step_fwd(cfg, w, _lib.ptr(wsp.rowptr), _lib.ptr(wsp.src), _lib.ptr(wsp.dst), nn_, ne_,
                                                  _lib.ptr(wsp.P), _lib.ptr(wsp.Q), _lib.ptr(wsp.R), 0, _lib.ptr(wsp.e), 0 if dead_e else _lib.ptr(wsp.e),
                                                  _lib.ptr(h_out), _lib.ptr(lg), guard, _lib.ptr(wsp.ws), wsp.ws_bytes, st), "mp_step_fwd")
            if h_out is not None:
                h = h_out
                hbuf = 2 if hbuf == 1 else 1
    if not trusted:      # one read back for all blocks, after the last launch
        bad = wsp.flags[:len(blocks)].amax(dim=0).cpu().tolist()
        if bad[1]:
            raise IndexError("edge_index has entries outside their frame block's node range (frames must be closed; ptr must match)")
        if bad[0]:
            if E < (1 << 31) - 4096:
                return forward_inference(model, x, edge_index, edge_attr, None, graph_ptr=graph_ptr, allow_fused=False, _no_block=True)
            raise NotImplementedError("edge_index is not source-major and the batch is too large for the sorting path")
    return [logits[i].view(E, 1) for i in range(k)]


def forward_host(model, x: torch.Tensor, edge_index: torch.Tensor, edge_attr: torch.Tensor, out: Optional[List[torch.Tensor]] = None,
                 device: Optional[torch.device] = None, n_chunks: int = 8, graph_ptr: Optional[torch.Tensor] = None,
                 caps=None) -> List[torch.Tensor]:
    """The same forward for HOST buffers (pinned memory recommended): the inputs are copied to the device on a
    side stream in row chunks, the node encoder of chunk i runs while chunk i+1 is in flight, and the per-step
    logits are copied back into ``out`` (host tensors [E,1]; allocated pinned when None).  Returns ``out``
    after the copies have completed."""
    if x.is_cuda or edge_index.is_cuda or edge_attr.is_cuda:
        raise ValueError("forward_host takes host tensors; call the module itself for device tensors")
    check_supported(model)
    dev = device if device is not None else next(model.parameters()).device
    if dev.type != "cuda":
        raise RuntimeError("gnn-cca_b200: the model must live on a CUDA device (no CPU path)")
    math = MATH_MODES[model.math]
    N, E = int(x.size(0)), int(edge_index.size(1))
    cur = torch.cuda.current_stream(dev)
    side = _side_stream(dev)
    side.wait_stream(cur)
    x_d = torch.empty((N, x.size(1)), dtype=torch.float32, device=dev)
    ei_d = torch.empty((2, E), dtype=torch.int64, device=dev)
    ea_d = torch.empty((E, edge_attr.size(1)), dtype=torch.float32, device=dev)
    bounds = [N * i // n_chunks for i in range(n_chunks + 1)]
    events = []
    with torch.cuda.stream(side):
        ei_d.copy_(edge_index, non_blocking=True)
        ea_d.copy_(edge_attr, non_blocking=True)
        ev0 = torch.cuda.Event(); ev0.record(side)
        for a, b in zip(bounds[:-1], bounds[1:]):
            if b > a:
                x_d[a:b].copy_(x[a:b], non_blocking=True)
            ev = torch.cuda.Event(); ev.record(side)
            events.append(ev)
    for t in (x_d, ei_d, ea_d):
        t.record_stream(side)
    cur.wait_event(ev0)
    plan = GraphPlan(ei_d, N)
    enc = model.encoder.node_mlp
    h = torch.empty((N, enc.layer_specs[-1].out_dim), dtype=torch.float32, device=dev)
    for (a, b), ev in zip(zip(bounds[:-1], bounds[1:]), events):
        cur.wait_event(ev)
        if b > a:
            mlp_forward_nograd(enc, x_d[a:b], math, out=h[a:b])
    if graph_ptr is not None and not graph_ptr.is_cuda:
        graph_ptr = graph_ptr.to(dev, non_blocking=True)
    outs = forward_inference(model, x_d, ei_d, ea_d, plan, h_encoded=h, graph_ptr=graph_ptr, caps=caps)
    if out is None:
        out = [torch.empty((E, 1), dtype=torch.float32, pin_memory=True) for _ in outs]
    for o, d in zip(out, outs):
        o.copy_(d, non_blocking=True)
    cur.synchronize()
    return out


class HostPipeline:
    """``forward_host`` for a STREAM of host batches: two sets of device input buffers used alternately, so that the
    host->device copy of batch i+1 (side stream) runs while batch i is still computing and while its logits travel back
    (PCIe is full duplex).  ``submit`` returns at once with a handle; ``handle.wait()`` returns the host logits.

        pipe = gnncca.HostPipeline(model)
        pending = None
        for x, edge_index, edge_attr, ptr in batches:            # pinned host tensors
            h = pipe.submit(x, edge_index, edge_attr, graph_ptr=ptr, out=out_bufs[i % 2])
            if pending is not None:
                logits = pending.wait()
            pending = h
    """

    class Handle:
        def __init__(self, out, event):
            self.out, self.event = out, event

        def wait(self):
            self.event.synchronize()
            return self.out

    def __init__(self, model, n_chunks: int = 8, slots: int = 2):
        check_supported(model)
        self.model, self.n_chunks = model, n_chunks
        self.dev = next(model.parameters()).device
        if self.dev.type != "cuda":
            raise RuntimeError("gnn-cca_b200: the model must live on a CUDA device (no CPU path)")
        self.slots = [dict(key=None, free=None) for _ in range(slots)]
        self.turn = 0
        self.side = torch.cuda.Stream(device=self.dev)

    def submit(self, x, edge_index, edge_attr, out=None, graph_ptr=None, caps=None) -> "HostPipeline.Handle":
        if x.is_cuda or edge_index.is_cuda or edge_attr.is_cuda:
            raise ValueError("HostPipeline takes host tensors; call the module itself for device tensors")
        model, dev = self.model, self.dev
        math = MATH_MODES[model.math]
        N, E = int(x.size(0)), int(edge_index.size(1))
        cur = torch.cuda.current_stream(dev)
        slot = self.slots[self.turn % len(self.slots)]
        self.turn += 1
        key = (N, int(x.size(1)), E, int(edge_attr.size(1)))
        enc = model.encoder.node_mlp
        if slot["key"] != key:          # (re)allocate this slot's device buffers; ordered after whatever still uses the old ones
            if slot["free"] is not None:
                slot["free"].synchronize()
            slot.update(key=key, x=torch.empty((N, key[1]), dtype=torch.float32, device=dev), ei=torch.empty((2, E), dtype=torch.int64, device=dev),
                        ea=torch.empty((E, key[3]), dtype=torch.float32, device=dev),
                        h=torch.empty((N, enc.layer_specs[-1].out_dim), dtype=torch.float32, device=dev))
            self.side.wait_stream(cur)
        side = self.side
        if slot["free"] is not None:
            side.wait_event(slot["free"])       # the forward that last read these buffers has finished
        bounds = [N * i // self.n_chunks for i in range(self.n_chunks + 1)]
        events = []
        with torch.cuda.stream(side):
            slot["ei"].copy_(edge_index, non_blocking=True)
            slot["ea"].copy_(edge_attr, non_blocking=True)
            gp = graph_ptr.to(dev, non_blocking=True) if (graph_ptr is not None and not graph_ptr.is_cuda) else graph_ptr
            ev0 = torch.cuda.Event(); ev0.record(side)
            for a, b in zip(bounds[:-1], bounds[1:]):
                if b > a:
                    slot["x"][a:b].copy_(x[a:b], non_blocking=True)
                ev = torch.cuda.Event(); ev.record(side)
                events.append(ev)
        if gp is not None:
            gp.record_stream(cur)
        cur.wait_event(ev0)
        plan = GraphPlan(slot["ei"], N)
        for (a, b), ev in zip(zip(bounds[:-1], bounds[1:]), events):
            cur.wait_event(ev)
            if b > a:
                mlp_forward_nograd(enc, slot["x"][a:b], math, out=slot["h"][a:b])
        outs = forward_inference(model, slot["x"], slot["ei"], slot["ea"], plan, h_encoded=slot["h"], graph_ptr=gp, caps=caps)
        if out is None:
            out = [torch.empty((E, 1), dtype=torch.float32, pin_memory=True) for _ in outs]
        for o, d in zip(out, outs):
            o.copy_(d, non_blocking=True)
        done = torch.cuda.Event(); done.record(cur)
        slot["free"] = done
        return HostPipeline.Handle(out, done)


FAST_PLAN = TrustedPlan.__new__(TrustedPlan)     # what ``last_plan`` reports after the single-call path
FAST_PLAN.last_path = "fused"
FAST_PLAN.caps = None
BLOCKED_PLAN = TrustedPlan.__new__(TrustedPlan)  # ... and after the frame-blocked path (no batch-wide CSR exists)
BLOCKED_PLAN.last_path = "blocked"
BLOCKED_PLAN.caps = None

_SIDE_STREAMS = {}


def _side_stream(dev) -> "torch.cuda.Stream":
    key = (dev.type, dev.index if dev.index is not None else torch.cuda.current_device())
    if key not in _SIDE_STREAMS:
        _SIDE_STREAMS[key] = torch.cuda.Stream(device=dev)
    return _SIDE_STREAMS[key]


CACHE_PLANS = True          # a verified CSR is cached on the Batch object and reused while edge_index is unchanged
FUSED_NODE_ENCODER = True   # bench.py: the node encoder is one kernel in tf32 mode (algorithmic bytes N*(4K+128))


def node_encoder_fused_ok(mlp, inp) -> bool:
    """Shapes the tcgen05 encoder kernel covers: Linear+ReLU, Linear+ReLU, in % 32 == 0, hidden in {64,128,256}, out 32."""
    sp = mlp.layer_specs
    return (len(sp) == 2 and all(s.relu and s.bn is None for s in sp) and sp[0].in_dim % 32 == 0 and sp[0].out_dim in (64, 128)
            and sp[1].out_dim == 32 and inp.data_ptr() % 16 == 0)


_ENC_SCRATCH = {}


def encoder_scratch_bytes(rows: int, in_dim: int, hidden: int) -> int:
    """Bytes of scratch the node encoder's split-K form wants for ``rows`` rows (0: the one-launch form is the right one)."""
    key = (int(rows), int(in_dim), int(hidden))
    v = _ENC_SCRATCH.get(key)
    if v is None:
        if len(_ENC_SCRATCH) > 256:
            _ENC_SCRATCH.clear()
        v = _ENC_SCRATCH[key] = int(_lib.load().gnncca_node_encoder_scratch_bytes(*key))
    return v


def mlp_forward_nograd(mlp, inp, math, out=None):
    w0 = mlp.__dict__.get("_gnncca_w0")      # first-layer weight with the input's column normalisation folded in (MOTMPNet.forward)
    if math == _lib.MATH_TF32 and node_encoder_fused_ok(mlp, inp):
        lib = _lib.load()
        inp = inp.contiguous()
        rows = int(inp.size(0))
        l1, l2 = mlp.linear(0), mlp.linear(1)
        y = torch.empty((rows, l2.out_features), dtype=torch.float32, device=inp.device) if out is None else out
        assert y.is_contiguous() and y.shape == (rows, l2.out_features)
        # (inference keeps the one-launch form at every size: for small batches the split-K form of gnncca_node_encoder_ws is 3x faster
        # as a kernel, but the all-steps kernel that follows overlaps the one-launch encoder's tail as its programmatic dependent, and
        # end to end the two measure the same (cfg 1: 52 us either way); the training forward, which has no such overlap, uses it)
        _lib.check(lib.gnncca_node_encoder(_lib.ptr(inp), _lib.ptr(l1.weight if w0 is None else w0), _lib.ptr(l1.bias), _lib.ptr(l2.weight), _lib.ptr(l2.bias),
                                           _lib.ptr(y), rows, l1.in_features, l1.out_features, l2.out_features, _lib.MATH_TF32,
                                           _stream()), "node_encoder")
        return y
    math = _lib.MATH_FP32
    cur = inp
    n = len(mlp.layer_specs)
    for i, sp in enumerate(mlp.layer_specs):
        ok_tc = sp.in_dim % 32 == 0 and sp.out_dim % 16 == 0 and sp.out_dim <= 256 and sp.bn is None
        cur = linear(cur, mlp.linear(i), sp.relu, mlp.batchnorm(i), math if ok_tc else _lib.MATH_FP32,
                     out=out if i == n - 1 else None, weight=w0 if i == 0 else None)
    return cur


def small_mlp_rows_cls(model, e):
    cls = model.classifier.edge_mlp
    if any(sp.bn is not None for sp in cls.layer_specs):
        raise NotImplementedError("num_enc_steps == 0 with a BatchNorm classifier is not implemented")
    return small_mlp_rows(cls, e)


class FastPath:
    """Everything ``MOTMPNet.forward`` needs to run a trusted batch (``build_graph_batch``) with ONE library call
    (gnncca_mpn_forward_frames): the ctypes views of the configuration and of the live parameter pointers, built once
    per model and re-validated per call by comparing the parameters' device pointers."""

    def __init__(self, model):
        check_supported(model)
        enc_n, enc_e = model.encoder.node_mlp, model.encoder.edge_mlp
        sp = enc_n.layer_specs
        if not (len(sp) == 2 and all(s.relu and s.bn is None for s in sp) and sp[0].in_dim % 32 == 0 and sp[0].out_dim in (64, 128)
                and sp[1].out_dim == 32):
            raise NotImplementedError("node encoder shape without a fused kernel")
        if len(enc_e.layer_specs) != 1 or not enc_e.layer_specs[0].relu or enc_e.layer_specs[0].bn is not None:
            raise NotImplementedError("edge encoder must be a single Linear+ReLU")
        self.L, k = model.num_enc_steps, model.num_class_steps
        self.first = self.L - k + 1
        if self.L < 1 or self.first < 1 or model.math != "tf32":
            raise NotImplementedError("fast path needs L >= 1, k <= L and tf32 math")
        self.n_cls = self.L - self.first + 1
        self.fa = enc_e.layer_specs[0].in_dim
        # (dict, name) of every parameter / buffer slot of the module tree: a re-assigned nn.Parameter is seen too
        self.slots = [(d, n) for m in model.modules() for d in (m._parameters, m._buffers) for n, t in d.items() if t is not None]
        self.key = tuple(d[n].data_ptr() for d, n in self.slots)
        self.sb = StepBinding(model, _lib.MATH_TF32)
        l1, l2, le = enc_n.linear(0), enc_n.linear(1), enc_e.linear(0)
        self.prefix = (C.byref(self.sb.cfg), C.byref(self.sb.w), l1.weight.data_ptr(), l1.bias.data_ptr(), l2.weight.data_ptr(),
                       l2.bias.data_ptr(), l1.in_features, l1.out_features, le.weight.data_ptr(), le.bias.data_ptr(), self.fa)
        self.in_dim, self.hidden = l1.in_features, l1.out_features
        self.device = l1.weight.device
        self.fn = _lib.load().gnncca_mpn_forward_frames
        # The kernel raises flags[2] when a frame breaks the trust contract (not closed / not source-major / larger than caps) and
        # leaves that frame's logits as NaN.  The flags live in PINNED HOST memory that the device writes through its mapped
        # address: the normal case costs nothing (the word is never touched), and the host sees a violation of any earlier,
        # completed forward by reading one int at the start of the next call -- no copy, no synchronisation.
        self.flags = torch.zeros(_lib.CSR_NFLAGS, dtype=torch.int32).pin_memory()
        self.flags_np = self.flags.numpy()
        self.training = model.training

    def valid(self, model) -> bool:
        try:
            return self.training == model.training and self.key == tuple(d[n].data_ptr() for d, n in self.slots)
        except (KeyError, AttributeError):
            return False

    def run(self, model, x, edge_index, edge_attr, graph_ptr, edge_ptr, caps):
        """The logits list, or None when this batch is not eligible (the caller then takes the general path)."""
        N, E = x.size(0), edge_index.size(1)
        if not (x.is_cuda and x.dtype == torch.float32 and x.dim() == 2 and x.size(1) == self.in_dim and x.is_contiguous()
                and edge_attr.dtype == torch.float32 and edge_attr.dim() == 2 and edge_attr.size(1) == self.fa and edge_attr.is_contiguous()
                and edge_attr.size(0) == E and edge_index.dtype == torch.int64 and edge_index.is_contiguous() and E > 0
                and graph_ptr.is_cuda and graph_ptr.dtype == torch.int64 and edge_ptr.is_cuda and edge_ptr.dtype == torch.int64
                and x.device == self.device and x.data_ptr() % 16 == 0 and edge_attr.data_ptr() % 16 == 0):
            return None
        B = graph_ptr.numel() - 1
        cap = frame_capacity(model, N, E, B, caps)
        if cap is None or edge_ptr.numel() != B + 1:
            return None
        self.verify()
        h = torch.empty((N, 32), dtype=torch.float32, device=self.device)
        logits = torch.empty((self.n_cls, E), dtype=torch.float32, device=self.device)
        nb, scr = 0, None        # no encoder scratch: the one-launch form (see mlp_forward_nograd)
        w0 = model.encoder.node_mlp.__dict__.get("_gnncca_w0")
        prefix = self.prefix if w0 is None else self.prefix[:2] + (w0.data_ptr(),) + self.prefix[3:]
        rc = self.fn(*prefix, x.data_ptr(), edge_index.data_ptr(), edge_attr.data_ptr(), graph_ptr.data_ptr(), edge_ptr.data_ptr(),
                     N, E, B, self.L, self.first, cap[0], cap[1], h.data_ptr(), logits.data_ptr(), self.flags.data_ptr(),
                     scr.data_ptr() if nb else None, nb, _stream())
        if rc:
            _lib.check(rc, "mpn_forward_frames")
        return [logits[i].view(E, 1) for i in range(self.n_cls)]


def _fast_verify(self, wait: bool = False) -> None:
    """Raise if a forward launched through this binding found a frame that breaks the contract of a trusted batch (edges
    leaving their frame, sources not ascending, a frame larger than ``gnncca_caps``): its logits for that frame are NaN.
    ``wait``: synchronise the device first (otherwise only forwards that have already completed are covered)."""
    if wait:
        torch.cuda.synchronize(self.device)
    if self.flags_np[2]:
        self.flags_np[2] = 0
        raise RuntimeError("gnn-cca_b200: an earlier forward on a batch marked gnncca_trusted met a frame that is not a closed, "
                           "source-major frame within gnncca_caps (edge_index / ptr / edge_ptr edited after build_graph_batch?); the "
                           "logits of that frame were left as NaN.  Clear data.gnncca_trusted to use the checked path.")


FastPath.verify = _fast_verify


def fast_path(model):
    """Cached ``FastPath`` of a model (None when its configuration has no single-call kernel path)."""
    fp = getattr(model, "_gnncca_fast", None)
    if fp is not None and fp is not False and fp.valid(model):
        return fp
    if fp is False and getattr(model, "_gnncca_fast_training", None) == model.training:
        return None
    try:
        fp = FastPath(model)
    except NotImplementedError:
        fp = False
        model.__dict__["_gnncca_fast_training"] = model.training
    model.__dict__["_gnncca_fast"] = fp
    return fp or None
