"""Training path of ``MOTMPNet.forward``: a ``torch.autograd.Function`` whose forward runs the per-step CUDA
pipeline keeping every step's latents, and whose backward runs the transposed kernels of ``csrc/train.cu``
(reference: autograd through models/mpn.py:250-299 under the loss of train.py:80-97, ``loss.backward()`` at
train.py:493).  Inputs ``x`` / ``edge_attr`` get no gradient (the CNN runs under ``no_grad``, train.py:248)."""
from __future__ import annotations

import ctypes as C
import os
from typing import List

import torch

from . import _lib
from .plan import GraphPlan, _stream

P = _lib.ptr
FUSED_BWD = os.environ.get("GNNCCA_FUSED_BWD", "1") != "0"
FUSED_FWD = os.environ.get("GNNCCA_FUSED_TRAIN_FWD", "1") != "0"      # the all-steps forward of small frames in training
FUSED_BWD_MAX_EDGES = int(os.environ.get("GNNCCA_FUSED_BWD_MAX_EDGES", "3500"))     # per frame: above it the per-step chain runs


def _blocks(lib, n, kind):
    return int(lib.gnncca_train_blocks(int(n), kind))


def _bn_mode(model):
    bn = model.classifier.edge_mlp.batchnorm(0)
    if bn is None:
        return None, "none"
    return bn, ("train" if model.training else "eval")


class _Seg:
    """Collects gnncca_fin_seg entries and flushes them in tables of at most MAX_FIN_SEGS."""

    def __init__(self, lib):
        self.lib, self.items = lib, []

    def add(self, partial, n_blocks, width, src_off, src_stride, rows, cols, dst, dst_stride, accumulate=0):
        self.items.append(_lib.FinSeg(P(partial), dst if isinstance(dst, int) else P(dst), n_blocks, width, src_off, src_stride, rows, cols,
                                      dst_stride, accumulate))

    def flush(self):
        for i in range(0, len(self.items), _lib.MAX_FIN_SEGS):
            chunk = self.items[i:i + _lib.MAX_FIN_SEGS]
            arr = (_lib.FinSeg * len(chunk))(*chunk)
            _lib.check(self.lib.gnncca_finalize_grads(arr, len(chunk), _stream()), "finalize_grads")
        self.items = []


class _MPNFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, model, x, edge_index, edge_attr, plan, trusted, frames, *params):
        from . import engine
        lib = _lib.load()
        engine.check_supported(model)
        engine._require_cuda(x, "data.x")
        engine._require_cuda(edge_attr, "data.edge_attr")
        dev = x.device
        N, E = int(x.size(0)), int(edge_index.size(1))
        if plan is None:
            plan = GraphPlan(edge_index, N)
        elif getattr(plan, "_deferred", False):
            plan.check_deferred()          # the flags an earlier forward asked for, if they have arrived (never blocks)
        guard = 0
        if trusted and not plan.resolved:
            # build_graph_batch output: source-major and in range by construction, so nothing is waited for.  The CSR flags still
            # guard every step launch (no-ops on an invalid CSR) and travel back asynchronously for check_deferred().
            plan._resolved = True
            plan._deferred = True
        if getattr(plan, "_deferred", False):
            guard = P(plan.flags)
        plan.resolve()
        enc_n, enc_e = model.encoder.node_mlp, model.encoder.edge_mlp
        if len(enc_e.layer_specs) != 1:
            raise NotImplementedError("training: the edge encoder must be a single Linear+ReLU (edge_fc_dims = [])")
        if int(edge_attr.size(1)) > 8:
            raise NotImplementedError("training: edge_in_dim > 8 has no kernel")
        sb = engine.StepBinding(model, _lib.MATH_FP32)
        L, k = model.num_enc_steps, model.num_class_steps
        first = L - k + 1
        if L < 1:
            raise NotImplementedError("training with num_enc_steps == 0 is not implemented")
        bn, bn_mode = _bn_mode(model)
        if bn is not None and bn.momentum is None:
            raise NotImplementedError("BatchNorm1d(momentum=None) (cumulative average) has no kernel")

        # transposed CSR (edges grouped by target) for the gQ reduction of the backward; kept with the plan, which the
        # module caches on the batch object while edge_index is unchanged
        if getattr(plan, "tperm", None) is None and trusted and plan.perm is None:
            # build_graph_batch output: symmetric, targets ascending inside every source segment -- the target-major permutation
            # is the reverse-edge table (one launch, no sort), its offsets are rowptr; flags[3] reports a broken promise
            tperm = torch.empty(E, dtype=torch.int32, device=dev)
            _lib.check(lib.gnncca_csr_reverse_edges(P(plan.rowptr), P(plan.src), P(plan.dst), N, E, P(tperm), P(plan.flags), _stream()),
                       "csr_reverse_edges")
            plan.tperm, plan.colptr = tperm, plan.rowptr
        if getattr(plan, "tperm", None) is None:
            key, other = plan.dst.clone(), plan.src.clone()
            tperm = torch.empty(E, dtype=torch.int32, device=dev)
            colptr = torch.empty(N + 1, dtype=torch.int32, device=dev)
            nbytes = int(lib.gnncca_csr_sort_workspace_bytes(N, E))
            ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
            _lib.check(lib.gnncca_csr_sort(N, E, P(key), P(other), P(tperm), P(colptr), P(ws), nbytes, _stream()), "csr_sort(dst)")
            plan.tperm, plan.colptr = tperm, colptr
        tperm, colptr = plan.tperm, plan.colptr

        # node encoder: every activation is kept for the backward.  math='tf32' (the module's default): both layers in the
        # tcgen05 kernel of the inference path, which also writes the hidden activations; math='fp32': layer by layer in FP32
        acts = [x.contiguous()]
        if model.math == "tf32" and engine.node_encoder_fused_ok(enc_n, acts[0]):
            l1, l2 = enc_n.linear(0), enc_n.linear(1)
            h1 = torch.empty((N, l1.out_features), dtype=torch.float32, device=dev)
            h2 = torch.empty((N, l2.out_features), dtype=torch.float32, device=dev)
            nb = engine.encoder_scratch_bytes(N, l1.in_features, l1.out_features)       # small batches: the split-K form
            scr = torch.empty(nb // 4, dtype=torch.float32, device=dev) if nb else None
            _lib.check(lib.gnncca_node_encoder_ws(P(acts[0]), P(l1.weight), P(l1.bias), P(l2.weight), P(l2.bias), P(h1), P(h2), N,
                                                  l1.in_features, l1.out_features, l2.out_features, P(scr), nb, _stream()), "node_encoder_ws")
            acts += [h1, h2]
        else:
            for i, sp in enumerate(enc_n.layer_specs):
                acts.append(engine.linear(acts[-1], enc_n.linear(i), sp.relu, None, _lib.MATH_FP32))
        h0 = acts[-1]
        ea = plan.to_sorted(edge_attr.contiguous())

        # eligibility of the one-launch paths (small closed frames with known offsets; see ctx.frames below)
        B = int(frames[0].numel()) - 1 if frames is not None else 0
        frames_ok = (frames is not None and FUSED_BWD and trusted and plan.perm is None and plan.colptr is plan.rowptr and bn_mode != "train"
                     and L <= _lib.FRAMES_BWD_MAX_STEPS and frames[2][1] <= FUSED_BWD_MAX_EDGES and B >= 8
                     and int(lib.gnncca_mp_frames_bwd_smem_bytes(int(frames[2][0]), int(frames[2][1]))) <= _lib.FRAMES_BWD_SMEM_LIMIT)
        cap = engine.frame_capacity(model, N, E, B, frames[2]) if (frames_ok and FUSED_FWD) else None
        if cap is not None and enc_e.layer_specs[0].relu and enc_e.layer_specs[0].bn is None:
            # the all-steps forward of small frames (gnncca_mp_fused_fwd, a frame's latents stay in shared memory across the steps)
            # with every step's latents written out for the backward; the projections are recomputed there
            n_cls = L - first + 1
            logits = torch.empty((n_cls, E), dtype=torch.float32, device=dev)
            h_dump = torch.empty((L + 1, N, sb.f_node), dtype=torch.float32, device=dev)
            e_dump = torch.empty((L + 1, E, sb.f_edge), dtype=torch.float32, device=dev)
            le = enc_e.linear(0)
            _lib.check(lib.gnncca_mp_fused_fwd(C.byref(sb.cfg), C.byref(sb.w), P(le.weight), P(le.bias), int(ea.size(1)), P(frames[0]), B, 0, 0,
                                               P(plan.rowptr), P(plan.src), P(plan.dst), N, E, P(h0), P(ea), L, first, cap[0], cap[1],
                                               P(logits), P(h_dump), P(e_dump), P(plan.flags), _stream()), "mp_fused_fwd")
            hs, es = [h_dump[t] for t in range(L + 1)], [e_dump[t] for t in range(L + 1)]
            affs = {}
            if bn_mode == "eval":
                invstd = torch.rsqrt(bn.running_var.detach() + bn.eps)
                scale = bn.weight.detach() * invstd
                aff = torch.cat([scale, bn.bias.detach() - bn.running_mean * scale, bn.running_mean.detach(), invstd]).contiguous()
                affs = {t: aff for t in range(first, L + 1)}
            if guard and not plan._requested and not torch.cuda.is_current_stream_capturing():
                plan.request_flags()
            ctx.model, ctx.plan, ctx.sb = model, plan, sb
            ctx.frames = (frames[0], frames[1], (int(frames[2][0]), int(frames[2][1])))
            ctx.saved = dict(acts=acts, ea=ea, hs=hs, es=es, affs=affs, proj=None, tperm=tperm, colptr=colptr, bn_mode=bn_mode, first=first,
                             names=[n for n, _ in model.named_parameters_fast()])
            ctx.set_materialize_grads(False)
            model.__dict__["_gnncca_train_plan"] = plan
            return tuple(logits[i].view(E, 1) for i in range(n_cls))
        e0 = engine.small_mlp_rows(enc_e, ea)

        hs, es, affs, outs = [h0], [e0], {}, []
        # the per-node projections of every step are kept for the backward (48 floats per node and step: nothing next to the
        # per-edge latents) instead of being recomputed there
        pw = 16 + sb.f_node
        proj = torch.empty((L, N * pw), dtype=torch.float32, device=dev)
        ws_bytes = int(lib.gnncca_mp_step_workspace_bytes(C.byref(sb.cfg), N, E))
        wsp = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        be = _blocks(lib, E, _lib.BLOCKS_PER_EDGE)
        for step in range(1, L + 1):
            Pn, Qn, Rn = proj[step - 1, :N * 8], proj[step - 1, N * 8:N * 16], proj[step - 1, N * 16:]
            if step == 1:      # later steps: projected by the previous step's fold launch (gnncca_mp_step_fwd_next), bit-equal
                _lib.check(lib.gnncca_node_project(C.byref(sb.cfg), C.byref(sb.w), P(h0), P(hs[-1]), N, P(Pn), P(Qn), P(Rn), _stream()), "node_project")
            e_out = torch.empty_like(e0)
            h_out = torch.empty((N, sb.f_node), dtype=torch.float32, device=dev) if step < L else None
            classify = step >= first
            inline = classify and bn_mode != "train"
            logits = torch.empty((E, 1), dtype=torch.float32, device=dev) if classify else None
            if step < L:
                Px, Qx, Rx = proj[step, :N * 8], proj[step, N * 8:N * 16], proj[step, N * 16:]
                _lib.check(lib.gnncca_mp_step_fwd_next(C.byref(sb.cfg), C.byref(sb.w), P(plan.rowptr), P(plan.src), P(plan.dst), N, E, P(Pn), P(Qn), P(Rn),
                                                       0, 4, 0, 0, P(e0) if model.reattach_initial_edges else 0, P(es[-1]), P(e_out), P(h_out),
                                                       P(logits) if inline else 0, P(h0), P(Px), P(Qx), P(Rx), guard, P(wsp), ws_bytes, _stream()),
                           "mp_step_fwd_next")
            else:
                _lib.check(lib.gnncca_mp_step_fwd(C.byref(sb.cfg), C.byref(sb.w), P(plan.rowptr), P(plan.src), P(plan.dst), N, E, P(Pn), P(Qn), P(Rn),
                                                  P(e0) if model.reattach_initial_edges else 0, P(es[-1]), P(e_out), P(h_out),
                                                  P(logits) if inline else 0, guard, P(wsp), ws_bytes, _stream()), "mp_step_fwd")
            es.append(e_out)
            hs.append(h_out)
            if classify and bn_mode == "train":       # BatchNorm1d with batch statistics over all E edges (mlp.py:16)
                aff = torch.empty(16, dtype=torch.float32, device=dev)
                part = torch.empty((be, 8), dtype=torch.float32, device=dev)
                sums = torch.empty(8, dtype=torch.float32, device=dev)
                for mode in (0, 1):
                    _lib.check(lib.gnncca_cls_moments(C.byref(sb.cfg), C.byref(sb.w), P(e_out), E, 0 if mode == 0 else aff[8:12].data_ptr(),
                                                      P(part), _stream()), "cls_moments")
                    seg = _Seg(lib); seg.add(part, be, 8, 0, 8, 1, 8, sums, 8); seg.flush()
                    _lib.check(lib.gnncca_cls_bn_prepare(mode, P(sums), E, P(bn.weight), P(bn.bias), float(bn.eps), float(bn.momentum),
                                                         P(bn.running_mean), P(bn.running_var), P(aff), _stream()), "cls_bn_prepare")
                bn.num_batches_tracked += 1
                _lib.check(lib.gnncca_cls_fwd(C.byref(sb.cfg), C.byref(sb.w), P(e_out), E, P(aff), P(logits), _stream()), "cls_fwd")
                affs[step] = aff
            elif classify and bn_mode == "eval":
                invstd = torch.rsqrt(bn.running_var.detach() + bn.eps)
                scale = bn.weight.detach() * invstd
                affs[step] = torch.cat([scale, bn.bias.detach() - bn.running_mean * scale, bn.running_mean.detach(), invstd]).contiguous()
            if classify:
                outs.append(plan.to_caller(logits))
        if guard and not plan._requested and not torch.cuda.is_current_stream_capturing():
            plan.request_flags()
        ctx.model, ctx.plan, ctx.sb = model, plan, sb
        # one-launch backward (gnncca_mp_frames_bwd): closed frames with known offsets, every frame small enough for one CTA,
        # the reverse-edge table in place of the sort by target, no batch statistics in the classifier
        ctx.frames = (frames[0], frames[1], (int(frames[2][0]), int(frames[2][1]))) if frames_ok else None
        ctx.saved = dict(acts=acts, ea=ea, hs=hs, es=es, affs=affs, proj=proj, tperm=tperm, colptr=colptr, bn_mode=bn_mode, first=first,
                         names=[n for n, _ in model.named_parameters_fast()])
        ctx.set_materialize_grads(False)
        model.__dict__["_gnncca_train_plan"] = plan
        return tuple(outs)

    @staticmethod
    def backward(ctx, *g_outs):
        from . import engine
        lib = _lib.load()
        model, plan, sb, sv = ctx.model, ctx.plan, ctx.sb, ctx.saved
        hs, es, acts, ea = sv["hs"], sv["es"], sv["acts"], sv["ea"]
        dev = ea.device
        N, E = plan.N, plan.E
        L, first, bn_mode = model.num_enc_steps, sv["first"], sv["bn_mode"]
        cfg, w = C.byref(sb.cfg), C.byref(sb.w)
        FN, FE = sb.f_node, sb.f_edge
        rn, re = bool(model.reattach_initial_nodes), bool(model.reattach_initial_edges)
        K, EIN = (2 * FN if rn else FN), (2 * FE if re else FE)
        edge_row, node_row = 2 * K + EIN, K + FE
        named = dict(model.named_parameters_fast())
        # every buffer that must start at zero comes out of ONE zero-filled allocation (one memset instead of ~30)
        sizes = {n: p.numel() for n, p in named.items()}
        em, nm = "MPNet.edge_model.edge_mlp.fc_layers.0.", "MPNet.node_model.node_mlp.fc_layers.0."
        cls = model.classifier.edge_mlp
        cpre = "classifier.edge_mlp.fc_layers."
        c1, c2 = cpre + "%d." % cls.layer_specs[0].linear, cpre + "%d." % cls.layer_specs[1].linear
        cbn = cpre + "%d." % cls.layer_specs[0].bn if cls.layer_specs[0].bn is not None else None

        be, bnw, bp = _blocks(lib, E, _lib.BLOCKS_PER_EDGE), _blocks(lib, N, _lib.BLOCKS_PER_NODE_WARP), _blocks(lib, N, _lib.BLOCKS_NODE_PROJECT)
        fused = ctx.frames is not None
        if fused:          # one partial row per frame and step
            be = bnw = bp = int(ctx.frames[0].numel()) - 1
        n_cls = L - first + 1
        shapes = [("pa_agg", (L, bnw, FN * FE)), ("pa_edge", (L, be, FE * EIN)), ("pa_proj", (L, bp, 44 * (K + 1))),
                  ("pa_cls", (n_cls, be, _lib.CLS_ACC)), ("gP", (N, 8)), ("gQ", (N, 8)),
                  ("gR", (N, FN)), ("gze", (E, 8)), ("gh_prev", (L, N, FN))]
        if bn_mode == "train":
            shapes.append(("pa_cls1", (n_cls, be, _lib.CLS_ACC)))
        if re:
            shapes.append(("ge0_acc", (E, FE)))
        if rn:
            shapes.append(("gh0_acc", (N, FN)))
        shapes += [("grad/" + n, tuple(p.shape)) for n, p in named.items()]
        offs, total = {}, 0
        for name, shp in shapes:
            cnt = 1
            for d in shp:
                cnt *= d
            offs[name] = (total, cnt, shp)
            total += (cnt + 3) & ~3          # 16-byte aligned pieces
        zeros = torch.zeros(total, dtype=torch.float32, device=dev)
        zb = {name: zeros[o:o + c].view(shp) for name, (o, c, shp) in offs.items()}
        grads = {n: zb["grad/" + n] for n in named}
        pa_agg, pa_edge, pa_proj, pa_cls, pa_cls1 = zb["pa_agg"], zb["pa_edge"], zb["pa_proj"], zb["pa_cls"], zb.get("pa_cls1")
        gP, gQ, gR, gze = zb["gP"], zb["gQ"], zb["gR"], zb["gze"]
        proj = sv["proj"]
        ge0_acc, gh0_acc, gh_prevs = zb.get("ge0_acc"), zb.get("gh0_acc"), zb["gh_prev"]
        gh, ge_next = None, None
        is_max = model.node_agg_name == "max"
        if fused:
            B = be
            tmp = torch.empty(2 * E * FE, dtype=torch.float32, device=dev)
            ge_a, ge_b = tmp[:E * FE], tmp[E * FE:]
            steps = (_lib.FrameBwdStep * L)()
            keep_g = []
            for t in range(1, L + 1):
                g = None
                if t >= first and g_outs[t - first] is not None:
                    g = g_outs[t - first].contiguous().view(E)
                    keep_g.append(g)
                aff = sv["affs"].get(t)
                ci = t - first
                steps[t - 1] = _lib.FrameBwdStep(proj[t - 1, N * 16:].data_ptr() if proj is not None else 0, P(es[t]), P(es[t - 1]), P(hs[t]), P(hs[t - 1]), P(g), P(aff),
                                                 P(pa_agg[t - 1]), P(pa_edge[t - 1]), P(pa_proj[t - 1]), P(pa_cls[ci]) if g is not None else 0,
                                                 P(gh_prevs[t - 1]))
            _lib.check(lib.gnncca_mp_frames_bwd(cfg, w, P(ctx.frames[0]), P(ctx.frames[1]), B, P(plan.rowptr), P(plan.src), P(sv["tperm"]), N, E,
                                                ctx.frames[2][0], ctx.frames[2][1], steps, L, P(es[0]) if re else 0, P(hs[0]) if rn else 0,
                                                P(ge_a), P(ge_b), P(ge0_acc), P(gh0_acc), _stream()), "mp_frames_bwd")
            gh, ge_next = gh_prevs[0], ge_a.view(E, FE)
        for t in (range(L, 0, -1) if not fused else ()):
            ge_cls = None
            if t >= first and g_outs[t - first] is not None:
                ci = t - first
                g = plan.to_sorted(g_outs[ci].contiguous().view(E, 1)).contiguous()
                aff = sv["affs"].get(t)
                ge_cls = torch.empty((E, FE), dtype=torch.float32, device=dev)
                if bn_mode == "train":
                    _lib.check(lib.gnncca_cls_bwd(cfg, w, P(es[t]), E, P(aff), 1, P(g), 0, 0, P(pa_cls1[ci]), _stream()), "cls_bwd pass 1")
                    sums = torch.empty(_lib.CLS_ACC, dtype=torch.float32, device=dev)
                    seg = _Seg(lib); seg.add(pa_cls1[ci], be, _lib.CLS_ACC, 0, _lib.CLS_ACC, 1, _lib.CLS_ACC, sums, _lib.CLS_ACC); seg.flush()
                    _lib.check(lib.gnncca_cls_bwd(cfg, w, P(es[t]), E, P(aff), 2, P(g), P(sums), P(ge_cls), P(pa_cls[ci]), _stream()), "cls_bwd pass 2")
                else:
                    _lib.check(lib.gnncca_cls_bwd(cfg, w, P(es[t]), E, P(aff), 0, P(g), 0, P(ge_cls), P(pa_cls[ci]), _stream()), "cls_bwd")
            Rn = proj[t - 1, N * 16:]            # R of step t, kept by the forward
            if gh is not None:
                gh_eff = torch.empty((2, N, FN), dtype=torch.float32, device=dev) if is_max else None   # gh / ties | recomputed max
                _lib.check(lib.gnncca_node_agg_bwd(cfg, w, P(plan.rowptr), N, P(Rn), P(es[t]), P(gh), P(hs[t]), P(gR), P(gh_eff), P(pa_agg[t - 1]),
                                                   _stream()), "node_agg_bwd")
                if is_max:
                    gh = gh_eff[0]      # ties at the max share the gradient
            else:
                gR.zero_()
            ge_prev = torch.empty((E, FE), dtype=torch.float32, device=dev)
            _lib.check(lib.gnncca_edge_bwd(cfg, w, P(plan.rowptr), P(plan.src), E, P(Rn), P(es[t]), P(es[t - 1]), P(es[0]) if re else 0, P(gh),
                                           (P(gh_eff[1]) if is_max else P(hs[t])) if gh is not None else 0, P(ge_next), P(ge_cls), P(gze), P(ge_prev), P(ge0_acc),
                                           P(pa_edge[t - 1]), _stream()), "edge_bwd")
            _lib.check(lib.gnncca_pq_reduce(P(plan.rowptr), P(sv["colptr"]), P(sv["tperm"]), P(gze), N, P(gP), P(gQ), _stream()), "pq_reduce")
            gh_prev = gh_prevs[t - 1]
            _lib.check(lib.gnncca_node_project_bwd(cfg, w, P(gP), P(gQ), P(gR), P(hs[0]), P(hs[t - 1]), N, P(gh_prev), P(gh0_acc), P(pa_proj[t - 1]),
                                                   _stream()), "node_project_bwd")
            gh, ge_next = gh_prev, ge_prev
        ge0 = ge_next + ge0_acc if re else ge_next
        gh0 = gh + gh0_acc if rn else gh

        seg = _Seg(lib)
        gWe, gWn = grads[em + "weight"], grads[nm + "weight"]
        seg.add(pa_agg, L * bnw, FN * FE, 0, FE, FN, FE, gWn.data_ptr() + 4 * K, node_row)
        seg.add(pa_edge, L * be, FE * EIN, 0, EIN, FE, EIN, gWe.data_ptr() + 4 * 2 * K, edge_row)
        wp = 44 * (K + 1)
        seg.add(pa_proj, L * bp, wp, 0, K + 1, FE, K, gWe, edge_row)
        seg.add(pa_proj, L * bp, wp, FE * (K + 1), K + 1, FE, K, gWe.data_ptr() + 4 * K, edge_row)
        seg.add(pa_proj, L * bp, wp, 2 * FE * (K + 1), K + 1, FN, K, gWn, node_row)
        seg.add(pa_proj, L * bp, wp, K, K + 1, FE, 1, grads[em + "bias"], 1)
        seg.add(pa_proj, L * bp, wp, 2 * FE * (K + 1) + K, K + 1, FN, 1, grads[nm + "bias"], 1)
        A = _lib.CLS_ACC
        HC = sb.cfg.cls_hidden
        seg.add(pa_cls, n_cls * be, A, 0, FE, HC, FE, grads[c1 + "weight"], FE)
        seg.add(pa_cls, n_cls * be, A, HC * FE, HC, 1, HC, grads[c1 + "bias"], HC)
        last = pa_cls1 if bn_mode == "train" else pa_cls
        seg.add(last, n_cls * be, A, HC * FE + HC, HC, 1, HC, grads[c2 + "weight"], HC)
        seg.add(last, n_cls * be, A, HC * FE + 2 * HC, 1, 1, 1, grads[c2 + "bias"], 1)
        if cbn is not None:
            seg.add(last, n_cls * be, A, HC * FE + 2 * HC + 1, HC, 1, HC, grads[cbn + "weight"], HC)
            seg.add(last, n_cls * be, A, HC * FE + 3 * HC + 1, HC, 1, HC, grads[cbn + "bias"], HC)
        # edge encoder
        enc_e = model.encoder.edge_mlp
        fa = int(ea.size(1))
        be = _blocks(lib, E, _lib.BLOCKS_PER_EDGE)       # (the fused backward counted frames)
        pa_enc = torch.empty((be, FE * 8 + FE), dtype=torch.float32, device=dev)
        _lib.check(lib.gnncca_edge_enc_bwd(P(ea), fa, P(es[0]), P(ge0), E, P(pa_enc), _stream()), "edge_enc_bwd")
        epre = "encoder.edge_mlp.fc_layers.%d." % enc_e.layer_specs[0].linear
        seg.add(pa_enc, be, FE * 8 + FE, 0, 8, FE, fa, grads[epre + "weight"], fa)
        seg.add(pa_enc, be, FE * 8 + FE, FE * 8, FE, 1, FE, grads[epre + "bias"], FE)
        # node encoder, last layer first
        enc_n = model.encoder.node_mlp
        keep = [pa_agg, pa_edge, pa_proj, pa_cls, pa_cls1, pa_enc]
        gy = gh0.contiguous()
        rows = N
        chunks = _blocks(lib, rows, _lib.BLOCKS_ROW_CHUNKS)
        for i in range(len(enc_n.layer_specs) - 1, -1, -1):
            sp, lin = enc_n.layer_specs[i], enc_n.linear(i)
            # row splits of the weight-gradient GEMM: all its CTAs in ONE wave (3 per SM) with as few iterations each as that allows
            tiles = ((sp.in_dim + 63) // 64) * ((sp.out_dim + 63) // 64)
            splits = max(1, min(rows // 64, 444 // tiles))
            gz = torch.empty((rows, sp.out_dim), dtype=torch.float32, device=dev)
            gx = torch.empty((rows, sp.in_dim), dtype=torch.float32, device=dev) if i > 0 else None
            gbp = torch.empty((chunks, sp.out_dim), dtype=torch.float32, device=dev)
            gwp = torch.empty((splits, sp.out_dim * sp.in_dim), dtype=torch.float32, device=dev)
            _lib.check(lib.gnncca_linear_bwd(P(acts[i]), P(lin.weight), P(acts[i + 1]), P(gy), int(sp.relu), rows, sp.in_dim, sp.out_dim, P(gz), P(gx),
                                             P(gbp), P(gwp), splits, _lib.MATH_TF32 if model.math == "tf32" else _lib.MATH_FP32, _stream()),
                           "linear_bwd")
            npre = "encoder.node_mlp.fc_layers.%d." % sp.linear
            seg.add(gwp, splits, sp.out_dim * sp.in_dim, 0, sp.in_dim, sp.out_dim, sp.in_dim, grads[npre + "weight"], sp.in_dim)
            seg.add(gbp, chunks, sp.out_dim, 0, sp.out_dim, 1, sp.out_dim, grads[npre + "bias"], sp.out_dim)
            keep += [gz, gbp, gwp]
            gy = gx
        seg.flush()
        out: List = [None, None, None, None, None, None, None]
        for n in sv["names"]:
            out.append(grads[n] if named[n].requires_grad else None)
        return tuple(out)


def forward_training(model, x, edge_index, edge_attr, plan=None, trusted=False, return_plan=False, frames=None):
    """``MOTMPNet.forward`` with an autograd graph: returns the list of per-step logits [E,1] (and, on request, the
    CSR plan it used, which carries the transposed CSR of the backward: the module caches it on the batch object)."""
    params = [p for _, p in model.named_parameters_fast()]
    outs = list(_MPNFunction.apply(model, x, edge_index, edge_attr, plan, trusted, frames, *params))
    if return_plan:
        return outs, model.__dict__.pop("_gnncca_train_plan", None)
    return outs
