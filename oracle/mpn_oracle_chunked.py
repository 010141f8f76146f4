"""The oracle's ``motmpnet_forward`` evaluated in chunks of SOURCE rows (TEST INFRASTRUCTURE ONLY).

``oracle/mpn_oracle.py`` materialises ``cat[h[row], h[col], e]`` for all E edges at once, like the reference
(``models/mpn.py:68,97``); for BASELINE config 5 (62.9 M edges) that is 17.6 GB in fp32.  This file runs the SAME
functions (``mpn_oracle.mlp_forward`` over the same ``state_dict`` keys, ``scatter_*`` with the same edge order) over
contiguous chunks of a source-major edge list, so that nothing larger than one chunk's ``[E_c, 70]`` exists, in the dtype
of its inputs (float64 for the giant-graph parity check of ``bench.py`` / ``tests``).  Every node's out-edges lie in one
chunk, so each segment is reduced in the same order as in the unchunked oracle: in the same dtype the two agree bit for
bit (``tests/test_oracle_golden.py::test_chunked_oracle_equals_oracle``).
"""
from __future__ import annotations

from typing import Dict, List

import torch

from . import mpn_oracle as mo

Tensor = torch.Tensor


def cast_state_dict(sd: Dict[str, Tensor], dtype) -> Dict[str, Tensor]:
    return {k: (v.to(dtype) if v.is_floating_point() else v) for k, v in sd.items()}


def source_chunks(row: Tensor, n_nodes: int, max_edges: int) -> List[tuple]:
    """[(e0, e1)] covering a non-decreasing ``row``: whole source segments, about ``max_edges`` edges each."""
    E = int(row.numel())
    if E == 0:
        return []
    if bool((row[1:] < row[:-1]).any()):
        raise ValueError("chunked oracle needs source-major edges (edge_index[0] non-decreasing)")
    rowptr = torch.searchsorted(row.contiguous(), torch.arange(n_nodes + 1, dtype=row.dtype))
    out, e0 = [], 0
    while e0 < E:
        goal = min(E, e0 + max_edges)
        v = int(torch.searchsorted(rowptr, torch.tensor(goal, dtype=rowptr.dtype), right=False))   # first node boundary >= goal
        e1 = int(rowptr[min(v, n_nodes)])
        if e1 <= e0:
            e1 = int(rowptr[min(v + 1, n_nodes)])
        out.append((e0, e1))
        e0 = e1
    return out


def motmpnet_forward_chunked(spec: mo.MPNSpec, sd: Dict[str, Tensor], x: Tensor, edge_index: Tensor, edge_attr: Tensor,
                             max_edges: int = 1 << 21, out_dtype=None) -> Dict[str, List[Tensor]]:
    """``MOTMPNet.forward`` (mpn.py:250-299), eval mode, source-major ``edge_index``; see the module docstring."""
    row, col = edge_index[0], edge_index[1]
    N, E = int(x.size(0)), int(row.numel())
    dt = x.dtype
    chunks = source_chunks(row, N, max_edges)
    h = mo.mlp_forward(sd, "encoder.node_mlp", spec.enc_node, x) if spec.enc_node is not None else x
    h0 = h
    e_all = None
    for a, b in chunks:     # encoder, mpn.py:128-142
        ec = mo.mlp_forward(sd, "encoder.edge_mlp", spec.enc_edge, edge_attr[a:b]) if spec.enc_edge is not None else edge_attr[a:b]
        if e_all is None:
            e_all = torch.empty((E, ec.size(1)), dtype=dt)
        e_all[a:b] = ec
    if e_all is None:
        e_all = torch.empty((0, int(edge_attr.size(1))), dtype=dt)
    e0_all = e_all.clone() if spec.reattach_edges else None
    first = spec.num_enc_steps - spec.num_class_steps + 1      # mpn.py:277
    od = out_dtype or dt
    outs: List[Tensor] = []

    def classify(ee):
        return mo.mlp_forward(sd, "classifier.edge_mlp", spec.cls_edge, ee) if spec.cls_edge is not None else ee

    for step in range(1, spec.num_enc_steps + 1):
        hh = torch.cat((h0, h), dim=1) if spec.reattach_nodes else h       # mpn.py:285
        h_new = torch.zeros((N, h.size(1)), dtype=dt)
        logits = torch.empty((E, 1), dtype=od) if step >= first else None
        for a, b in chunks:
            r, c = row[a:b], col[a:b]
            e = e_all[a:b]
            if spec.reattach_edges:
                e = torch.cat((e0_all[a:b], e), dim=1)                     # mpn.py:283
            e = mo.mlp_forward(sd, "MPNet.edge_model.edge_mlp", spec.edge_model, torch.cat([hh[r], hh[c], e], dim=1))   # mpn.py:68
            m = mo.mlp_forward(sd, "MPNet.node_model.node_mlp", spec.node_model, torch.cat([hh[r], e], dim=1))          # mpn.py:97
            v0, v1 = int(r[0]), int(r[-1]) + 1
            loc = r - v0
            if spec.agg == "mean":
                h_new[v0:v1] = mo.scatter_mean(m, loc, 0, v1 - v0)
            elif spec.agg == "max":
                h_new[v0:v1] = mo.scatter_max(m, loc, 0, v1 - v0)[0]
            else:
                h_new[v0:v1] = mo.scatter_add(m, loc, 0, v1 - v0)
            e_all[a:b] = e
            if logits is not None:
                logits[a:b] = classify(e).to(od)
        h = h_new
        if logits is not None:
            outs.append(logits)
    if spec.num_enc_steps == 0:
        outs.append(torch.cat([classify(e_all[a:b]).to(od) for a, b in chunks]) if chunks else torch.empty((0, 1), dtype=od))
    return {"classified_edges": outs}


def motmpnet_step(spec: mo.MPNSpec, sd: Dict[str, Tensor], h: Tensor, h0: Tensor, e: Tensor, e0: Tensor, row: Tensor, col: Tensor,
                  rows_out: Tensor, classify: bool = True):
    """ONE iteration of the loop of ``MOTMPNet.forward`` (mpn.py:279-293) for a subset of the edges: the reattach
    concatenations (:282-285), ``MetaLayer`` (:44-54: edge model on the OLD nodes, node model on the NEW edges, aggregation by
    ``row``) and the classifier (:290-293), in the dtype of the inputs.

    ``h`` / ``h0``: current / initial node latents of ALL nodes; ``e`` / ``e0``: current / initial latents of the given edges
    only; ``row`` / ``col``: their GLOBAL endpoints; ``rows_out``: the source nodes whose out-edges are ALL among the given
    edges (their new latents are exact).  Returns (e', h'[rows_out], logits or None).  Used for the per-step parity check of
    graphs too large for an end-to-end CPU evaluation: the inputs of step t are the DEVICE's outputs of step t-1."""
    ee = torch.cat((e0, e), dim=1) if spec.reattach_edges else e
    hh = torch.cat((h0, h), dim=1) if spec.reattach_nodes else h
    e_new = mo.mlp_forward(sd, "MPNet.edge_model.edge_mlp", spec.edge_model, torch.cat([hh[row], hh[col], ee], dim=1))
    m = mo.mlp_forward(sd, "MPNet.node_model.node_mlp", spec.node_model, torch.cat([hh[row], e_new], dim=1))
    n = int(h.size(0))
    if spec.agg == "mean":
        h_new = mo.scatter_mean(m, row, 0, n)
    elif spec.agg == "max":
        h_new = mo.scatter_max(m, row, 0, n)[0]
    else:
        h_new = mo.scatter_add(m, row, 0, n)
    logits = None
    if classify:
        logits = mo.mlp_forward(sd, "classifier.edge_mlp", spec.cls_edge, e_new) if spec.cls_edge is not None else e_new
    return e_new, h_new[rows_out], logits
