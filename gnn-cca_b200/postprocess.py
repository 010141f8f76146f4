"""Edge decisions, pruning of one-directional edges and identity clusters on the device.

Replaces the networkx / list-comprehension block of the reference's evaluation loop, inference.py:286-312,331-335:

    predictions = (sigmoid(preds) >= 0.5) * 1                                              inference.py:289-291
    predictions, active = utils.remove_edges_single_direction(active, predictions, edges)  utils.py:387-404
    ID_pred, n_clusters = utils.compute_SCC_and_Clusters(nx.DiGraph(active), num_nodes)    utils.py:295-317

with one call (two launches, ``gnncca_frame_clusters``).  The rounding and splitting heuristics (utils.py:25-173,
319-386) stay host code; ``frame_clusters(kind="flags")`` re-clusters their output.  Cluster ids are dense in order
of each cluster's smallest node: the SAME PARTITION as the reference's (whose numbering follows networkx's traversal
order and only feeds permutation-invariant metrics, inference.py:347-352).  No CPU fallback.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib

KINDS = {"logits": 0, "labels": 1, "flags": 2}
MAX_FRAME_NODES = 64


class Clusters:
    """``predictions`` [E] int64 (after pruning), ``ids`` [N] int64, ``n_clusters`` 0-dim int64 tensor — all on the device."""

    def __init__(self, predictions, ids, n_clusters):
        self.predictions, self.ids, self.n_clusters = predictions, ids, n_clusters

    def active_edges(self, edge_index: torch.Tensor):
        """The reference's ``predicted_active_edges`` list of (src, dst) tuples (inference.py:300), built on request."""
        keep = self.predictions.bool()
        return [tuple(p) for p in edge_index[:, keep].t().tolist()]


def edge_ptr_of(edge_index: torch.Tensor, ptr: torch.Tensor) -> torch.Tensor:
    """Per-frame edge offsets of a batch whose edges are stored frame after frame (PyG ``Batch.from_data_list``,
    train.py:450).  Frame membership of every edge is verified by the kernel."""
    frame = torch.bucketize(edge_index[0], ptr[1:], right=True)
    return torch.searchsorted(frame, torch.arange(ptr.numel(), device=ptr.device, dtype=frame.dtype))


def frame_clusters(decision: torch.Tensor, edge_index: torch.Tensor, ptr: torch.Tensor, edge_ptr: Optional[torch.Tensor] = None,
                   kind: str = "logits", prune: bool = False, want_predictions: bool = True) -> Clusters:
    lib = _lib.load()
    if kind not in KINDS:
        raise ValueError("kind must be one of %s" % sorted(KINDS))
    if not edge_index.is_cuda:
        raise RuntimeError("gnncca post-processing kernels need CUDA tensors (there is no CPU path)")
    dev = edge_index.device
    if edge_index.dtype != torch.int64 or edge_index.dim() != 2 or edge_index.size(0) != 2:
        raise ValueError("edge_index must be int64 [2, E]")
    edge_index = edge_index.contiguous()
    E = edge_index.size(1)
    want = torch.int64 if kind == "flags" else torch.float32
    decision = decision.detach().reshape(-1)
    if decision.dtype != want:
        decision = decision.to(want)
    decision = decision.contiguous()
    if decision.numel() != E or decision.device != dev:
        raise ValueError("one decision value per edge, on the device of edge_index")
    ptr = ptr.to(device=dev, dtype=torch.int64).contiguous()
    G = ptr.numel() - 1
    if G < 0:
        raise ValueError("ptr must hold n_graphs + 1 node offsets")
    edge_ptr = edge_ptr_of(edge_index, ptr) if edge_ptr is None else edge_ptr.to(device=dev, dtype=torch.int64).contiguous()
    if edge_ptr.numel() != G + 1:
        raise ValueError("edge_ptr must hold n_graphs + 1 edge offsets")
    # one host read for the sizes the caller's tensors do not carry (N and a consistency check of the offsets)
    N, e_end = (int(v) for v in torch.stack([ptr[-1], edge_ptr[-1]]).tolist()) if G > 0 else (0, 0)
    if e_end != E:
        raise ValueError("edge_ptr[-1] != number of edges: edges are not stored frame after frame")
    pred = torch.empty(E, dtype=torch.int64, device=dev) if want_predictions else None
    ids = torch.empty(N, dtype=torch.int64, device=dev)
    tail = torch.zeros(4, dtype=torch.int64, device=dev)            # [0] n_clusters, [1] flags (2 x int32)
    scratch = torch.empty(N + G + 1, dtype=torch.int32, device=dev)
    flags = tail[1:2].view(torch.int32)
    stream = torch.cuda.current_stream(dev).cuda_stream
    _lib.check(lib.gnncca_frame_clusters(decision.data_ptr(), KINDS[kind], edge_index.data_ptr(), E, ptr.data_ptr(), edge_ptr.data_ptr(), G, N,
                                         int(bool(prune)), _lib.ptr(pred), ids.data_ptr(), tail.data_ptr(), scratch.data_ptr(),
                                         flags.data_ptr(), stream), "gnncca_frame_clusters")
    res = Clusters(pred, ids, tail[0])
    res._flags = flags
    return res


def check(res: Clusters) -> Clusters:
    """Raise if the kernel refused a frame (synchronises: reads 8 bytes)."""
    f = res._flags.tolist()
    if f[1]:
        raise NotImplementedError("a frame has more than %d nodes: no clustering kernel for it" % MAX_FRAME_NODES)
    if f[0]:
        raise IndexError("an edge leaves its frame (edge_index / ptr / edge_ptr are inconsistent)")
    return res


def decide_and_cluster(outputs, data_batch, prune: bool = True):
    """inference.py:286-312 in one call: ``(predictions, ID_pred, n_clusters_pred)`` of the LAST classified step, with
    ``ID_pred`` a numpy int array [num_nodes] and ``n_clusters_pred`` an int as the reference's metrics code expects
    (inference.py:347-352); ``predictions`` stays a device tensor (input of the rounding / splitting heuristics)."""
    logits = outputs["classified_edges"][-1]
    res = check(frame_clusters(logits, data_batch.edge_index, data_batch.ptr, getattr(data_batch, "edge_ptr", None), "logits", prune))
    return res.predictions, res.ids.cpu().numpy(), int(res.n_clusters)


def ground_truth_clusters(data_batch):
    """inference.py:297-299: clusters of the edges with label 1 -> ``(ID_GT, n_clusters_GT)``."""
    res = check(frame_clusters(data_batch.edge_labels, data_batch.edge_index, data_batch.ptr, getattr(data_batch, "edge_ptr", None), "labels",
                               False, want_predictions=False))
    return res.ids.cpu().numpy(), int(res.n_clusters)
