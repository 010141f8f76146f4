"""CPU restatement of the reference's per-frame graph + edge-attribute construction
(TEST INFRASTRUCTURE ONLY).

The reference has no function for this: the code is written inline, three times, inside its epoch
loops (``train.py:257-259,286-361,450``; ``train.py:582-584,616-692,790``; ``inference.py:189-190,
199-279``) and cannot be imported as a unit.  This file restates it line by line (citations below
use the ``train.py`` copy).  It is pinned by ``tests/golden/graph_*.npz``: those vectors were produced
by driving the reference's *own* ``inference.validate_GNN_cross_camera_association`` loop on CPU
(``oracle/gen_golden.py``), so the restatement is checked against the real code, integers bit-exact.

Third-party pieces restated from their published semantics (absent from this image):
``torch_geometric`` 2.0.1 ``Data`` / ``Batch.from_data_list`` (``env_gnn.yml:87``; call sites
``train.py:357,450``) and ``sklearn.metrics.pairwise.paired_distances`` 0.24.2 (``train.py:322,325``).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import numpy as np
import torch
import torch.nn.functional as F


# --------------------------------------------------------------------------------------
# torch_geometric.data.Data / Batch (2.0.1), the subset the reference touches
# --------------------------------------------------------------------------------------
class Data:
    """Attribute bag; ``num_nodes`` = ``x.size(0)`` (PyG infers it from ``x``)."""

    def __init__(self, x=None, edge_index=None, edge_attr=None, y=None, **kwargs):
        self.x, self.edge_index, self.edge_attr, self.y = x, edge_index, edge_attr, y
        self._extra = list(kwargs)
        for k, v in kwargs.items():
            setattr(self, k, v)

    @property
    def keys(self):
        return [k for k in ["x", "edge_index", "edge_attr", "y"] + self._extra if getattr(self, k) is not None]

    @property
    def num_nodes(self):
        return int(self.x.size(0))


class Batch(Data):
    """Disjoint union: every attribute is concatenated on dim 0, except ``edge_index`` which is
    concatenated on dim 1 after adding the running node count; adds ``batch`` [N], ``ptr`` [B+1]."""

    @classmethod
    def from_data_list(cls, data_list: Sequence[Data]) -> "Batch":
        keys = data_list[0].keys
        out = cls()
        out._extra = [k for k in keys if k not in ("x", "edge_index", "edge_attr", "y")]
        cum, ptr, batch_vec = 0, [0], []
        cols: Dict[str, list] = {k: [] for k in keys}
        for g, d in enumerate(data_list):
            n = d.num_nodes
            for k in keys:
                v = getattr(d, k)
                cols[k].append(v + cum if k == "edge_index" else v)
            batch_vec.append(torch.full((n,), g, dtype=torch.long))
            cum += n
            ptr.append(cum)
        for k in keys:
            setattr(out, k, torch.cat(cols[k], dim=1 if k == "edge_index" else 0))
        out.batch = torch.cat(batch_vec)
        out.ptr = torch.tensor(ptr, dtype=torch.long)
        out.num_graphs = len(data_list)
        return out


# --------------------------------------------------------------------------------------
# sklearn paired_distances (0.24.2): euclidean = row_norms(X - Y), manhattan = |X - Y|.sum(-1)
# --------------------------------------------------------------------------------------
def paired_distances(x: np.ndarray, y: np.ndarray, metric: str = "euclidean") -> np.ndarray:
    d = np.asarray(x, dtype=np.float64) - np.asarray(y, dtype=np.float64)
    if metric == "euclidean":
        return np.sqrt(np.einsum("ij,ij->i", d, d))
    if metric == "manhattan":
        return np.abs(d).sum(axis=-1)
    raise ValueError(metric)


# --------------------------------------------------------------------------------------
# train.py:257-259 -- batch-wide COLUMN normalisation (dim=0), as the reference does it
# --------------------------------------------------------------------------------------
def column_normalize(emb: torch.Tensor) -> torch.Tensor:
    return F.normalize(emb, p=2, dim=0)


# --------------------------------------------------------------------------------------
# train.py:292-299 -- edge enumeration for one frame
# --------------------------------------------------------------------------------------
def enumerate_edges(id_cam: np.ndarray, node: np.ndarray) -> torch.Tensor:
    """For each camera id ascending: cartesian_prod(nodes in cam, nodes not in cam); cat; transpose.
    Returns int64 [2, e] in GLOBAL node ids (rebased later, train.py:356)."""
    id_cam = np.asarray(id_cam)
    node = np.asarray(node)
    blocks = []
    for cam in np.unique(id_cam):
        ids_in = torch.from_numpy(node[id_cam == cam])
        ids_out = torch.from_numpy(node[id_cam != cam])
        # cartesian_prod of two 1-D tensors -> [len(in)*len(out), 2], first factor slowest
        a = ids_in.repeat_interleave(ids_out.numel())
        b = ids_out.repeat(ids_in.numel())
        blocks.append(torch.stack([a, b], dim=1))
    return torch.cat(blocks, dim=0).T.contiguous()


def build_frame(frame: dict, first_node: int, reid_embeds: torch.Tensor, node_embeds: torch.Tensor,
                max_dist: float, mode: str = "full") -> Data:
    """train.py:286-357 for one frame ``g``.  ``frame`` holds 1-D arrays id, id_cam, xw, yw in
    DataFrame row order.  ``mode``: 'full' | 'only_appearance' | 'only_dist' (train.py:335-344)."""
    n = len(frame["id"])
    node = np.arange(first_node, first_node + n)                                   # :288
    x_g = torch.stack([node_embeds[i] for i in node])                              # :291
    ei = enumerate_edges(frame["id_cam"], node)                                    # :292-299
    ei_np = ei.numpy()
    emb_dist = F.pairwise_distance(reid_embeds[ei[0]], reid_embeds[ei[1]]).view(-1, 1)    # :306
    emb_cos = F.cosine_similarity(reid_embeds[ei[0]], reid_embeds[ei[1]]).view(-1, 1)     # :308
    xw = np.asarray(frame["xw"], dtype=np.float64)
    yw = np.asarray(frame["yw"], dtype=np.float64)
    p1 = np.stack([xw[ei_np[0] - first_node], yw[ei_np[0] - first_node]], axis=1)  # :313-319
    p2 = np.stack([xw[ei_np[1] - first_node], yw[ei_np[1] - first_node]], axis=1)
    sp_l2 = torch.from_numpy(paired_distances(p1, p2)[:, None] / max_dist)         # :322-323 (float64)
    sp_l1 = torch.from_numpy(paired_distances(p1, p2, "manhattan")[:, None] / max_dist)  # :325-326
    if mode == "only_appearance":
        edge_attr = torch.cat((emb_dist, emb_cos), dim=1)                          # :337
    elif mode == "only_dist":
        edge_attr = torch.cat((sp_l2.type(torch.float32), sp_l1.type(torch.float32)), dim=1)  # :340
    else:
        edge_attr = torch.cat((sp_l2.type(torch.float32), sp_l1.type(torch.float32), emb_dist, emb_cos), dim=1)  # :344
    ids = np.asarray(frame["id"])
    labels = torch.from_numpy((ids[ei_np[0] - first_node] == ids[ei_np[1] - first_node]).astype(np.int64)).type(torch.float)  # :348-351
    ei = ei - torch.min(ei)                                                        # :356
    return Data(x=x_g, edge_index=ei, y=torch.from_numpy(ids.astype(np.int64)), edge_attr=edge_attr, edge_labels=labels)  # :357


def build_batch(frames: List[dict], reid_embeds: torch.Tensor, node_embeds: torch.Tensor,
                max_dist: Sequence[float], mode: str = "full", l2norm: bool = True) -> Batch:
    """train.py:257-259 + the per-frame loop :271-361 + Batch.from_data_list :450."""
    if l2norm:
        reid_embeds = column_normalize(reid_embeds)
        node_embeds = column_normalize(node_embeds)
    first, out = 0, []
    for g, fr in enumerate(frames):
        out.append(build_frame(fr, first, reid_embeds, node_embeds, float(max_dist[g]), mode))
        first += len(fr["id"])
    return Batch.from_data_list(out)


# --------------------------------------------------------------------------------------
# GT parsing rules for fixtures (libs/datasets.py:89-134, libs/utils.py:211-225)
# --------------------------------------------------------------------------------------
def homography_image_to_world(xi: np.ndarray, yi: np.ndarray, H: np.ndarray):
    """utils.py:211-225 applied elementwise: prj = H @ [xi, yi, 1]; (prj0/prj2, prj1/prj2)."""
    xi = np.asarray(xi, dtype=np.float64)
    yi = np.asarray(yi, dtype=np.float64)
    p0 = H[0, 0] * xi + H[0, 1] * yi + H[0, 2]
    p1 = H[1, 0] * xi + H[1, 1] * yi + H[1, 2]
    p2 = H[2, 0] * xi + H[2, 1] * yi + H[2, 2]
    return p0 / p2, p1 / p2
