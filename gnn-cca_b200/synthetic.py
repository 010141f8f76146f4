"""Synthetic EPFL-shaped detections: C cameras x D identities per frame, every identity
seen by every camera, camera-major rows; embeddings non-negative like pooled post-ReLU CNN features."""
from __future__ import annotations

import torch


def make_detections(n_cams: int, n_dets: int, n_graphs: int, seed: int, device, node_dim: int = 2048, reid_dim: int = 256,
                    max_dist: float = 50.83):
    n = n_cams * n_dets
    N = n * n_graphs
    g = torch.Generator(device=device).manual_seed(seed)
    k = torch.arange(n, device=device)
    id_cam = (k // n_dets).repeat(n_graphs).to(torch.int32)
    ids = (k % n_dets).repeat(n_graphs).to(torch.int64)
    centers = torch.rand((n_graphs, n_dets, 2), generator=g, device=device, dtype=torch.float64) * torch.tensor([360.0, 288.0], device=device, dtype=torch.float64)
    pts = centers[:, None, :, :].expand(n_graphs, n_cams, n_dets, 2) + 5.0 * torch.randn((n_graphs, n_cams, n_dets, 2), generator=g, device=device, dtype=torch.float64)
    pts = pts.reshape(N, 2)
    node = torch.relu(torch.randn((N, node_dim), generator=g, device=device))
    reid = torch.relu(torch.randn((N, reid_dim), generator=g, device=device))
    graph_ptr = torch.arange(0, N + 1, n, device=device, dtype=torch.int64)
    md = torch.full((n_graphs,), max_dist, dtype=torch.float64, device=device)
    return dict(graph_ptr=graph_ptr, id_cam=id_cam, ids=ids, xw=pts[:, 0].contiguous(), yw=pts[:, 1].contiguous(),
                max_dist=md, reid_embeds=reid, node_embeds=node)


def edges_per_graph(n_cams: int, n_dets: int) -> int:
    n = n_cams * n_dets
    return n * n - n_cams * n_dets * n_dets
