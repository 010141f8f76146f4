"""GPU parity of edge decisions + pruning + clustering (gnncca_frame_clusters) against vectors made by
running the reference's own utils functions with networkx (tests/golden/post_*.npz) and against the CPU oracle."""
import types

import numpy as np
import pytest
import torch

from conftest import golden_cases, load_golden
from oracle import post_oracle

pytestmark = pytest.mark.gpu


def dev_batch(z, dev, with_edge_ptr=True):
    b = types.SimpleNamespace(edge_index=torch.from_numpy(z["edge_index"]).to(dev), ptr=torch.from_numpy(z["ptr"]).to(dev),
                              edge_labels=torch.from_numpy(z["labels"]).to(dev))
    if with_edge_ptr:
        b.edge_ptr = torch.from_numpy(z["edge_ptr"]).to(dev)
    return b


@pytest.mark.parametrize("with_edge_ptr", [True, False])
@pytest.mark.parametrize("case", golden_cases("post_"))
def test_decisions_pruning_and_clusters_match_the_reference_functions(case, with_edge_ptr, cuda_device):
    import gnncca
    from gnncca import postprocess as pp
    z = load_golden("post_" + case)
    b = dev_batch(z, cuda_device, with_edge_ptr)
    logits = torch.from_numpy(z["logits"]).to(cuda_device).view(-1, 1)
    ep = getattr(b, "edge_ptr", None)
    # raw decisions (inference.py:289-302)
    n0 = gnncca._lib.load().gnncca_launch_count()
    r = pp.check(pp.frame_clusters(logits, b.edge_index, b.ptr, ep, "logits", prune=False))
    assert gnncca._lib.load().gnncca_launch_count() - n0 == 2
    assert r.predictions.dtype == torch.int64 and np.array_equal(r.predictions.cpu().numpy(), z["predictions"])     # bit-exact
    assert int(r.n_clusters) == int(z["n_raw"])
    assert np.array_equal(r.ids.cpu().numpy(), post_oracle.canonical(z["ids_raw"]))                                 # same partition
    # with pruning (inference.py:307-310), through the drop-in call
    pred, ids, k = pp.decide_and_cluster({"classified_edges": [logits * 0, logits]}, b, prune=True)
    assert np.array_equal(pred.cpu().numpy(), z["pruned"])
    assert k == int(z["n_pruned"]) and isinstance(ids, np.ndarray) and np.array_equal(ids, post_oracle.canonical(z["ids_pruned"]))
    # ground-truth clusters (inference.py:297-299)
    ids, k = pp.ground_truth_clusters(b)
    assert k == int(z["n_gt"]) and np.array_equal(ids, post_oracle.canonical(z["ids_gt"]))
    # re-clustering of host-edited decisions (after rounding / splitting): kind "flags"
    r2 = pp.check(pp.frame_clusters(torch.from_numpy(z["pruned"]).to(cuda_device), b.edge_index, b.ptr, ep, "flags", prune=True))
    assert np.array_equal(r2.predictions.cpu().numpy(), z["pruned"]) and int(r2.n_clusters) == int(z["n_pruned"])
    assert r.active_edges(b.edge_index) == [(int(z["edge_index"][0][e]), int(z["edge_index"][1][e])) for e in np.nonzero(z["predictions"])[0]]


def test_cfg2_sized_batch_against_the_oracle(cuda_device):
    """1024 frames of 4 cameras x 6 detections with directed noise: asymmetric decisions, chains, 64-node frames."""
    import gnncca
    from gnncca import postprocess as pp
    syn = __import__("importlib").import_module("gnn-cca_b200.synthetic")
    b = gnncca.build_graph_batch(**syn.make_detections(4, 6, 256, seed=5, device=cuda_device, max_dist=1e9))
    g = torch.Generator().manual_seed(9)
    logits = ((b.edge_labels.cpu() * 2 - 1) * 1.5 + torch.randn(b.num_edges, generator=g) * 2.0).to(cuda_device)
    ei = b.edge_index.cpu().numpy()
    for prune in (False, True):
        r = pp.check(pp.frame_clusters(logits, b.edge_index, b.ptr, b.edge_ptr, "logits", prune=prune))
        ref_pred = post_oracle.decisions(logits.cpu().numpy())
        if prune:
            ref_pred = post_oracle.remove_edges_single_direction(ref_pred, ei)
        ref_ids, ref_k = post_oracle.clusters(b.num_nodes, ref_pred, ei)
        assert np.array_equal(r.predictions.cpu().numpy(), ref_pred)
        assert int(r.n_clusters) == ref_k and np.array_equal(r.ids.cpu().numpy(), ref_ids)


def test_directed_cycles_and_wide_frames(cuda_device):
    """Hand-made frames: a directed 3-cycle (one SCC without any bidirectional pair), a chain (no SCC), a 64-node ring
    (the widest frame the kernel takes) and an empty frame; pruning dissolves the cycle and the ring."""
    from gnncca import postprocess as pp
    edges, ptr, eptr = [], [0], [0]
    edges += [(0, 1), (1, 2), (2, 0), (3, 4), (4, 3), (2, 3)]; ptr.append(6); eptr.append(len(edges))          # nodes 0..5 (5 isolated)
    edges += [(6, 7), (7, 8)]; ptr.append(9); eptr.append(len(edges))                                            # chain
    ptr.append(9); eptr.append(len(edges))                                                                       # empty frame
    edges += [(9 + i, 9 + (i + 1) % 64) for i in range(64)]; ptr.append(73); eptr.append(len(edges))            # ring
    ei = torch.tensor(edges, dtype=torch.int64).t().contiguous().to(cuda_device)
    flags = torch.ones(len(edges), dtype=torch.int64, device=cuda_device)
    P, EP = torch.tensor(ptr, device=cuda_device), torch.tensor(eptr, device=cuda_device)
    r = pp.check(pp.frame_clusters(flags, ei, P, EP, "flags", prune=False))
    ref_ids, ref_k = post_oracle.clusters(73, np.ones(len(edges), np.int64), ei.cpu().numpy())
    assert ref_k == 1 + 1 + 1 + 3 + 1 and int(r.n_clusters) == ref_k
    assert np.array_equal(r.ids.cpu().numpy(), ref_ids)
    r = pp.check(pp.frame_clusters(flags, ei, P, EP, "flags", prune=True))
    kept = post_oracle.remove_edges_single_direction(np.ones(len(edges), np.int64), ei.cpu().numpy())
    assert kept.sum() == 2 and np.array_equal(r.predictions.cpu().numpy(), kept)
    ref_ids, ref_k = post_oracle.clusters(73, kept, ei.cpu().numpy())
    assert int(r.n_clusters) == ref_k == 72 and np.array_equal(r.ids.cpu().numpy(), ref_ids)


def test_bad_frames_are_refused(cuda_device):
    from gnncca import postprocess as pp
    ei = torch.tensor([[0, 1, 2], [1, 0, 3]], dtype=torch.int64, device=cuda_device)      # edge 2 -> 3 crosses the frame cut at 3
    x = torch.ones(3, device=cuda_device)
    with pytest.raises(IndexError):
        pp.check(pp.frame_clusters(x, ei, torch.tensor([0, 3, 4], device=cuda_device), torch.tensor([0, 3, 3], device=cuda_device)))
    big = torch.tensor([[0], [64]], dtype=torch.int64, device=cuda_device)
    with pytest.raises(NotImplementedError):
        pp.check(pp.frame_clusters(torch.ones(1, device=cuda_device), big, torch.tensor([0, 65], device=cuda_device)))
    with pytest.raises(ValueError):
        pp.frame_clusters(x, ei, torch.tensor([0, 4], device=cuda_device), torch.tensor([0, 2], device=cuda_device))
    with pytest.raises(RuntimeError):
        pp.frame_clusters(torch.ones(3), ei.cpu(), torch.tensor([0, 4]))
