"""GPU parity of graph construction + CSR against vectors produced by the reference's own loop."""
import numpy as np
import pytest
import torch

from conftest import golden_cases, load_golden

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("case", golden_cases("graph_"))
def test_build_graph_batch_matches_reference_loop(case, cuda_device):
    import gnncca
    z = load_golden("graph_" + case)
    t = lambda k, dt=None: torch.from_numpy(z[k]).to(cuda_device) if dt is None else torch.from_numpy(z[k]).to(cuda_device, dt)
    b = gnncca.build_graph_batch(t("graph_ptr"), t("id_cam"), t("id"), t("xw"), t("yw"), t("max_dist"), t("reid"), t("node"),
                                 mode=str(z["mode"]), l2norm=True)
    assert b.edge_index.dtype == torch.int64
    assert np.array_equal(b.edge_index.cpu().numpy(), z["out_edge_index"])       # bit-exact
    assert np.array_equal(b.edge_labels.cpu().numpy(), z["out_edge_labels"])     # bit-exact
    assert np.array_equal(b.y.cpu().numpy(), z["out_y"])
    assert np.array_equal(b.ptr.cpu().numpy(), z["out_ptr"])
    assert np.array_equal(b.batch.cpu().numpy(), z["out_batch"])
    # floats: column-normalised embeddings and the four attributes, fp32 tolerance
    np.testing.assert_allclose(b.x.cpu().numpy(), z["out_x"], rtol=2e-6, atol=1e-9)
    np.testing.assert_allclose(b.edge_attr.cpu().numpy(), z["out_edge_attr"], rtol=1e-5, atol=2e-6)


@pytest.mark.parametrize("sorted_rows", [True, False])
def test_csr_is_bit_exact(sorted_rows, cuda_device):
    import gnncca
    rng = np.random.default_rng(7)
    N, E = 1000, 20000
    row = rng.integers(0, N, E)
    row[row == 17] = 18            # node 17 has no out-edge
    if sorted_rows:
        row = np.sort(row)
    col = rng.integers(0, N, E)
    ei = torch.from_numpy(np.stack([row, col])).to(cuda_device)
    plan = gnncca.GraphPlan(ei, N).resolve()
    order = np.argsort(row, kind="stable")
    assert (plan.perm is None) == sorted_rows
    if not sorted_rows:
        assert np.array_equal(plan.perm.cpu().numpy(), order)
    assert np.array_equal(plan.src.cpu().numpy(), row[order])
    assert np.array_equal(plan.dst.cpu().numpy(), col[order])
    rp = np.concatenate([[0], np.cumsum(np.bincount(row, minlength=N))])
    assert np.array_equal(plan.rowptr.cpu().numpy(), rp)


def test_csr_rejects_out_of_range(cuda_device):
    import gnncca
    ei = torch.tensor([[0, 1, 2], [1, 2, 5]], device=cuda_device)
    with pytest.raises(IndexError):
        gnncca.GraphPlan(ei, 3).resolve()


def test_full_size_properties_cfg2(cuda_device):
    """BASELINE cfg 2 (4 cams x 6, 1024 frames) at full size through size-independent properties:
    edge count closed form, source-major order, both directions present, labels symmetric."""
    import gnncca
    from importlib import import_module
    syn = import_module("gnn-cca_b200.synthetic")
    det = syn.make_detections(4, 6, 1024, seed=1236, device=cuda_device, node_dim=64, max_dist=44.23)
    b = gnncca.build_graph_batch(**det)
    E = 1024 * syn.edges_per_graph(4, 6)
    assert b.num_edges == E == 442368
    src, dst = b.edge_index[0], b.edge_index[1]
    assert bool((src[1:] >= src[:-1]).all())
    assert bool((det["id_cam"][src] != det["id_cam"][dst]).all())
    key = src * b.num_nodes + dst
    rkey = dst * b.num_nodes + src
    assert torch.equal(torch.sort(key).values, torch.sort(rkey).values)        # every edge has its reverse
    lab = b.edge_labels
    assert torch.equal(lab, (det["ids"][src] == det["ids"][dst]).float())
    assert abs(float(lab.mean()) - 3 / 18) < 1e-6                              # C-1 positives per node
    assert bool((b.edge_attr[:, :3] >= 0).all()) and bool((b.edge_attr[:, 3].abs() <= 1 + 1e-5).all())


@pytest.mark.parametrize("reid_dim,mode,lazy", [(256, "full", True), (64, "full", False), (6, "full", True), (20, "only_appearance", True),
                                                (256, "only_dist", True), (1028, "full", True)])
def test_every_launch_structure_of_build_graph_batch_matches_the_oracle(reid_dim, mode, lazy, cuda_device):
    """The single prepare call (ReID norms fused into the frame-table launches, node norms on the side stream, totals polled from
    pinned memory, ReID rows normalised while staged) and the separate-launch path taken for ReID widths it does not accept
    (6: not a multiple of 4; 1028: wider than 1024) give what the CPU restatement of the reference loop gives; ragged frames of
    up to ~60 detections so that one-tile frames, diagonal and off-diagonal tiles of larger frames and an empty camera all occur.
    Two builds in a row: the second one finds the previous batch's size as its output-size guess (wrong on purpose)."""
    import gnncca
    from oracle import graph_oracle
    rng = np.random.default_rng(reid_dim)
    for frames_n in (7, 5):
        frames, ptr = [], [0]
        for _ in range(frames_n):
            counts = rng.integers(0, 15, 5)
            counts[rng.integers(0, 5)] += 1
            cam = np.repeat(np.arange(5), counts).astype(np.int64)
            n = len(cam)
            frames.append(dict(id=rng.integers(0, 12, n).astype(np.int64), id_cam=cam, xw=rng.random(n) * 300, yw=rng.random(n) * 200))
            ptr.append(ptr[-1] + n)
        N = ptr[-1]
        g = torch.Generator().manual_seed(N)
        reid = torch.relu(torch.randn((N, reid_dim), generator=g))
        node = torch.relu(torch.randn((N, 96), generator=g))
        md = [50.83 + k for k in range(frames_n)]
        ref = graph_oracle.build_batch(frames, reid, node, md, mode=mode)
        cat = lambda k, dt: torch.from_numpy(np.concatenate([f[k] for f in frames])).to(cuda_device, dt)
        b = gnncca.build_graph_batch(torch.tensor(ptr, device=cuda_device), cat("id_cam", torch.int32), cat("id", torch.int64),
                                     cat("xw", torch.float64), cat("yw", torch.float64), torch.tensor(md, dtype=torch.float64, device=cuda_device),
                                     reid.to(cuda_device), node.to(cuda_device), mode=mode, lazy_x=lazy)
        assert (b.raw_x() is not None) == lazy
        assert torch.equal(b.edge_index.cpu(), ref.edge_index)
        assert torch.equal(b.edge_labels.cpu(), ref.edge_labels.float())
        np.testing.assert_allclose(b.edge_attr.cpu().numpy(), ref.edge_attr.numpy(), rtol=1e-5, atol=2e-6)
        np.testing.assert_allclose(b.x.cpu().numpy(), ref.x.numpy(), rtol=2e-6, atol=1e-9)
        assert b.gnncca_caps[0] == max(len(f["id"]) for f in frames)


def test_reverse_edge_table_equals_the_stable_sort_by_target(cuda_device):
    """gnncca_csr_reverse_edges on a build_graph_batch output (symmetric, source-major, targets ascending) gives the permutation
    the radix sort by target gives, with rowptr as offsets; an edge without its reverse raises flags[3]."""
    import importlib
    import gnncca
    _lib = importlib.import_module("gnn-cca_b200._lib")
    syn = importlib.import_module("gnn-cca_b200.synthetic")
    lib = _lib.load()
    b = gnncca.build_graph_batch(**syn.make_detections(5, 7, 9, seed=3, device=cuda_device))
    N, E = b.num_nodes, b.num_edges
    plan = gnncca.GraphPlan(b.edge_index, N).resolve()
    assert plan.perm is None
    st = torch.cuda.current_stream().cuda_stream
    tperm = torch.empty(E, dtype=torch.int32, device=cuda_device)
    flags = torch.zeros(4, dtype=torch.int32, device=cuda_device)
    _lib.check(lib.gnncca_csr_reverse_edges(_lib.ptr(plan.rowptr), _lib.ptr(plan.src), _lib.ptr(plan.dst), N, E, _lib.ptr(tperm), _lib.ptr(flags), st), "rev")
    key, other = plan.dst.clone(), plan.src.clone()
    want = torch.empty(E, dtype=torch.int32, device=cuda_device)
    colptr = torch.empty(N + 1, dtype=torch.int32, device=cuda_device)
    nbytes = int(lib.gnncca_csr_sort_workspace_bytes(N, E))
    ws = torch.empty(nbytes, dtype=torch.uint8, device=cuda_device)
    _lib.check(lib.gnncca_csr_sort(N, E, _lib.ptr(key), _lib.ptr(other), _lib.ptr(want), _lib.ptr(colptr), _lib.ptr(ws), nbytes, st), "sort")
    assert torch.equal(tperm, want) and torch.equal(colptr, plan.rowptr) and int(flags.sum()) == 0
    # drop one direction of one pair: no longer symmetric
    ei = b.edge_index[:, 1:].contiguous()
    plan2 = gnncca.GraphPlan(ei, N).resolve()
    t2 = torch.empty(E - 1, dtype=torch.int32, device=cuda_device)
    _lib.check(lib.gnncca_csr_reverse_edges(_lib.ptr(plan2.rowptr), _lib.ptr(plan2.src), _lib.ptr(plan2.dst), N, E - 1, _lib.ptr(t2), _lib.ptr(flags), st), "rev")
    assert int(flags[3]) == 1
