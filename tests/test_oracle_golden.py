"""Pin the oracle restatements against vectors produced by the reference itself (tests/golden)."""
import json

import numpy as np
import pytest
import torch

from conftest import golden_cases, load_golden
from oracle import graph_oracle, mpn_oracle


def _sd(z):
    return {k[3:]: torch.from_numpy(v) for k, v in z.items() if k.startswith("sd/")}


@pytest.mark.parametrize("case", golden_cases("mpn_"))
def test_mpn_oracle_matches_reference_vectors(case):
    z = load_golden("mpn_" + case)
    spec = mpn_oracle.MPNSpec(json.loads(str(z["config"])), str(z["arch"]))
    sd = _sd(z)
    out = mpn_oracle.motmpnet_forward(spec, sd, torch.from_numpy(z["x"]), torch.from_numpy(z["edge_index"]),
                                      torch.from_numpy(z["edge_attr"]), training=bool(z["train"]))
    res = out["classified_edges"]
    assert len(res) == int(z["n_steps"])
    for i, r in enumerate(res):
        ref = torch.from_numpy(z["logits/%d" % i])
        assert r.shape == ref.shape and r.dtype == torch.float32
        # same torch CPU ops in the same order -> expected bit-equal; allow 1 ulp-ish slack for BN restatement
        torch.testing.assert_close(r, ref, rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("case", golden_cases("graph_"))
def test_graph_oracle_matches_reference_loop(case):
    z = load_golden("graph_" + case)
    ptr = z["graph_ptr"]
    frames = [dict(id=z["id"][a:b], id_cam=z["id_cam"][a:b], xw=z["xw"][a:b], yw=z["yw"][a:b])
              for a, b in zip(ptr[:-1], ptr[1:])]
    b = graph_oracle.build_batch(frames, torch.from_numpy(z["reid"]), torch.from_numpy(z["node"]), z["max_dist"],
                                 mode=str(z["mode"]), l2norm=True)  # inference.py:189-190 normalises unconditionally
    assert b.edge_index.dtype == torch.int64
    assert np.array_equal(b.edge_index.numpy(), z["out_edge_index"])          # bit-exact
    assert np.array_equal(b.edge_labels.numpy(), z["out_edge_labels"])        # bit-exact
    assert np.array_equal(b.y.numpy(), z["out_y"])
    assert np.array_equal(b.ptr.numpy(), z["out_ptr"])
    assert np.array_equal(b.batch.numpy(), z["out_batch"])
    assert np.array_equal(b.x.numpy(), z["out_x"])
    assert np.array_equal(b.edge_attr.numpy(), z["out_edge_attr"])            # same CPU ops -> bit-equal


def test_edge_enumeration_closed_form_when_camera_major():
    """with camera-major rows the enumeration is the row-major mask cam[i] != cam[j]."""
    for case in ("terrace_real", "laboratory_real", "synthetic_terrace_full"):
        z = load_golden("graph_" + case)
        ptr = z["graph_ptr"]
        parts = []
        for a, b in zip(ptr[:-1], ptr[1:]):
            cam = z["id_cam"][a:b]
            assert np.all(np.diff(cam) >= 0)
            i, j = np.nonzero(cam[:, None] != cam[None, :])
            parts.append(np.stack([i + a, j + a]))
        assert np.array_equal(np.concatenate(parts, axis=1), z["out_edge_index"])
        assert np.all(np.diff(z["out_edge_index"][0]) >= 0)


def test_scatter_restatements():
    src = torch.tensor([[1., -2.], [3., 4.], [-5., 6.], [7., -8.]])
    idx = torch.tensor([2, 0, 2, 0])
    assert torch.equal(mpn_oracle.scatter_add(src, idx, 0, 4), torch.tensor([[10., -4.], [0., 0.], [-4., 4.], [0., 0.]]))
    assert torch.equal(mpn_oracle.scatter_mean(src, idx, 0, 4), torch.tensor([[5., -2.], [0., 0.], [-2., 2.], [0., 0.]]))
    v, a = mpn_oracle.scatter_max(src, idx, 0, 4)
    assert torch.equal(v, torch.tensor([[7., 4.], [0., 0.], [1., 6.], [0., 0.]]))
    assert torch.equal(a, torch.tensor([[3, 1], [4, 4], [0, 2], [4, 4]]))


@pytest.mark.parametrize("case", [c for c in golden_cases("mpn_") if c.endswith("_train")])
def test_mpn_oracle_gradients_match_reference_vectors(case):
    """train.py:80-97,493: loss = sum over classified steps of BCEWithLogits; backward to every parameter."""
    z = load_golden("mpn_" + case)
    spec = mpn_oracle.MPNSpec(json.loads(str(z["config"])), str(z["arch"]))
    sd = _sd(z)
    for k, v in sd.items():
        if v.is_floating_point() and "running_" not in k:
            v.requires_grad_(True)
    out = mpn_oracle.motmpnet_forward(spec, sd, torch.from_numpy(z["x"]), torch.from_numpy(z["edge_index"]),
                                      torch.from_numpy(z["edge_attr"]), training=True)
    labels = torch.from_numpy(z["labels"])
    crit = torch.nn.BCEWithLogitsLoss()
    loss = sum(crit(r.view(-1), labels) for r in out["classified_edges"])
    loss.backward()
    assert abs(loss.item() - float(z["loss"])) < 1e-6
    n = 0
    for k, v in z.items():
        if k.startswith("grad/"):
            torch.testing.assert_close(sd[k[5:]].grad, torch.from_numpy(v), rtol=1e-5, atol=1e-7)
            n += 1
    assert n >= 12


def test_oracle_matches_live_reference_when_present():
    """In the build container the unmodified reference module is importable: compare directly."""
    from oracle import ref_shim
    if not ref_shim.reference_available():
        pytest.skip("/root/reference not present (GPU box)")
    import copy
    mpn = ref_shim.import_reference("models.mpn")
    z = load_golden("mpn_shipped_bdnet_train")
    params = json.loads(str(z["config"]))
    torch.manual_seed(123)
    model = mpn.MOTMPNet(copy.deepcopy(params), None, "bdnet_market").eval()
    spec = mpn_oracle.MPNSpec(params, "bdnet_market")
    x, ei, ea = torch.from_numpy(z["x"]), torch.from_numpy(z["edge_index"]), torch.from_numpy(z["edge_attr"])
    with torch.no_grad():
        ref = model(graph_oracle.Data(x=x, edge_index=ei, edge_attr=ea))["classified_edges"]
        got = mpn_oracle.motmpnet_forward(spec, model.state_dict(), x, ei, ea)["classified_edges"]
    assert len(ref) == len(got) == 3
    for a, b in zip(ref, got):
        assert torch.equal(a, b)


@pytest.mark.parametrize("case", golden_cases("loss_"))
def test_loss_oracle_matches_the_reference_function(case):
    """oracle/loss_oracle.py against compute_loss_acc (train.py:51-208) executed unmodified (oracle/gen_golden.py gen_loss)."""
    from oracle import loss_oracle

    z = load_golden("loss_" + case)
    k = int(z["k"])
    logits = [torch.from_numpy(z["logits%d" % s]).requires_grad_(True) for s in range(k)]
    pw = None if float(z["pos_weight"]) < 0 else float(z["pos_weight"])
    r = loss_oracle.compute_loss_acc(logits, torch.from_numpy(z["labels"]), str(z["criterion"]) or "BCE", pw, str(z["mode"]))
    torch.testing.assert_close(r["loss"].detach(), torch.from_numpy(z["loss"]), rtol=1e-6, atol=1e-7)
    for name in ("loss_class1", "loss_class0"):
        torch.testing.assert_close(r[name].detach(), torch.from_numpy(z[name]), rtol=1e-6, atol=1e-7, equal_nan=True)
    for name in ("precision_class1", "precision_class0", "precision_all"):
        assert r[name] == float(z[name]), name                                     # integer counts: exact
    grads = torch.autograd.grad(r["loss"], logits)
    for s in range(k):
        assert np.array_equal(r["probs"][s].numpy(), z["probs%d" % s].reshape(-1))
        torch.testing.assert_close(grads[s], torch.from_numpy(z["grad%d" % s]), rtol=1e-6, atol=1e-9)


@pytest.mark.parametrize("case", golden_cases("post_"))
def test_post_oracle_matches_the_reference_functions(case):
    """oracle/post_oracle.py against utils.remove_edges_single_direction / compute_SCC_and_Clusters run unmodified with
    networkx (oracle/gen_golden.py gen_post).  Cluster ids are compared as partitions (canonical numbering)."""
    from oracle import post_oracle

    z = load_golden("post_" + case)
    n = int(z["ptr"][-1])
    pred = post_oracle.decisions(z["logits"])
    assert np.array_equal(pred, z["predictions"])
    ids, k = post_oracle.clusters(n, pred, z["edge_index"])
    assert k == int(z["n_raw"]) and np.array_equal(ids, post_oracle.canonical(z["ids_raw"]))
    pruned = post_oracle.remove_edges_single_direction(pred, z["edge_index"])
    assert np.array_equal(pruned, z["pruned"])
    ids, k = post_oracle.clusters(n, pruned, z["edge_index"])
    assert k == int(z["n_pruned"]) and np.array_equal(ids, post_oracle.canonical(z["ids_pruned"]))
    ids, k = post_oracle.clusters(n, (z["labels"] == 1).astype(np.int64), z["edge_index"])
    assert k == int(z["n_gt"]) and np.array_equal(ids, post_oracle.canonical(z["ids_gt"]))


@pytest.mark.parametrize("case", golden_cases("dets_"))
def test_detections_oracle_matches_the_reference_source(case):
    """oracle/detections_oracle.py against the reference's own homography function and valid-frame loop on real EPFL ground truth."""
    from oracle import detections_oracle

    z = load_golden("dets_" + case)
    xw, yw = detections_oracle.world_coordinates(z["xmin"], z["xmax"], z["ymax"], z["id_cam"], z["homographies"])
    np.testing.assert_allclose(xw, z["xw"], rtol=1e-12, atol=0)
    np.testing.assert_allclose(yw, z["yw"], rtol=1e-12, atol=0)
    assert np.array_equal(detections_oracle.valid_frames(z["frame"], z["id_cam"], z["id"]), z["frames_valid"])
    rows = np.concatenate([detections_oracle.frame_rows(z["frame"], f) for f in z["some_frames"]])
    assert np.array_equal(rows, z["some_rows"])
