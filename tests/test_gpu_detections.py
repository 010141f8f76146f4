"""GPU parity of the detection-table kernels (gnncca_foot_to_world, gnncca_frame_valid) against real EPFL
ground truth processed by the reference's own source (tests/golden/dets_*.npz) and against the CPU oracle."""
import numpy as np
import pytest
import torch

from conftest import golden_cases, load_golden
from oracle import detections_oracle

pytestmark = pytest.mark.gpu
KEYS = ("id", "xmin", "ymin", "xmax", "ymax", "frame", "id_cam")


@pytest.mark.parametrize("case", golden_cases("dets_"))
def test_real_sequences_match_the_reference(case, cuda_device):
    import gnncca
    from gnncca import detections as D
    z = load_golden("dets_" + case)
    n0 = gnncca._lib.load().gnncca_launch_count()
    seq = D.prepare_sequence(**{k: z[k] for k in KEYS}, homographies=z["homographies"], device=cuda_device)
    # world coordinates: the same float64 operations in the same order as the oracle -> bit-equal; reference dot product (BLAS): 1e-12
    oxw, oyw = detections_oracle.world_coordinates(z["xmin"], z["xmax"], z["ymax"], z["id_cam"], z["homographies"])
    assert np.array_equal(seq.xw.cpu().numpy(), oxw) and np.array_equal(seq.yw.cpu().numpy(), oyw)
    np.testing.assert_allclose(seq.xw.cpu().numpy(), z["xw"], rtol=1e-12, atol=0)
    np.testing.assert_allclose(seq.yw.cpu().numpy(), z["yw"], rtol=1e-12, atol=0)
    assert np.array_equal(seq.frames_valid.cpu().numpy(), z["frames_valid"])                 # bit-exact
    tab = seq.gather(z["some_frames"].tolist())
    assert np.array_equal(tab["rows"].cpu().numpy(), z["some_rows"])                          # frame after frame, table order inside
    counts = [int((z["frame"] == f).sum()) for f in z["some_frames"]]
    assert tab["graph"]["graph_ptr"].cpu().tolist() == np.concatenate([[0], np.cumsum(counts)]).tolist()
    assert np.array_equal(tab["graph"]["ids"].cpu().numpy(), z["id"][z["some_rows"]])
    assert np.array_equal(tab["graph"]["id_cam"].cpu().numpy(), z["id_cam"][z["some_rows"]])


def test_gathered_frames_feed_graph_construction(cuda_device):
    """Detection table -> gather -> build_graph_batch: the same batch as building it from the oracle's per-frame tables."""
    import gnncca
    from gnncca import detections as D
    from oracle import graph_oracle
    z = load_golden("dets_terrace")
    seq = D.prepare_sequence(**{k: z[k] for k in KEYS}, homographies=z["homographies"], device=cuda_device)
    frames = z["frames_valid"][5:45:4].tolist()
    tab = seq.gather(frames)
    rows = tab["rows"]
    g = torch.Generator().manual_seed(2)
    M = len(z["id"])
    reid, node = torch.relu(torch.randn(M, 64, generator=g)), torch.relu(torch.randn(M, 96, generator=g))
    b = gnncca.build_graph_batch(max_dist=torch.full((len(frames),), 50.83, dtype=torch.float64), reid_embeds=reid[rows.cpu()].to(cuda_device),
                                 node_embeds=node[rows.cpu()].to(cuda_device), **tab["graph"])
    oxw, oyw = detections_oracle.world_coordinates(z["xmin"], z["xmax"], z["ymax"], z["id_cam"], z["homographies"])
    fr = []
    for f in frames:
        r = detections_oracle.frame_rows(z["frame"], f)
        fr.append(dict(id=z["id"][r], id_cam=z["id_cam"][r], xw=oxw[r], yw=oyw[r]))
    rr = rows.cpu()
    ref = graph_oracle.build_batch(fr, reid[rr], node[rr], [50.83] * len(frames), mode="full", l2norm=True)
    assert torch.equal(b.edge_index.cpu(), ref.edge_index) and torch.equal(b.edge_labels.cpu(), ref.edge_labels)
    np.testing.assert_allclose(b.edge_attr.cpu().numpy(), ref.edge_attr.numpy(), rtol=1e-5, atol=2e-6)


def test_synthetic_tables_and_errors(cuda_device):
    from gnncca import detections as D
    rng = np.random.default_rng(4)
    M = 5000
    t = dict(id=rng.integers(0, 9, M), xmin=rng.integers(0, 300, M), ymin=rng.integers(0, 100, M), frame=rng.integers(10, 700, M),
             id_cam=np.sort(rng.integers(0, 4, M)))
    t["xmax"] = t["xmin"] + rng.integers(5, 60, M)
    t["ymax"] = t["ymin"] + rng.integers(20, 180, M)
    H = rng.normal(size=(4, 3, 3)) + np.eye(3) * 5
    seq = D.prepare_sequence(**t, homographies=H, device=cuda_device)
    oxw, oyw = detections_oracle.world_coordinates(t["xmin"], t["xmax"], t["ymax"], t["id_cam"], H)
    assert np.array_equal(seq.xw.cpu().numpy(), oxw) and np.array_equal(seq.yw.cpu().numpy(), oyw)
    assert np.array_equal(seq.frames_valid.cpu().numpy(), detections_oracle.valid_frames(t["frame"], t["id_cam"], t["id"]))
    f = [int(v) for v in seq.frames_valid[:20].tolist()]
    rows = np.concatenate([detections_oracle.frame_rows(t["frame"], x) for x in f])
    assert np.array_equal(seq.gather(f)["rows"].cpu().numpy(), rows)
    with pytest.raises(IndexError):
        seq.gather([9])
    bad = dict(t); bad["id_cam"] = t["id_cam"] + 2
    with pytest.raises(IndexError):
        D.prepare_sequence(**bad, homographies=H, device=cuda_device)
    with pytest.raises(RuntimeError):
        D.prepare_sequence(**t, homographies=H, device="cpu")
