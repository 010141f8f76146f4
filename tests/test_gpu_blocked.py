"""Frame-blocked execution of large batches (engine.forward_blocked; BASELINE config 4 runs through it), the trust contract of
``build_graph_batch`` outputs (interleaved cameras are NOT source-major), and the violation checks of the trusted paths."""
import copy
import importlib

import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import mpn_oracle

pytestmark = pytest.mark.gpu

engine = importlib.import_module("gnn-cca_b200.engine")
cfgmod = importlib.import_module("gnn-cca_b200.config")
syn = importlib.import_module("gnn-cca_b200.synthetic")


def _model(dev, math="fp32", agg="sum", arch="bdnet_market", extra=None):
    import gnncca
    params = cfgmod.graph_net_params()
    params["node_agg_fn"] = agg
    if extra:
        params.update(extra)
    torch.manual_seed(0)
    return gnncca.MOTMPNet(copy.deepcopy(params), None, arch, math=math).to(dev).eval(), params


def _ragged_batch(dev, frames=23, seed=3, node_dim=512, cams=5, max_dets=40):
    """Frames of 5 cameras with 1..max_dets detections per camera (camera-major rows): a few hundred nodes per frame, too large
    for the fused small-frame kernel, ragged edge counts so that block boundaries fall anywhere inside a reduction tile."""
    import gnncca
    rng = np.random.default_rng(seed)
    cam, ids, ptr = [], [], [0]
    for _ in range(frames):
        counts = rng.integers(1, max_dets + 1, cams)
        cam.append(np.repeat(np.arange(cams), counts))
        ids.append(np.concatenate([rng.permutation(max_dets)[:c] for c in counts]))
        ptr.append(ptr[-1] + int(counts.sum()))
    N = ptr[-1]
    g = torch.Generator().manual_seed(seed)
    return gnncca.build_graph_batch(graph_ptr=torch.tensor(ptr, device=dev), id_cam=torch.from_numpy(np.concatenate(cam)).to(dev),
                                    ids=torch.from_numpy(np.concatenate(ids)).to(dev), xw=torch.rand(N, generator=g, dtype=torch.float64).to(dev) * 300,
                                    yw=torch.rand(N, generator=g, dtype=torch.float64).to(dev) * 200,
                                    max_dist=torch.full((frames,), 50.83, dtype=torch.float64, device=dev),
                                    reid_embeds=torch.relu(torch.randn(N, 256, generator=g)).to(dev),
                                    node_embeds=torch.relu(torch.randn(N, node_dim, generator=g)).to(dev))


@pytest.mark.parametrize("agg", ["sum", "max"])
@pytest.mark.parametrize("target", [4000, 30000, 10 ** 9])
def test_blocked_forward_is_bit_identical_to_the_unblocked_steps(target, agg, cuda_device):
    model, _ = _model(cuda_device, agg=agg)
    b = _ragged_batch(cuda_device)
    assert b.gnncca_trusted and b.gnncca_caps[0] > 32
    with torch.no_grad():
        whole = engine.forward_inference(model, b.x, b.edge_index, b.edge_attr, graph_ptr=b.ptr, trusted=True, _no_block=True, allow_fused=False)
        for trusted, eptr in ((True, b.edge_ptr), (False, None)):      # offsets given / found by a search in edge_index[0]
            got = engine.forward_blocked(model, b.x, b.edge_index, b.edge_attr, b.ptr, eptr, trusted=trusted, target_edges=target)
            assert len(got) == 3
            for a, c in zip(got, whole):
                assert a.shape == c.shape and torch.equal(a, c)


def test_blocked_forward_matches_the_oracle_and_is_what_the_module_runs_above_the_threshold(cuda_device, monkeypatch):
    model, params = _model(cuda_device)
    b = _ragged_batch(cuda_device, frames=9, seed=5)
    monkeypatch.setattr(engine, "BLOCKED_MIN_EDGES", 1000)
    monkeypatch.setattr(engine, "BLOCK_EDGES", 20000)
    with torch.no_grad():
        out = model(b)["classified_edges"]
    assert model.last_plan is engine.BLOCKED_PLAN and len(b._gnncca_blocks[1]) > 1
    spec = mpn_oracle.MPNSpec(params, "bdnet_market")
    sd = {k: v.cpu() for k, v in model.state_dict().items()}
    with torch.no_grad():
        ref = mpn_oracle.motmpnet_forward(spec, sd, b.x.cpu(), b.edge_index.cpu(), b.edge_attr.cpu())["classified_edges"]
    for o, r in zip(out, ref):
        err = ((o.cpu() - r).abs() / (1 + r.abs() + 0.01 * r.abs().max())).max()
        assert float(err) <= 1e-4, float(err)


def test_blocked_forward_flags_frames_that_are_not_closed(cuda_device):
    model, _ = _model(cuda_device)
    b = _ragged_batch(cuda_device, frames=6, seed=7)
    ei = b.edge_index.clone()
    ei[1, 5] = b.num_nodes - 1           # an edge of the first frame that ends in the last frame
    with torch.no_grad(), pytest.raises(IndexError):
        engine.forward_blocked(model, b.x, ei, b.edge_attr, b.ptr, b.edge_ptr, trusted=False, target_edges=5000)


# ------------------------------------------------------------------------------------------------------------
# the trust contract
# ------------------------------------------------------------------------------------------------------------
def _interleaved_batch(dev):
    import gnncca
    z = load_golden("graph_synthetic_interleaved_cams")
    t = lambda k: torch.from_numpy(z[k]).to(dev)
    return gnncca.build_graph_batch(t("graph_ptr"), t("id_cam"), t("id"), t("xw"), t("yw"), t("max_dist"), t("reid"), t("node"), mode="full"), z


def test_interleaved_camera_frames_are_not_trusted_and_run_on_the_checked_path(cuda_device):
    """Rows that are not in ascending camera order give an edge_index[0] that is not non-decreasing (the reference enumerates
    camera by camera, train.py:292-299): such a batch must NOT be marked source-major.  Forward and training step against the
    oracle on exactly that batch."""
    import gnncca
    b, z = _interleaved_batch(cuda_device)
    src = b.edge_index[0]
    assert not bool((src[1:] >= src[:-1]).all())          # the premise
    assert b.gnncca_trusted is False
    assert np.array_equal(b.edge_index.cpu().numpy(), z["out_edge_index"])
    params = cfgmod.graph_net_params()
    params["encoder_feats_dict"]["nodes"]["tiny"] = dict(node_in_dim=64, node_fc_dims=[128], node_out_dim=32, dropout_p=0, use_batchnorm=False)
    spec = mpn_oracle.MPNSpec(params, "tiny")
    for math in ("fp32", "tf32"):
        torch.manual_seed(1)
        model = gnncca.MOTMPNet(copy.deepcopy(params), None, "tiny", math=math).to(cuda_device).eval()
        sd = {k: v.detach().cpu() for k, v in model.state_dict().items()}
        with torch.no_grad():
            out = model(b)["classified_edges"]
            ref = mpn_oracle.motmpnet_forward(spec, sd, b.x.cpu(), b.edge_index.cpu(), b.edge_attr.cpu())["classified_edges"]
        assert model.last_plan.perm is not None          # it was sorted
        tol = 1e-4 if math == "fp32" else 1e-3
        for o, r in zip(out, ref):
            assert float(((o.cpu() - r).abs() / (1 + r.abs())).max()) <= tol
    # training step (fp32): loss and every gradient
    torch.manual_seed(1)
    model = gnncca.MOTMPNet(copy.deepcopy(params), None, "tiny", math="fp32").to(cuda_device).train()
    sdr = {k: v.detach().cpu().clone().requires_grad_(v.is_floating_point()) for k, v in model.state_dict().items()}
    out = model(b)["classified_edges"]
    crit = torch.nn.BCEWithLogitsLoss()
    loss = sum(crit(o.view(-1), b.edge_labels) for o in out)
    loss.backward()
    ref = mpn_oracle.motmpnet_forward(spec, sdr, b.x.cpu(), b.edge_index.cpu(), b.edge_attr.cpu(), training=True)["classified_edges"]
    lref = sum(crit(r.view(-1), b.edge_labels.cpu()) for r in ref)
    lref.backward()
    assert abs(float(loss) - float(lref)) <= 1e-5 * (1 + abs(float(lref)))
    for name, p in model.named_parameters():
        g, r = p.grad.cpu().double(), sdr[name].grad.double()
        assert float(((g - r).abs() / (1 + r.abs() + 0.01 * r.abs().max())).max()) <= 2e-4, name


def test_trusted_fast_path_reports_a_broken_contract(cuda_device):
    """A batch from build_graph_batch whose edge_index is edited afterwards: the single-call path cannot check it up front
    (nothing is read back), but the frame is skipped on the device, its logits are NaN, and the NEXT call (or verify()) raises."""
    import gnncca
    params = cfgmod.graph_net_params()
    torch.manual_seed(0)
    model = gnncca.MOTMPNet(copy.deepcopy(params), None, "resnet50").to(cuda_device).eval()
    b = gnncca.build_graph_batch(**syn.make_detections(4, 6, 20, seed=2, device=cuda_device))
    with torch.no_grad():
        good = model(b)["classified_edges"]
        assert model.last_plan is engine.FAST_PLAN
        engine.fast_path(model).verify(wait=True)         # nothing to report
        e0, e1 = int(b.edge_ptr[3]), int(b.edge_ptr[4])
        b.edge_index[1, e0 + 2] = 0                        # frame 3 now has an edge that leaves it
        bad = model(b)["classified_edges"]
        torch.cuda.synchronize()
        for k in range(3):
            assert bool(torch.isnan(bad[k][e0:e1]).all())
            assert torch.equal(bad[k][:e0], good[k][:e0]) and torch.equal(bad[k][e1:], good[k][e1:])
        with pytest.raises(RuntimeError):
            model(b)
        model(b)                                           # the flag was consumed: the call goes through again (and raises next time)
        with pytest.raises(RuntimeError):
            engine.fast_path(model).verify(wait=True)
