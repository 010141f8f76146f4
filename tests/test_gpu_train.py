"""GPU parity of the training path (forward in train mode + backward kernels) against loss / gradients / BatchNorm
buffers produced by the reference itself (tests/golden/mpn_*_train.npz, see oracle/gen_golden.py)."""
import json

import numpy as np
import pytest
import torch

from conftest import golden_cases, load_golden
from test_gpu_mpn import D, build_model, close, frame_ptr

pytestmark = pytest.mark.gpu

GRAD_TOL = 2e-4   # |got - ref| <= 2e-4 * (1 + |ref| + 0.01 max|ref|): fp32 sums over E edges in a different (fixed) order


def run_train_case(z, dev, with_ptr=False):
    model = build_model(z, dev).train()
    ei = torch.from_numpy(z["edge_index"])
    ptr = frame_ptr(ei, z["x"].shape[0]).to(dev) if with_ptr else None
    data = D(torch.from_numpy(z["x"]).to(dev), ei.to(dev), torch.from_numpy(z["edge_attr"]).to(dev), ptr)
    labels = torch.from_numpy(z["labels"]).to(dev)
    out = model(data)["classified_edges"]
    crit = torch.nn.BCEWithLogitsLoss()
    loss = sum(crit(r.view(-1), labels) for r in out)          # train.py:80-97
    loss.backward()                                            # train.py:493
    return model, out, loss


@pytest.mark.parametrize("case", [c for c in golden_cases("mpn_") if c.endswith("_train")])
def test_loss_and_gradients_match_reference_vectors(case, cuda_device):
    z = load_golden("mpn_" + case)
    model, out, loss = run_train_case(z, cuda_device)
    assert len(out) == int(z["n_steps"])
    for i, o in enumerate(out):
        assert o.shape == (z["edge_index"].shape[1], 1) and o.requires_grad
        assert close(o, torch.from_numpy(z["logits/%d" % i])) <= 1e-4, (case, "logits", i)
    assert abs(loss.item() - float(z["loss"])) <= 1e-5 * (1 + abs(float(z["loss"])))
    n = 0
    for name, p in model.named_parameters():
        ref = torch.from_numpy(z["grad/" + name])
        assert p.grad is not None and p.grad.shape == ref.shape, name
        err = close(p.grad, ref)
        assert err <= GRAD_TOL, (case, name, err)
        n += 1
    assert n >= 12
    for k, v in z.items():                                      # BatchNorm running statistics after the step
        if k.startswith("sd_after/"):
            got = model.state_dict()[k[len("sd_after/"):]].cpu()
            if "num_batches" in k:
                assert int(got) == int(v)
            else:
                np.testing.assert_allclose(got.numpy(), v, rtol=1e-5, atol=1e-6)


def test_backward_is_bit_reproducible(cuda_device):
    z = load_golden("mpn_shipped_bdnet_train")
    grads = []
    for _ in range(3):
        model, _, _ = run_train_case(z, cuda_device)
        grads.append({n: p.grad.clone() for n, p in model.named_parameters()})
    for g in grads[1:]:
        for n in g:
            assert torch.equal(g[n], grads[0][n]), n


def test_sgd_step_changes_the_next_forward(cuda_device):
    """optimizer.step mutates the parameters in place (train.py:494): the kernels must read the live tensors."""
    z = load_golden("mpn_shipped_bdnet_train")
    model = build_model(z, cuda_device).train()
    data = D(torch.from_numpy(z["x"]).to(cuda_device), torch.from_numpy(z["edge_index"]).to(cuda_device),
             torch.from_numpy(z["edge_attr"]).to(cuda_device))
    labels = torch.from_numpy(z["labels"]).to(cuda_device)
    opt = torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4)
    crit = torch.nn.BCEWithLogitsLoss()
    losses = []
    for _ in range(8):
        opt.zero_grad()
        loss = sum(crit(r.view(-1), labels) for r in model(data)["classified_edges"])
        loss.backward()
        opt.step()
        losses.append(loss.item())
    assert losses[-1] < losses[0]


def test_cfg3_shape_gradients_match_oracle_autograd(cuda_device):
    """BASELINE cfg 3 shape (4 cams x 10, batch 64, fwd + BCE + bwd) against torch autograd through the CPU oracle."""
    import copy
    from importlib import import_module
    import gnncca
    from oracle import mpn_oracle
    from test_gpu_mpn import synthetic_graph
    params = import_module("gnn-cca_b200.config").graph_net_params()
    torch.manual_seed(0)
    model = gnncca.MOTMPNet(copy.deepcopy(params), None, "bdnet_market").train()
    spec = mpn_oracle.MPNSpec(params, "bdnet_market")
    x, ei, ea = synthetic_graph(4, 10, 64, seed=11)
    labels = (torch.rand(ei.size(1), generator=torch.Generator().manual_seed(1)) < 0.1).float()
    sd = {k: v.detach().clone().requires_grad_(v.is_floating_point()) for k, v in model.state_dict().items()}
    ref = mpn_oracle.motmpnet_forward(spec, sd, x, ei, ea, training=True)["classified_edges"]
    crit = torch.nn.BCEWithLogitsLoss()
    sum(crit(r.view(-1), labels) for r in ref).backward()
    model = model.to(cuda_device)
    out = model(D(x.to(cuda_device), ei.to(cuda_device), ea.to(cuda_device)))["classified_edges"]
    sum(crit(r.view(-1), labels.to(cuda_device)) for r in out).backward()
    for name, p in model.named_parameters():
        assert close(p.grad, sd[name].grad) <= GRAD_TOL, name


def test_default_tf32_training_step_stays_within_tf32_tolerance_of_the_reference(cuda_device):
    """math='tf32' (the module default): the node encoder of the training forward runs on tcgen05 kind::tf32 and keeps its
    hidden activations; the backward is FP32.  Loss and gradients against the reference's own (fp32) step."""
    TF32_GRAD_TOL = 2e-2
    z = load_golden("mpn_shipped_bdnet_train")
    from test_gpu_mpn import build_model
    model = build_model(z, cuda_device, math="tf32").train()
    data = D(torch.from_numpy(z["x"]).to(cuda_device), torch.from_numpy(z["edge_index"]).to(cuda_device),
             torch.from_numpy(z["edge_attr"]).to(cuda_device))
    labels = torch.from_numpy(z["labels"]).to(cuda_device)
    import gnncca
    n0 = gnncca._lib.load().gnncca_launch_count()
    out = model(data)["classified_edges"]
    crit = torch.nn.BCEWithLogitsLoss()
    loss = sum(crit(r.view(-1), labels) for r in out)
    loss.backward()
    assert abs(loss.item() - float(z["loss"])) <= 1e-3 * (1 + abs(float(z["loss"])))
    for name, p in model.named_parameters():
        ref = torch.from_numpy(z["grad/" + name]).double()
        err = ((p.grad.cpu().double() - ref).abs() / (1 + ref.abs() + 0.01 * ref.abs().max())).max()
        assert float(err) <= TF32_GRAD_TOL, (name, float(err))


@pytest.mark.parametrize("agg,extra", [("sum", None), ("mean", None), ("max", None),
                                       ("sum", {"reattach_initial_nodes": True, "reattach_initial_edges": True})])
def test_one_launch_backward_of_small_frames_matches_the_oracle_and_the_per_step_chain(agg, extra, cuda_device, monkeypatch):
    """Batches from build_graph_batch (closed frames, offsets known) run their message-passing backward as ONE launch
    (gnncca_mp_frames_bwd) with the reverse-edge table instead of the sort by target: loss and every gradient against torch
    autograd through the CPU oracle, and against the per-step kernel chain on the same batch; ragged frames."""
    import copy
    from importlib import import_module
    import gnncca
    from oracle import mpn_oracle
    afn = import_module("gnn-cca_b200.autograd_fn")
    params = import_module("gnn-cca_b200.config").graph_net_params()
    params["node_agg_fn"] = agg
    if extra:
        params.update(extra)
    rng = np.random.default_rng(5)
    cam, ids, ptr = [], [], [0]
    for _ in range(19):
        counts = rng.integers(1, 9, 4)
        cam.append(np.repeat(np.arange(4), counts))
        ids.append(np.concatenate([rng.permutation(8)[:c] for c in counts]))
        ptr.append(ptr[-1] + int(counts.sum()))
    N = ptr[-1]
    g = torch.Generator().manual_seed(3)
    b = gnncca.build_graph_batch(graph_ptr=torch.tensor(ptr, device=cuda_device), id_cam=torch.from_numpy(np.concatenate(cam)).to(cuda_device),
                                 ids=torch.from_numpy(np.concatenate(ids)).to(cuda_device), xw=torch.rand(N, generator=g, dtype=torch.float64).to(cuda_device) * 300,
                                 yw=torch.rand(N, generator=g, dtype=torch.float64).to(cuda_device) * 200,
                                 max_dist=torch.full((19,), 50.83, dtype=torch.float64, device=cuda_device),
                                 reid_embeds=torch.relu(torch.randn(N, 64, generator=g)).to(cuda_device),
                                 node_embeds=torch.relu(torch.randn(N, 512, generator=g)).to(cuda_device))
    assert b.gnncca_trusted
    crit = torch.nn.BCEWithLogitsLoss()
    lib = import_module("gnn-cca_b200._lib").load()
    grads, launches = {}, {}
    for fused in (True, False):
        monkeypatch.setattr(afn, "FUSED_BWD", fused)
        torch.manual_seed(0)
        model = gnncca.MOTMPNet(copy.deepcopy(params), None, "bdnet_market", math="fp32").to(cuda_device).train()
        if hasattr(b, "gnncca_plan"):
            del b.gnncca_plan
        out = model(b)["classified_edges"]
        loss = sum(crit(o.view(-1), b.edge_labels) for o in out)
        n0 = lib.gnncca_launch_count()
        loss.backward()
        launches[fused] = lib.gnncca_launch_count() - n0
        grads[fused] = ({n: p.grad.clone() for n, p in model.named_parameters()}, float(loss.detach()))
        sd0 = {k: v.detach().cpu().clone() for k, v in model.state_dict().items()}
    assert launches[True] + 12 <= launches[False]        # the one-launch path really ran (18 launches of the chain became 1)
    spec = mpn_oracle.MPNSpec(params, "bdnet_market")
    sdr = {k: v.clone().requires_grad_(v.is_floating_point()) for k, v in sd0.items()}
    ref = mpn_oracle.motmpnet_forward(spec, sdr, b.x.cpu(), b.edge_index.cpu(), b.edge_attr.cpu(), training=True)["classified_edges"]
    lref = sum(crit(r.view(-1), b.edge_labels.cpu()) for r in ref)
    lref.backward()
    assert abs(grads[True][1] - float(lref)) <= 1e-5 * (1 + abs(float(lref)))
    for name in grads[True][0]:
        a, c, r = grads[True][0][name], grads[False][0][name], sdr[name].grad
        assert close(a, r) <= GRAD_TOL, name
        assert close(a, c.cpu()) <= 2e-5, name            # the same algebra, other groupings of the sums (and 3xTF32 projections)
