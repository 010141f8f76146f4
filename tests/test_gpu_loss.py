"""GPU parity of the fused loss + statistics kernels (gnncca_bce_stats_fwd/bwd) against vectors made
by running the reference's own compute_loss_acc (tests/golden/loss_*.npz) and against the CPU oracle at larger sizes."""
import types

import numpy as np
import pytest
import torch

from conftest import golden_cases, load_golden
from oracle import loss_oracle

pytestmark = pytest.mark.gpu

LOSS_RTOL = 2e-6      # fp32 means over E samples in a different (fixed) summation order; accumulators here are wider
GRAD_RTOL = 2e-6


class Focal(torch.nn.Module):
    """Stand-in with the attributes of libs/utils.py:264-293 FocalLoss_binary (the criterion object is only inspected)."""

    def __init__(self, focusing_param=5, balance_param=0.9, reduction="mean"):
        super().__init__()
        self.focusing_param, self.balance_param = focusing_param, balance_param
        self.cross_entropy = torch.nn.BCEWithLogitsLoss(reduction=reduction)


def criteria(name, pw, dev):
    if name == "Focal":
        return Focal(reduction="mean"), Focal(reduction="none")
    if name == "BCE_weighted":
        t = torch.tensor(pw, device=dev)
        return torch.nn.BCEWithLogitsLoss(reduction="mean", pos_weight=t), torch.nn.BCEWithLogitsLoss(reduction="none", pos_weight=t)
    if name == "BCE":
        return torch.nn.BCEWithLogitsLoss(reduction="mean"), torch.nn.BCEWithLogitsLoss(reduction="none")
    return "", ""


def rel(got, ref):
    got, ref = np.asarray(got, np.float64), np.asarray(ref, np.float64)
    if np.isnan(ref).any():
        assert np.array_equal(np.isnan(got), np.isnan(ref))
        return 0.0
    return float(np.max(np.abs(got - ref) / (1e-7 + np.abs(ref)))) if ref.size else 0.0


@pytest.mark.parametrize("case", golden_cases("loss_"))
def test_compute_loss_acc_matches_the_reference_function(case, cuda_device):
    import gnncca
    z = load_golden("loss_" + case)
    k = int(z["k"])
    logits = [torch.from_numpy(z["logits%d" % s]).to(cuda_device).requires_grad_(True) for s in range(k)]
    batch = types.SimpleNamespace(edge_labels=torch.from_numpy(z["labels"]).to(cuda_device))
    pw = None if float(z["pos_weight"]) < 0 else float(z["pos_weight"])
    c, cn = criteria(str(z["criterion"]), pw, cuda_device)
    n0 = gnncca._lib.load().gnncca_launch_count()
    loss, p1, p0, pa, l1, l0, probs = gnncca.compute_loss_acc({"classified_edges": logits}, batch, c, cn, str(z["mode"]))
    assert gnncca._lib.load().gnncca_launch_count() - n0 == 2
    assert loss.dtype == torch.float32 and loss.is_cuda and loss.requires_grad
    assert rel(loss.item(), z["loss"]) <= LOSS_RTOL
    assert rel(l1.item(), z["loss_class1"]) <= LOSS_RTOL and rel(l0.item(), z["loss_class0"]) <= LOSS_RTOL
    # counts are integers: the percentages are exactly the reference's
    assert (p1, p0, pa) == ([float(z["precision_class1"])], [float(z["precision_class0"])], [float(z["precision_all"])])
    assert len(probs) == k
    for s in range(k):
        np.testing.assert_allclose(probs[s].cpu().numpy(), z["probs%d" % s].reshape(-1), rtol=2e-6, atol=1e-7)
    n0 = gnncca._lib.load().gnncca_launch_count()
    loss.backward()
    assert gnncca._lib.load().gnncca_launch_count() - n0 == 1
    for s in range(k):
        g, ref = logits[s].grad.cpu().numpy(), z["grad%d" % s]
        assert g.shape == ref.shape
        assert np.max(np.abs(g - ref)) <= GRAD_RTOL * (np.abs(ref).max() + 1e-12), (case, s)


@pytest.mark.parametrize("name,pw", [("BCE", None), ("BCE_weighted", 3.3), ("Focal", None), ("", None)])
def test_large_batch_against_the_cpu_oracle(name, pw, cuda_device):
    """cfg-2 sized edge count (442 368 edges, 3 classified steps): many blocks, grid-stride tail."""
    import gnncca
    g = torch.Generator().manual_seed(11)
    E, k = 442368 + 5, 3
    labels = (torch.rand(E, generator=g) < 0.18).float()
    logits = [(torch.randn(E, 1, generator=g) * 3 + (labels.view(-1, 1) * 2 - 1) * s) for s in range(k)]
    mode = "train" if name else "validate"
    ref_in = [x.clone().requires_grad_(True) for x in logits]
    ref = loss_oracle.compute_loss_acc(ref_in, labels, name or "BCE", pw, mode)
    ref_g = torch.autograd.grad(ref["loss"], ref_in)
    dev_in = [x.to(cuda_device).requires_grad_(True) for x in logits]
    c, cn = criteria(name, pw, cuda_device)
    st_loss, p1, p0, pa, l1, l0, probs = gnncca.compute_loss_acc({"classified_edges": dev_in}, types.SimpleNamespace(edge_labels=labels.to(cuda_device)),
                                                                 c, cn, mode)
    # the CPU reference sums E fp32 terms in fp32: allow its own rounding (1e-5), counts stay exact
    assert rel(st_loss.item(), ref["loss"].item()) <= 1e-5
    assert rel(l1.item(), ref["loss_class1"].item()) <= 1e-5 and rel(l0.item(), ref["loss_class0"].item()) <= 1e-5
    assert (p1[0], p0[0], pa[0]) == (ref["precision_class1"], ref["precision_class0"], ref["precision_all"])
    st_loss.backward()
    for s in range(k):
        a, b = dev_in[s].grad.cpu(), ref_g[s]
        assert float((a - b).abs().max()) <= 1e-5 * float(b.abs().max())
        np.testing.assert_allclose(probs[s].cpu().numpy(), ref["probs"][s].numpy(), rtol=2e-6, atol=1e-7)


def test_statistics_are_bit_reproducible_and_stay_on_the_device(cuda_device):
    import gnncca
    g = torch.Generator().manual_seed(3)
    E = 100003
    labels = (torch.rand(E, generator=g) < 0.3).float().to(cuda_device)
    logits = [torch.randn(E, generator=g).to(cuda_device) * 4 for _ in range(3)]
    a = gnncca.loss_stats(logits, labels, pos_weight=2.0)
    b = gnncca.loss_stats(logits, labels, pos_weight=2.0)
    assert torch.equal(a.stats, b.stats) and a.probs is None
    # per-class mean probabilities (train.py:461-470) against torch
    for s in range(3):
        p = torch.sigmoid(logits[s])
        assert abs(a.mean_prob_class1[s].item() - p[labels == 1].double().mean().item()) <= 1e-6
        assert abs(a.mean_prob_class0[s].item() - p[labels == 0].double().mean().item()) <= 1e-6
    n1, n0, c1, c0 = a.counts()
    assert n1 == int((labels == 1).sum()) and n0 == int((labels == 0).sum())
    pred = (torch.sigmoid(logits[-1]) >= 0.5).float()
    assert c1 == int(((pred == labels) & (labels == 1)).sum()) and c0 == int(((pred == labels) & (labels == 0)).sum())


def test_unsupported_criteria_and_shapes_are_refused(cuda_device):
    import gnncca
    labels = torch.zeros(10, device=cuda_device)
    x = [torch.zeros(10, 1, device=cuda_device)]
    batch = types.SimpleNamespace(edge_labels=labels)
    with pytest.raises(NotImplementedError):
        gnncca.compute_loss_acc({"classified_edges": x}, batch, torch.nn.MSELoss(), torch.nn.MSELoss(reduction="none"), "train")
    with pytest.raises(NotImplementedError):
        gnncca.compute_loss_acc({"classified_edges": x}, batch, torch.nn.BCEWithLogitsLoss(), torch.nn.BCEWithLogitsLoss(), "train")
    with pytest.raises(ValueError):
        gnncca.compute_loss_acc({"classified_edges": x}, batch, "", "", "train")
    with pytest.raises(ValueError):
        gnncca.loss_stats([torch.zeros(9, device=cuda_device)], labels)
    with pytest.raises(NotImplementedError):
        gnncca.loss_stats([labels] * 9, labels)
    with pytest.raises(RuntimeError):
        gnncca.loss_stats([torch.zeros(10)], torch.zeros(10))


@pytest.mark.parametrize("case", ["shipped_bdnet_train", "reattach_train"])
def test_training_step_through_the_fused_loss_matches_reference_gradients(case, cuda_device):
    """MPN forward (train) -> compute_loss_acc -> backward: parameter gradients of the reference's own step."""
    import gnncca
    from test_gpu_mpn import D, build_model
    from test_gpu_train import GRAD_TOL
    z = load_golden("mpn_" + case)
    model = build_model(z, cuda_device).train()
    data = D(torch.from_numpy(z["x"]).to(cuda_device), torch.from_numpy(z["edge_index"]).to(cuda_device),
             torch.from_numpy(z["edge_attr"]).to(cuda_device))
    data.edge_labels = torch.from_numpy(z["labels"]).to(cuda_device)
    out = model(data)
    c, cn = criteria("BCE", None, cuda_device)
    loss = gnncca.compute_loss_acc(out, data, c, cn, "train")[0]
    loss.backward()
    assert abs(loss.item() - float(z["loss"])) <= 1e-5 * (1 + abs(float(z["loss"])))
    for name, p in model.named_parameters():
        ref = torch.from_numpy(z["grad/" + name])
        err = ((p.grad.cpu().double() - ref.double()).abs() / (1 + ref.abs().double() + 0.01 * ref.abs().max().double())).max()
        assert float(err) <= GRAD_TOL, (name, float(err))
