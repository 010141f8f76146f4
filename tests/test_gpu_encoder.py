"""tcgen05 kind::tf32 node encoder (gnncca_node_encoder) against a plain PyTorch fp32 reference of the same op
(models/mlp.py:12-23 applied as in models/mpn.py:128-142)."""
import importlib

import pytest
import torch

pytestmark = pytest.mark.gpu

TF32_TOL = 5e-3   # |got - ref| <= 5e-3 * (1 + |ref|): TF32 operands (10-bit mantissa), fp32 accumulate


def _run(rows, in_dim, hidden, dev, seed=0, scale=1.0, keep=None):
    lib_mod = importlib.import_module("gnn-cca_b200._lib")
    lib = lib_mod.load()
    g = torch.Generator().manual_seed(seed)
    x = torch.relu(torch.randn(rows, in_dim, generator=g)) * scale
    if keep is not None:
        x, rows = x[:keep].contiguous(), keep
    l1, l2 = torch.nn.Linear(in_dim, hidden), torch.nn.Linear(hidden, 32)
    with torch.no_grad():
        ref = torch.relu(l2(torch.relu(l1(x))))
    xd = x.to(dev)
    W1, b1, W2, b2 = (t.detach().to(dev).contiguous() for t in (l1.weight, l1.bias, l2.weight, l2.bias))
    h = torch.full((rows, 32), float("nan"), device=dev)
    rc = lib.gnncca_node_encoder(xd.data_ptr(), W1.data_ptr(), b1.data_ptr(), W2.data_ptr(), b2.data_ptr(), h.data_ptr(), rows,
                                 in_dim, hidden, 32, lib_mod.MATH_TF32, torch.cuda.current_stream().cuda_stream)
    lib_mod.check(rc, "node_encoder")
    torch.cuda.synchronize()
    return h.cpu(), ref


@pytest.mark.parametrize("rows,in_dim,hidden", [(128, 2048, 128), (24576, 2048, 128), (1, 512, 128), (777, 512, 128),
                                                 (300, 2048, 64), (5000, 256, 64), (129, 32, 128), (60000, 64, 128), (148 * 256 + 5, 32, 128),
                                                 (148 * 512 + 136, 32, 64)])
def test_node_encoder_tf32_matches_fp32_reference(rows, in_dim, hidden, cuda_device):
    torch.manual_seed(3)
    got, ref = _run(rows, in_dim, hidden, cuda_device)
    assert torch.isfinite(got).all()
    err = ((got - ref).abs() / (1 + ref.abs())).max().item()
    assert err <= TF32_TOL, err
    # tighter, scale-aware statement: error relative to the largest activation stays at TF32 rounding level
    assert (got - ref).abs().max().item() <= 2e-3 * max(1.0, ref.abs().max().item())


def test_node_encoder_tf32_is_deterministic_and_row_local(cuda_device):
    torch.manual_seed(4)
    a, _ = _run(1000, 2048, 128, cuda_device, seed=7)
    torch.manual_seed(4)     # same layer initialisation
    b, _ = _run(1000, 2048, 128, cuda_device, seed=7)
    assert torch.equal(a, b)
    torch.manual_seed(4)
    c, _ = _run(1000, 2048, 128, cuda_device, seed=7, keep=333)    # a shorter batch of the same rows
    assert torch.equal(a[:333], c)


def test_node_encoder_rejects_hidden_widths_without_a_kernel(cuda_device):
    """hidden = 256 does not fit one CTA's tensor memory next to the layer-2 accumulators: UNSUPPORTED, never a fallback."""
    with pytest.raises(NotImplementedError):
        _run(64, 64, 256, cuda_device)
