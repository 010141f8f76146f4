"""tcgen05 kind::tf32 node encoder (gnncca_node_encoder) against a plain PyTorch fp32 reference of the same op
(models/mlp.py:12-23 applied as in models/mpn.py:128-142)."""
import importlib

import pytest
import torch

pytestmark = pytest.mark.gpu

TF32_TOL = 1e-3   # |got - ref| <= 1e-3 * (1 + |ref|): TF32 operands (10-bit mantissa), fp32 accumulate; ~10x the measured error (1e-4)


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


@pytest.mark.parametrize("rows,in_dim,hidden", [(2560, 2048, 128), (777, 512, 64), (70000, 64, 128)])
def test_training_entry_also_writes_the_hidden_activations(rows, in_dim, hidden, cuda_device):
    """gnncca_node_encoder_train: same output bits as gnncca_node_encoder, plus h1 = ReLU(x W1^T + b1) as layer 2 read it."""
    lib_mod = importlib.import_module("gnn-cca_b200._lib")
    lib = lib_mod.load()
    g = torch.Generator().manual_seed(5)
    x = torch.relu(torch.randn(rows, in_dim, generator=g))
    torch.manual_seed(6)
    l1, l2 = torch.nn.Linear(in_dim, hidden), torch.nn.Linear(hidden, 32)
    with torch.no_grad():
        ref1 = torch.relu(l1(x))
    xd = x.to(cuda_device)
    W1, b1, W2, b2 = (t.detach().to(cuda_device).contiguous() for t in (l1.weight, l1.bias, l2.weight, l2.bias))
    h = torch.full((rows, 32), float("nan"), device=cuda_device)
    h_t = torch.full((rows, 32), float("nan"), device=cuda_device)
    h1 = torch.full((rows, hidden), float("nan"), device=cuda_device)
    st = torch.cuda.current_stream().cuda_stream
    lib_mod.check(lib.gnncca_node_encoder(xd.data_ptr(), W1.data_ptr(), b1.data_ptr(), W2.data_ptr(), b2.data_ptr(), h.data_ptr(), rows,
                                          in_dim, hidden, 32, lib_mod.MATH_TF32, st), "node_encoder")
    lib_mod.check(lib.gnncca_node_encoder_train(xd.data_ptr(), W1.data_ptr(), b1.data_ptr(), W2.data_ptr(), b2.data_ptr(), h1.data_ptr(),
                                                h_t.data_ptr(), rows, in_dim, hidden, 32, st), "node_encoder_train")
    torch.cuda.synchronize()
    assert torch.equal(h, h_t)
    got1 = h1.cpu()
    assert torch.isfinite(got1).all()
    assert ((got1 - ref1).abs() / (1 + ref1.abs())).max().item() <= TF32_TOL
    # layer 2 recomputed in fp32 from the SAVED activations reproduces the kernel's output to TF32 rounding of W2 only
    with torch.no_grad():
        again = torch.relu(l2(got1))
    assert ((h.cpu() - again).abs() / (1 + again.abs())).max().item() <= TF32_TOL


@pytest.mark.parametrize("rows,in_dim,hidden", [(2560, 2048, 128), (2304, 2048, 128), (8192, 2048, 128), (1000, 2048, 128), (130, 512, 64),
                                                 (7, 2048, 128), (24576, 2048, 128)])
def test_split_k_form_for_small_inputs_matches_the_reference_and_the_one_launch_form(rows, in_dim, hidden, cuda_device):
    """gnncca_node_encoder_ws: small inputs contract K in several CTAs per row tile and finish in a second launch.  Against the torch
    fp32 reference at the TF32 tolerance, against the one-launch form at fp32 rounding level (another summation order over K), h1
    included; large inputs (scratch_bytes == 0) give the one-launch form's bits."""
    lib_mod = importlib.import_module("gnn-cca_b200._lib")
    lib = lib_mod.load()
    g = torch.Generator().manual_seed(9)
    x = torch.relu(torch.randn(rows, in_dim, generator=g))
    torch.manual_seed(8)
    l1, l2 = torch.nn.Linear(in_dim, hidden), torch.nn.Linear(hidden, 32)
    with torch.no_grad():
        ref1 = torch.relu(l1(x))
        ref = torch.relu(l2(ref1))
    dev = cuda_device
    xd = x.to(dev)
    W1, b1, W2, b2 = (t.detach().to(dev).contiguous() for t in (l1.weight, l1.bias, l2.weight, l2.bias))
    st = torch.cuda.current_stream().cuda_stream
    splits = int(lib.gnncca_node_encoder_splits(rows, in_dim))
    nbytes = int(lib.gnncca_node_encoder_scratch_bytes(rows, in_dim, hidden))
    assert (nbytes > 0) == (splits > 1) and (splits > 1) == (rows <= 9472)
    scratch = torch.full((max(nbytes, 16) // 4,), float("nan"), device=dev)
    outs = {}
    for name, use in (("ws", True), ("one", False)):
        h = torch.full((rows, 32), float("nan"), device=dev)
        h1 = torch.full((rows, hidden), float("nan"), device=dev)
        lib_mod.check(lib.gnncca_node_encoder_ws(xd.data_ptr(), W1.data_ptr(), b1.data_ptr(), W2.data_ptr(), b2.data_ptr(), h1.data_ptr(), h.data_ptr(),
                                                 rows, in_dim, hidden, 32, scratch.data_ptr() if use else None, nbytes if use else 0, st), "node_encoder_ws")
        torch.cuda.synchronize()
        outs[name] = (h.cpu(), h1.cpu())
    for h, h1 in outs.values():
        assert torch.isfinite(h).all() and torch.isfinite(h1).all()
        assert ((h - ref).abs() / (1 + ref.abs())).max().item() <= TF32_TOL
        assert ((h1 - ref1).abs() / (1 + ref1.abs())).max().item() <= TF32_TOL
    if splits > 1:      # the same products, summed in another order; h1 is rounded to tf32 (2^-11 relative) after the sum
        assert ((outs["ws"][1] - outs["one"][1]).abs() / (1 + outs["one"][1].abs())).max().item() <= 6e-4
        assert ((outs["ws"][0] - outs["one"][0]).abs() / (1 + outs["one"][0].abs())).max().item() <= 6e-4
    else:
        assert torch.equal(outs["ws"][0], outs["one"][0]) and torch.equal(outs["ws"][1], outs["one"][1])
