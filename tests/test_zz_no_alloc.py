"""``gnncca_colnorm_static``: the column normalisation with its scratch in a file-scope device array (no allocation inside the
call, include/gnncca.h).  Collected last on purpose: a failure here hides nothing.

The checks go through the C ABI: the existing ``gnncca_colnorm`` (caller scratch) is the comparison, bit for bit -- both run the
same launches -- on two calls in a row with DIFFERENT shapes, so that the second call finds the first one's partial sums in
the shared array; the float64 definition of ``F.normalize(x, p=2, dim=0)`` bounds both at the 1e-6 the graph tests use."""
import importlib

import pytest
import torch

pytestmark = pytest.mark.gpu

_lib = importlib.import_module("gnn-cca_b200._lib")


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _existing(lib, emb):
    rows, cols = emb.shape
    scratch = torch.empty(int(lib.gnncca_colnorm_scratch_bytes(rows, cols)) // 4 + 1, dtype=torch.float32, device=emb.device)
    out = torch.empty_like(emb)
    _lib.check(lib.gnncca_colnorm(_lib.ptr(emb), _lib.ptr(out), rows, cols, _lib.ptr(scratch), _stream()), "colnorm")
    return out


def test_static_scratch_colnorm_equals_the_caller_scratch_entry_on_two_calls_in_a_row(cuda_device):
    lib = _lib.load()
    g = torch.Generator(device=cuda_device).manual_seed(11)
    # large first (fills 41 x 2048 words of the array), then smaller / ragged / unaligned-column shapes that reuse its front
    shapes = [(5120, 2048), (777, 256), (130, 6), (5120, 2048), (1, 512)]
    embs = [torch.relu(torch.randn(s, generator=g, device=cuda_device)) for s in shapes]
    embs[2][:, 3] = 0                                           # a zero column: the 1e-12 clamp
    want = [_existing(lib, e) for e in embs]
    outs = [torch.full_like(e, float("nan")) for e in embs]     # every output buffer exists before the calls under test
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats(cuda_device)["allocation.all.allocated"]
    for e, o in zip(embs, outs):                                # back to back on one stream: ordered, no host sync between
        rc = lib.gnncca_colnorm_static(_lib.ptr(e), _lib.ptr(o), e.size(0), e.size(1), _stream())
        assert rc == _lib.OK, lib.gnncca_last_cuda_error().decode()
    torch.cuda.synchronize()
    assert torch.cuda.memory_stats(cuda_device)["allocation.all.allocated"] == before      # torch's allocator was not touched
    for e, o, w in zip(embs, outs, want):
        assert torch.equal(o, w)
        ref = e.double() / e.double().norm(dim=0).clamp_min(1e-12)
        assert float((o.double() - ref).abs().max()) <= 1e-6
    # in place, like the existing entry allows
    e = embs[1].clone()
    assert lib.gnncca_colnorm_static(_lib.ptr(e), _lib.ptr(e), e.size(0), e.size(1), _stream()) == _lib.OK
    assert torch.equal(e, want[1])


def test_static_scratch_colnorm_rejects_sizes_beyond_the_array(cuda_device):
    lib = _lib.load()
    cap = 1 << 20                                               # GNNCCA_COLNORM_STATIC_MAX_FLOATS
    cols = 2048
    rows_ok = (cap // cols - 1) * 128                           # (ceil(rows / 128) + 1) * cols == cap
    assert lib.gnncca_colnorm_scratch_bytes(rows_ok, cols) == 4 * cap
    assert lib.gnncca_colnorm_scratch_bytes(rows_ok + 1, cols) > 4 * cap
    # the size check comes before anything touches the pointers: one byte of storage stands in for the [rows, cols] matrices
    dummy = torch.zeros(4, dtype=torch.float32, device=cuda_device)
    rc = lib.gnncca_colnorm_static(_lib.ptr(dummy), _lib.ptr(dummy), rows_ok + 1, cols, _stream())
    assert rc == _lib.ERR_WORKSPACE
    msg = lib.gnncca_last_cuda_error().decode()
    assert "gnncca_colnorm_static" in msg and str(cap) in msg
    with pytest.raises(_lib.GnnccaError):
        _lib.check(rc, "colnorm_static")
    torch.cuda.synchronize()
    # the largest accepted shape runs (512 MiB in, 512 MiB out) and agrees with the existing entry
    emb = torch.rand((rows_ok, cols), device=cuda_device)
    out = torch.empty_like(emb)
    assert lib.gnncca_colnorm_static(_lib.ptr(emb), _lib.ptr(out), rows_ok, cols, _stream()) == _lib.OK
    assert torch.equal(out, _existing(lib, emb))
    assert lib.gnncca_colnorm_static(None, None, 0, cols, _stream()) == _lib.OK            # empty input: nothing to do
    assert lib.gnncca_colnorm_static(None, _lib.ptr(out), 5, cols, _stream()) == _lib.ERR_BAD_ARG
