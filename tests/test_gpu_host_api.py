"""The host-buffer entry points (``MOTMPNet.forward_host`` and ``HostPipeline``): same logits as the device call and as the
reference's vectors, also when batches of different shapes alternate through the pipeline's slots."""
import pytest
import torch

from conftest import golden_cases, load_golden
from test_gpu_mpn import FP32_RTOL, D, build_model, close, frame_ptr

pytestmark = pytest.mark.gpu

CASES = [c for c in golden_cases("mpn_") if c.endswith("_eval")]


def host_inputs(z):
    return tuple(torch.from_numpy(z[k]).pin_memory() for k in ("x", "edge_index", "edge_attr"))


@pytest.mark.parametrize("case", CASES)
def test_forward_host_matches_reference_vectors(case, cuda_device):
    z = load_golden("mpn_" + case)
    model = build_model(z, cuda_device).eval()
    with torch.no_grad():
        out = model.forward_host(*host_inputs(z))
    assert len(out) == int(z["n_steps"])
    for i, o in enumerate(out):
        ref = torch.from_numpy(z["logits/%d" % i])
        assert not o.is_cuda and o.shape == ref.shape
        assert close(o, ref) <= FP32_RTOL, (case, i, close(o, ref))


def test_forward_host_rejects_device_tensors(cuda_device):
    z = load_golden("mpn_" + CASES[0])
    model = build_model(z, cuda_device).eval()
    x, ei, ea = host_inputs(z)
    with pytest.raises(ValueError):
        model.forward_host(x.to(cuda_device), ei, ea)


def test_pipeline_equals_device_forward_bitwise(cuda_device):
    """Batches of two shapes pushed through the two-slot pipeline, several rounds: every result equals the device call's
    bits (same kernels, same order), whatever is still in flight in the other slot."""
    import gnncca
    zs = [load_golden("mpn_" + c) for c in CASES]
    by_cfg = {}
    for z in zs:
        by_cfg.setdefault((str(z["config"]), str(z["arch"])), []).append(z)
    group = max(by_cfg.values(), key=len)           # cases sharing one architecture (different graphs)
    z0 = group[0]
    model = build_model(z0, cuda_device).eval()
    want = []
    with torch.no_grad():
        for z in group:
            data = D(*(torch.from_numpy(z[k]).to(cuda_device) for k in ("x", "edge_index", "edge_attr")))
            want.append([o.cpu() for o in model(data)["classified_edges"]])
        pipe = gnncca.HostPipeline(model)
        hosts = [host_inputs(z) for z in group]
        pending = []
        for rnd in range(3):
            for i, h in enumerate(hosts):
                pending.append((i, pipe.submit(*h)))
                if len(pending) > 1:
                    j, hd = pending.pop(0)
                    got = hd.wait()
                    for a, b in zip(got, want[j]):
                        assert torch.equal(a, b), (rnd, j)
        for j, hd in pending:
            for a, b in zip(hd.wait(), want[j]):
                assert torch.equal(a, b)


def test_pipeline_with_frame_offsets(cuda_device):
    """The trusted-frames route (graph_ptr given) through the pipeline."""
    import gnncca
    z = load_golden("mpn_" + CASES[0])
    model = build_model(z, cuda_device).eval()
    x, ei, ea = host_inputs(z)
    ptr = frame_ptr(ei, x.shape[0]).pin_memory()
    with torch.no_grad():
        pipe = gnncca.HostPipeline(model)
        a = pipe.submit(x, ei, ea, graph_ptr=ptr)
        b = pipe.submit(x, ei, ea, graph_ptr=ptr)
        c = pipe.submit(x, ei, ea, graph_ptr=ptr)
        outs = [h.wait() for h in (a, b, c)]
    for i in range(int(z["n_steps"])):
        ref = torch.from_numpy(z["logits/%d" % i])
        for o in outs:
            assert close(o[i], ref) <= FP32_RTOL
        assert torch.equal(outs[0][i], outs[2][i])
