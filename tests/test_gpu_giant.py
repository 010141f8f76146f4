"""One large cross-camera graph (the shape of BASELINE config 5, scaled to what the CPU oracle evaluates end to end in seconds):
the per-step pipeline, unpartitioned and edge-partitioned, against the float64 chunked oracle -- full-degree segments (1920
out-edges per node) that span dozens of reduction tiles and CTAs."""
import copy
import importlib

import pytest
import torch

from oracle import mpn_oracle, mpn_oracle_chunked as moc

pytestmark = pytest.mark.gpu

engine = importlib.import_module("gnn-cca_b200.engine")
mg = importlib.import_module("gnn-cca_b200.multigpu")
cfgmod = importlib.import_module("gnn-cca_b200.config")


_CACHE = {}


def _inputs(C, D, dev, seed=1239):
    N, deg = C * D, C * D - D
    g = torch.Generator(device=dev).manual_seed(seed)
    x = torch.relu(torch.randn((N, 2048), generator=g, device=dev))
    x = x / x.norm(dim=0).clamp_min(1e-12)
    ei = mg.complete_multipartite_slice(C, D, (0, N), dev)
    eid = torch.arange(0, N * deg, device=dev, dtype=torch.float32)
    ea = torch.stack([torch.frac(eid * 0.61803398875), torch.frac(eid * 0.41421356237), torch.frac(eid * 0.73205080757),
                      torch.frac(eid * 0.23606797750)], dim=1).contiguous()
    return x, ei, ea


@pytest.mark.parametrize("math,tol", [("fp32", 1e-4), ("tf32", 1e-3)])
def test_giant_graph_matches_the_float64_oracle_end_to_end(math, tol, cuda_device):
    import gnncca
    C, D = 16, 128                      # N = 2048, E = 3 932 160, degree 1920
    params = cfgmod.graph_net_params()
    torch.manual_seed(0)
    model = gnncca.MOTMPNet(copy.deepcopy(params), None, "resnet50", math=math).to(cuda_device).eval()
    x, ei, ea = _inputs(C, D, cuda_device)
    N, deg = C * D, C * D - D
    with torch.no_grad():
        full = engine.forward_inference(model, x, ei, ea)
        # the same through 4 virtual ranks in lock step on this GPU (the exchange replaced by a sum of the partial matrices)
        plan = importlib.import_module("gnn-cca_b200.plan").GraphPlan(ei, N).resolve()
        ranges = mg.balanced_source_ranges(plan.rowptr, 4)
        for shard in (False, True):      # replicated encoder / encoder and first projection on the owned rows + one more exchange
            gens = [mg.edge_partitioned_steps(model, x, ei[:, a * deg:b * deg].contiguous(), ea[a * deg:b * deg].contiguous(), a * deg,
                                              owned=(a, b) if shard else None, shard_encoder=shard) for a, b in ranges]
            parts, outs = [next(gn) for gn in gens], [None] * 4
            rounds = 0
            while any(o is None for o in outs):
                h = torch.stack(parts).sum(0)
                rounds += 1
                nxt = []
                for r, gn in enumerate(gens):
                    try:
                        nxt.append(gn.send(h))
                    except StopIteration as done:
                        outs[r] = done.value
                parts = nxt
            assert rounds == (4 if shard else 3)
            for k in range(3):
                assert torch.equal(torch.cat([o[k] for o in outs]), full[k])
    if "ref" not in _CACHE:            # same seed, same inputs for both math modes: one oracle evaluation (~15 s of CPU)
        spec = mpn_oracle.MPNSpec(params, "resnet50")
        sd64 = moc.cast_state_dict({k: v.detach().cpu() for k, v in model.state_dict().items()}, torch.float64)
        with torch.no_grad():
            _CACHE["ref"] = moc.motmpnet_forward_chunked(spec, sd64, x.cpu().double(), ei.cpu(), ea.cpu().double())["classified_edges"]
    ref = _CACHE["ref"]
    for o, r in zip(full, ref):
        err = float(((o.cpu().double() - r).abs() / (1 + r.abs() + 0.01 * r.abs().max())).max())
        assert err <= tol, err
        sure = r.view(-1).abs() > tol * (1 + float(r.abs().max()))
        assert torch.equal(o.cpu().view(-1)[sure] >= 0, r.view(-1)[sure] >= 0)
