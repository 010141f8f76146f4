"""Multi-GPU paths on the GPU: frame shards reproduce the single-GPU result bit for bit; the edge-partitioned giant
graph path (virtual ranks driven in lock step on one GPU, the exchange replaced by a sum of the partial matrices)
reproduces the unpartitioned result bit for bit; and, when at least two GPUs are visible, the same through NCCL."""
import copy
import importlib
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _model(dev, agg="sum"):
    import gnncca
    params = importlib.import_module("gnn-cca_b200.config").graph_net_params()
    params["node_agg_fn"] = agg
    torch.manual_seed(0)
    return gnncca.MOTMPNet(copy.deepcopy(params), None, "bdnet_market", math="fp32").to(dev).eval()


def test_frame_shards_equal_the_single_gpu_result(cuda_device):
    import gnncca
    mg = importlib.import_module("gnn-cca_b200.multigpu")
    syn = importlib.import_module("gnn-cca_b200.synthetic")
    model = _model(cuda_device)
    batch = gnncca.build_graph_batch(**syn.make_detections(4, 6, 37, seed=5, device=cuda_device, node_dim=512))
    with torch.no_grad():
        full = model(batch)["classified_edges"]
        for world in (2, 3, 8):
            parts = [model(mg.shard_batch(batch, r, world))["classified_edges"] for r in range(world)]
            for k in range(3):
                assert torch.equal(torch.cat([p[k] for p in parts]), full[k])


@pytest.mark.parametrize("shape", [(6, 40), (4, 37), (16, 64)])
@pytest.mark.parametrize("agg", ["sum", "max"])
def test_edge_partitioned_giant_graph_equals_unpartitioned(agg, shape, cuda_device):
    """(6, 40): 240 nodes, degree 200: segments span tiles and rank boundaries fall mid-tile, on 16-byte aligned edge offsets (the
    chunks of a partition arrive as bulk copies).  (4, 37): degree 111, odd edge offsets: every chunk of a partition takes the
    register path of the step kernel and the pair stores are unaligned.  (16, 64): degree 960, ~1 M edges, several tiles per CTA."""
    mg = importlib.import_module("gnn-cca_b200.multigpu")
    engine = importlib.import_module("gnn-cca_b200.engine")
    model = _model(cuda_device, agg)
    C, Dd = shape
    N = C * Dd
    g = torch.Generator().manual_seed(3)
    x = torch.relu(torch.randn(N, 512, generator=g))
    x = (x / x.norm(dim=0).clamp_min(1e-12)).to(cuda_device)
    ei = mg.complete_multipartite_slice(C, Dd, (0, N), cuda_device)
    ea = torch.rand(ei.size(1), 4, generator=g).to(cuda_device)
    with torch.no_grad():
        full = engine.forward_inference(model, x, ei, ea)
        plan = importlib.import_module("gnn-cca_b200.plan").GraphPlan(ei, N).resolve()
        for world in (2, 4):
            ranges = mg.balanced_source_ranges(plan.rowptr, world)
            bounds = [(int(plan.rowptr[a]), int(plan.rowptr[b])) for a, b in ranges]
            gens = [mg.edge_partitioned_steps(model, x, ei[:, e0:e1].contiguous(), ea[e0:e1].contiguous(), e0) for e0, e1 in bounds]
            parts = [next(gn) for gn in gens]
            outs = [None] * world
            while any(o is None for o in outs):
                for r, (a, b) in enumerate(ranges):           # a rank only ever writes the rows it owns
                    mask = torch.ones(N, dtype=torch.bool, device=cuda_device); mask[a:b] = False
                    assert float(parts[r][mask].abs().sum()) == 0.0
                h = torch.stack(parts).sum(0)
                nxt = []
                for r, gn in enumerate(gens):
                    try:
                        nxt.append(gn.send(h))
                    except StopIteration as done:
                        outs[r] = done.value
                parts = nxt
            for k in range(3):
                assert torch.equal(torch.cat([o[k] for o in outs]), full[k])


def test_edge_partitioned_over_nccl_when_two_gpus_are_visible():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29631", os.path.join(ROOT, "tools", "giant_graph.py"), "--cams", "6", "--dets", "40", "--check"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "CHECK OK" in r.stdout
