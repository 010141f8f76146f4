"""Edge-partitioned forward with W virtual ranks on ONE GPU against the unpartitioned forward (bit equality), reporting where they differ:
    python tools/partition_check.py cams dets world [agg]"""
import copy, importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
gnncca = importlib.import_module("gnn-cca_b200")
mg = importlib.import_module("gnn-cca_b200.multigpu")
engine = importlib.import_module("gnn-cca_b200.engine")
cfgmod = importlib.import_module("gnn-cca_b200.config")
C, Dd, world = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
agg = sys.argv[4] if len(sys.argv) > 4 else "sum"
dev = torch.device("cuda:0")
params = cfgmod.graph_net_params(); params["node_agg_fn"] = agg
torch.manual_seed(0)
model = gnncca.MOTMPNet(copy.deepcopy(params), None, "resnet50").to(dev).eval()
N = C * Dd
g = torch.Generator().manual_seed(3)
x = torch.relu(torch.randn(N, 2048, generator=g)); x = (x / x.norm(dim=0).clamp_min(1e-12)).to(dev)
ei = mg.complete_multipartite_slice(C, Dd, (0, N), dev)
ea = torch.rand(ei.size(1), 4, generator=g).to(dev)
with torch.no_grad():
    full = engine.forward_inference(model, x, ei, ea)
    plan = importlib.import_module("gnn-cca_b200.plan").GraphPlan(ei, N).resolve()
    ranges = mg.balanced_source_ranges(plan.rowptr, world)
    bounds = [(int(plan.rowptr[a]), int(plan.rowptr[b])) for a, b in ranges]
    gens = [mg.edge_partitioned_steps(model, x, ei[:, e0:e1].contiguous(), ea[e0:e1].contiguous(), e0) for e0, e1 in bounds]
    parts = [next(gn) for gn in gens]
    outs = [None] * world
    step = 0
    while any(o is None for o in outs):
        h = torch.stack(parts).sum(0)
        step += 1
        nxt = []
        for r, gn in enumerate(gens):
            try:
                nxt.append(gn.send(h))
            except StopIteration as done:
                outs[r] = done.value
        parts = nxt
    bad = 0
    for k in range(len(full)):
        got = torch.cat([o[k] for o in outs])
        if not torch.equal(got, full[k]):
            d = (got != full[k]).view(-1).nonzero().view(-1)
            bad += 1
            print("logits %d differ at %d edges, first %s, tile %s chunk-in-tile %s, max |d| %.3e" % (
                k, d.numel(), d[:4].tolist(), (d[:4] // 512).tolist(), ((d[:4] % 512) // 64).tolist(), float((got - full[k]).abs().max())))
print("bounds", bounds[:3], "phases", [b[0] % 512 for b in bounds])
print("PARTITION CHECK", "OK" if bad == 0 else "FAILED")
