"""Per-stage device times of ONE rank's share of the cfg-5 partitioned forward, run alone on one GPU (no exchange partner: the
assembled Q is this rank's partial, so the numbers are timings only).  python tools/partition_stages.py [world rank]"""
import copy, importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
gnncca = importlib.import_module("gnn-cca_b200"); engine = importlib.import_module("gnn-cca_b200.engine")
mg = importlib.import_module("gnn-cca_b200.multigpu"); cfgmod = importlib.import_module("gnn-cca_b200.config")
world, rank = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (8, 3)
dev = torch.device("cuda:0")
wl = bench.WORKLOADS["cfg5"]
N, deg = wl["cams"] * wl["dets"], wl["cams"] * wl["dets"] - wl["dets"]
torch.manual_seed(0)
model = gnncca.MOTMPNet(copy.deepcopy(cfgmod.graph_net_params()), None, "resnet50").to(dev).eval()
ranges = mg.balanced_source_ranges(torch.arange(0, N * deg + 1, deg, dtype=torch.int64), world)
n0, n1 = ranges[rank]
x, ei, ea = bench.giant_inputs(wl, (n0, n1), dev)
with torch.no_grad():
    profs = []
    for i in range(8):
        pr = engine.StageTimer()
        mg.forward_edge_partitioned(model, x, ei, ea, ranges, edge_base=n0 * deg, prof=pr)
        pr.mark("end"); profs.append(pr)
    torch.cuda.synchronize()
acc = {}
for pr in profs[4:]:
    for nm, tt in pr.stages():
        acc.setdefault(nm, []).append(tt)
tot = 0
for k, v in acc.items():
    if k not in ("begin", "end"):
        print("%-14s %8.1f us" % (k, 1e3 * sum(v) / len(v))); tot += sum(v) / len(v)
print("sum %.1f us for E_local = %d (world %d rank %d)" % (tot * 1e3, ei.size(1), world, rank))
