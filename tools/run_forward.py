"""A few inference forwards of one synthetic workload (development aid for ncu captures):
    python tools/run_forward.py [cams dets frames [n_forwards]]"""
import copy, importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
gnncca = importlib.import_module("gnn-cca_b200")
cfgmod = importlib.import_module("gnn-cca_b200.config")
syn = importlib.import_module("gnn-cca_b200.synthetic")
cams, dets, frames = (int(v) for v in (sys.argv[1:4] if len(sys.argv) >= 4 else (8, 64, 64)))
n = int(sys.argv[4]) if len(sys.argv) > 4 else 3
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = gnncca.MOTMPNet(copy.deepcopy(cfgmod.graph_net_params()), None, "resnet50").to(dev).eval()
b = gnncca.build_graph_batch(**syn.make_detections(cams, dets, frames, seed=3, device=dev, max_dist=50.83))
with torch.no_grad():
    for i in range(n):
        out = model(b)["classified_edges"]
torch.cuda.synchronize()
a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
with torch.no_grad():
    a.record()
    for i in range(n):
        out = model(b)["classified_edges"]
    z.record()
torch.cuda.synchronize()
print("N=%d E=%d path=%s %.3f ms/forward" % (b.num_nodes, b.num_edges, model.last_plan.last_path, a.elapsed_time(z) / n))
