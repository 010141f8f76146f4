"""Two inference forwards of one cross-camera graph through the per-step pipeline (for compute-sanitizer):  python tools/one_forward.py cams dets"""
import copy, importlib, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
gnncca = importlib.import_module("gnn-cca_b200")
mg = importlib.import_module("gnn-cca_b200.multigpu")
engine = importlib.import_module("gnn-cca_b200.engine")
cfgmod = importlib.import_module("gnn-cca_b200.config")
C, Dd = int(sys.argv[1]), int(sys.argv[2])
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = gnncca.MOTMPNet(copy.deepcopy(cfgmod.graph_net_params()), None, "resnet50").to(dev).eval()
N = C * Dd
g = torch.Generator().manual_seed(3)
x = torch.relu(torch.randn(N, 2048, generator=g)); x = (x / x.norm(dim=0).clamp_min(1e-12)).to(dev)
ei = mg.complete_multipartite_slice(C, Dd, (0, N), dev)
ea = torch.rand(ei.size(1), 4, generator=g).to(dev)
with torch.no_grad():
    a = engine.forward_inference(model, x, ei, ea)
    b = engine.forward_inference(model, x, ei, ea)
torch.cuda.synchronize()
print("equal:", [bool(torch.equal(u, v)) for u, v in zip(a, b)])
