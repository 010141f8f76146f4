"""Inference forwards with every torch.empty the engine makes poisoned (NaN / 0xFF): an uninitialised read shows up as NaN logits.
    python tools/poison_forward.py cams dets"""
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
real_empty = torch.empty
def poisoned(*a, **k):
    t = real_empty(*a, **k)
    if t.is_cuda:
        if t.dtype.is_floating_point: t.fill_(float("nan"))
        elif t.dtype == torch.uint8: t.fill_(255)
    return t
with torch.no_grad():
    ref = engine.forward_inference(model, x, ei, ea)
    torch.empty = poisoned
    outs = [engine.forward_inference(model, x, ei, ea) for _ in range(3)]
    torch.empty = real_empty
torch.cuda.synchronize()
for k in range(len(ref)):
    print("logits", k, "nan:", [int(torch.isnan(o[k]).sum()) for o in outs], "differ from first run:", [int((o[k] != ref[k]).sum()) for o in outs])
