import copy, importlib, os, sys
import torch
sys.path.insert(0, "/root/repo")
gnncca = importlib.import_module("gnn-cca_b200")
mg = importlib.import_module("gnn-cca_b200.multigpu")
engine = importlib.import_module("gnn-cca_b200.engine")
cfgmod = importlib.import_module("gnn-cca_b200.config")
C, Dd = int(sys.argv[1]), int(sys.argv[2])
dev = torch.device("cuda:0")
params = cfgmod.graph_net_params()
torch.manual_seed(0)
model = gnncca.MOTMPNet(copy.deepcopy(params), None, "resnet50").to(dev).eval()
N = C * Dd
g = torch.Generator().manual_seed(3)
x = torch.relu(torch.randn(N, 2048, generator=g)); x = (x / x.norm(dim=0).clamp_min(1e-12)).to(dev)
ei = mg.complete_multipartite_slice(C, Dd, (0, N), dev)
ea = torch.rand(ei.size(1), 4, generator=g).to(dev)
with torch.no_grad():
    math = engine.MATH_MODES[model.math]
    hs = [engine.mlp_forward_nograd(model.encoder.node_mlp, x, math).clone() for _ in range(6)]
    print("encoder reproducible:", all(torch.equal(hs[0], h) for h in hs), [int((hs[0] != h).sum()) for h in hs])
    outs = [engine.forward_inference(model, x, ei, ea) for _ in range(4)]
    for k in range(3):
        print("full forward logits", k, "reproducible:", [bool(torch.equal(outs[0][k], o[k])) for o in outs], [int((outs[0][k] != o[k]).sum()) for o in outs])
    # one rank holding everything through the partition driver
    gen = mg.edge_partitioned_steps(model, x, ei, ea, 0)
    part = next(gen); res = None
    while res is None:
        try: part = gen.send(part)
        except StopIteration as done: res = done.value
    for k in range(3):
        print("driver(1 rank) vs full", k, bool(torch.equal(res[k], outs[0][k])), int((res[k] != outs[0][k]).sum()))
