"""Host-side profile of one cfg-3 training step (development aid): python tools/prof_train.py"""
import copy, cProfile, importlib, io, os, pstats, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
gnncca = importlib.import_module("gnn-cca_b200")
cfgmod = importlib.import_module("gnn-cca_b200.config")
syn = importlib.import_module("gnn-cca_b200.synthetic")
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = gnncca.MOTMPNet(copy.deepcopy(cfgmod.graph_net_params()), None, "resnet50").to(dev).train()
opt = torch.optim.SGD(model.parameters(), lr=1e-3, momentum=0.9, weight_decay=1e-4)
crit = torch.nn.BCEWithLogitsLoss()
batches = [gnncca.build_graph_batch(**syn.make_detections(4, 10, 64, seed=7 + s, device=dev)) for s in range(4)]

def step(b):
    opt.zero_grad(set_to_none=True)
    out = model(b)["classified_edges"]
    loss = gnncca.loss_stats(out, b.edge_labels).loss
    loss.backward()
    opt.step()

import sys
NW = int(sys.argv[1]) if len(sys.argv) > 1 else 8
for i in range(NW):
    step(batches[i % 4])
torch.cuda.synchronize()
if len(sys.argv) > 2 and sys.argv[2] == "only":      # for an ncu launch list: NW steps and out
    sys.exit(0)
t0 = time.perf_counter()
for i in range(40):
    step(batches[i % 4])
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print("host %.3f ms/step, with final sync %.3f ms/step" % ((t1 - t0) / 40 * 1e3, (t2 - t0) / 40 * 1e3))
pr = cProfile.Profile(); pr.enable()
for i in range(40):
    step(batches[i % 4])
pr.disable(); torch.cuda.synchronize()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(28); print(s.getvalue()[:6000])
