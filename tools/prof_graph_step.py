"""cfg-3 training step through GraphedTrainStep (development aid for ncu launch lists): python tools/prof_graph_step.py [replays]"""
import copy, importlib, os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
gnncca = importlib.import_module("gnn-cca_b200")
cfgmod = importlib.import_module("gnn-cca_b200.config")
syn = importlib.import_module("gnn-cca_b200.synthetic")
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = gnncca.MOTMPNet(copy.deepcopy(cfgmod.graph_net_params()), None, "resnet50").to(dev).train()
opt = torch.optim.SGD(model.parameters(), lr=1e-3, momentum=0.9, weight_decay=1e-4, fused=True)
batches = [gnncca.build_graph_batch(**syn.make_detections(4, 10, 64, seed=7 + s, device=dev, max_dist=153.54)) for s in range(4)]
step = gnncca.GraphedTrainStep(model)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 50
step(batches[0]); opt.step()
torch.cuda.synchronize()
print("MARK replay start", flush=True)
t0 = time.perf_counter()
for i in range(n):
    step(batches[i % 4]); opt.step()
torch.cuda.synchronize()
print("%.3f ms/step over %d replays (%d kernels of libgnncca per replay)" % ((time.perf_counter() - t0) / n * 1e3, n, step.launches_replayed // step.replays))
