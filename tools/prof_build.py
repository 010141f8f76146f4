"""Host + device profile of build_graph_batch at cfg 2 (development aid)."""
import cProfile, importlib, io, os, pstats, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
gnncca = importlib.import_module("gnn-cca_b200")
syn = importlib.import_module("gnn-cca_b200.synthetic")
dev = torch.device("cuda:0")
cams, dets, frames = (int(v) for v in (sys.argv[1:4] if len(sys.argv) >= 4 else (4, 6, 1024)))
det = syn.make_detections(cams, dets, frames, seed=3, device=dev, max_dist=44.23)
for _ in range(3):
    b = gnncca.build_graph_batch(**det)
torch.cuda.synchronize()
if len(sys.argv) > 4 and sys.argv[4] == "only":
    sys.exit(0)
t0 = time.perf_counter()
for _ in range(20):
    b = gnncca.build_graph_batch(**det)
torch.cuda.synchronize()
print("N=%d E=%d: %.3f ms per build" % (b.num_nodes, b.num_edges, (time.perf_counter() - t0) / 20 * 1e3))
pr = cProfile.Profile(); pr.enable()
for _ in range(20):
    b = gnncca.build_graph_batch(**det)
pr.disable(); torch.cuda.synchronize()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(18); print(s.getvalue()[:3500])
