"""Step-kernel micro-benchmark: gnncca_mp_step_fwd (+ fold) launched back to back on a cfg-4-shaped block, CUDA events around the
launches.  python tools/stepbench.py [cams dets frames]   (development aid; bench.py is the number of record)"""
import copy, ctypes as C, importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
gnncca = importlib.import_module("gnn-cca_b200"); engine = importlib.import_module("gnn-cca_b200.engine")
cfgmod = importlib.import_module("gnn-cca_b200.config"); syn = importlib.import_module("gnn-cca_b200.synthetic")
_lib = gnncca._lib
cams, dets, frames = (int(v) for v in (sys.argv[1:4] if len(sys.argv) >= 4 else (8, 64, 64)))
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = gnncca.MOTMPNet(copy.deepcopy(cfgmod.graph_net_params()), None, "resnet50").to(dev).eval()
b = gnncca.build_graph_batch(**syn.make_detections(cams, dets, frames, seed=3, device=dev, max_dist=50.83))
N, E = b.num_nodes, b.num_edges
lib = _lib.load()
plan = gnncca.GraphPlan(b.edge_index, N).resolve()
sb = engine.StepBinding(model, _lib.MATH_TF32)
h = torch.rand(N, 32, device=dev)
P, Q, R = torch.empty(N, 8, device=dev), torch.empty(N, 8, device=dev), torch.empty(N, 32, device=dev)
st = torch.cuda.current_stream().cuda_stream
_lib.check(lib.gnncca_node_project(C.byref(sb.cfg), C.byref(sb.w), 0, h.data_ptr(), N, P.data_ptr(), Q.data_ptr(), R.data_ptr(), st), "proj")
e = torch.rand(E, 6, device=dev); e2 = torch.empty_like(e)
h_out = torch.empty(N, 32, device=dev); lg = torch.empty(E, 1, device=dev)
wsb = int(lib.gnncca_mp_step_workspace_bytes(C.byref(sb.cfg), N, E)); ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
le = model.encoder.edge_mlp.linear(0)
def full():   # steps 2, 3
    _lib.check(lib.gnncca_mp_step_fwd(C.byref(sb.cfg), C.byref(sb.w), plan.rowptr.data_ptr(), plan.src.data_ptr(), plan.dst.data_ptr(), N, E, P.data_ptr(), Q.data_ptr(),
                                      R.data_ptr(), 0, e.data_ptr(), e2.data_ptr(), h_out.data_ptr(), lg.data_ptr(), 0, ws.data_ptr(), wsb, st), "step")
def full_next():   # steps 2, 3 with the next step's projections in the fold launch
    _lib.check(lib.gnncca_mp_step_fwd_next(C.byref(sb.cfg), C.byref(sb.w), plan.rowptr.data_ptr(), plan.src.data_ptr(), plan.dst.data_ptr(), N, E, P.data_ptr(), Q.data_ptr(),
                                           R.data_ptr(), 0, 4, 0, 0, 0, e.data_ptr(), e2.data_ptr(), h_out.data_ptr(), lg.data_ptr(), 0, P2.data_ptr(), Q2.data_ptr(), R2.data_ptr(),
                                           0, ws.data_ptr(), wsb, st), "step_next")
P2, Q2, R2 = torch.empty_like(P), torch.empty_like(Q), torch.empty_like(R)
def first():  # step 1 (edge encoder folded in)
    _lib.check(lib.gnncca_mp_step_fwd_attr(C.byref(sb.cfg), C.byref(sb.w), plan.rowptr.data_ptr(), plan.src.data_ptr(), plan.dst.data_ptr(), N, E, P.data_ptr(), Q.data_ptr(),
                                           R.data_ptr(), b.edge_attr.data_ptr(), 4, le.weight.data_ptr(), le.bias.data_ptr(), 0, e2.data_ptr(), h_out.data_ptr(), 0, 0,
                                           ws.data_ptr(), wsb, st), "step1")
def last():   # step 4 (classify only)
    _lib.check(lib.gnncca_mp_step_fwd(C.byref(sb.cfg), C.byref(sb.w), plan.rowptr.data_ptr(), plan.src.data_ptr(), plan.dst.data_ptr(), N, E, P.data_ptr(), Q.data_ptr(),
                                      R.data_ptr(), 0, e.data_ptr(), 0, 0, lg.data_ptr(), 0, ws.data_ptr(), wsb, st), "step4")
def proj():
    _lib.check(lib.gnncca_node_project(C.byref(sb.cfg), C.byref(sb.w), 0, h.data_ptr(), N, P.data_ptr(), Q.data_ptr(), R.data_ptr(), st), "proj")
peak = 6554.9
for name, fn, byt in (("step2 (+fold)", full, E * 68 + N * 256), ("step2 (+fold+project)", full_next, E * 68 + N * 256), ("step1 attr (+fold)", first, E * 56 + N * 256), ("step4 classify", last, E * 36), ("node_project", proj, N * 320)):
    for _ in range(3): fn()
    a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    for _ in range(20): fn()
    z.record(); torch.cuda.synchronize()
    t = a.elapsed_time(z) / 20
    print("%-20s %8.1f us  %6.2f G edges/s  %5.1f%% of HBM peak on %d MB" % (name, t * 1e3, E / t / 1e6, byt / t / 1e6 / peak * 100, byt / 1e6))
proj(); Pr, Qr, Rr = P.clone(), Q.clone(), R.clone()
full_next()
_lib.check(lib.gnncca_node_project(C.byref(sb.cfg), C.byref(sb.w), 0, h_out.data_ptr(), N, Pr.data_ptr(), Qr.data_ptr(), Rr.data_ptr(), st), "proj")
torch.cuda.synchronize()
print("fused projection == node_project:", bool(torch.equal(P2, Pr)), bool(torch.equal(Q2, Qr)), bool(torch.equal(R2, Rr)))
print("N", N, "E", E, "checksum", float(h_out.double().sum()), float(lg.double().sum()))
