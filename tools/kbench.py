"""Kernel-level timings (CUDA events around back-to-back launches, inputs rotated through > L2-sized sets).
Usage: python tools/kbench.py [cfg2|cfg1|cfg3] -- development aid, not the bench of record (bench.py is)."""
import copy
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

gnncca = importlib.import_module("gnn-cca_b200")
engine = importlib.import_module("gnn-cca_b200.engine")
cfgmod = importlib.import_module("gnn-cca_b200.config")
syn = importlib.import_module("gnn-cca_b200.synthetic")
_lib = gnncca._lib


def timeit(fn, n=30, warm=3):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(n):
        fn(i)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3   # us


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
    wl = bench.WORKLOADS[name]
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    model = gnncca.MOTMPNet(copy.deepcopy(cfgmod.graph_net_params()), None, "resnet50", math="tf32").to(dev).eval()
    nset = 3
    bs = [gnncca.build_graph_batch(**syn.make_detections(wl["cams"], wl["dets"], wl["frames"], seed=wl["seed"] + s, device=dev,
                                                           max_dist=wl["max_dist"])) for s in range(nset)]
    N, E = bs[0].num_nodes, bs[0].num_edges
    print("workload", name, "N", N, "E", E)
    with torch.no_grad():
        enc = model.encoder.node_mlp
        hs = [torch.empty((N, 32), device=dev) for _ in range(nset)]
        t = timeit(lambda i: engine.mlp_forward_nograd(enc, bs[i % nset].x, _lib.MATH_TF32, out=hs[i % nset]))
        nb = N * (bs[0].x.size(1) * 4 + 128)
        print("node_encoder_tf32   %8.1f us   %7.1f GB/s  (%.1f%% of 6545)" % (t, nb / t / 1e3, nb / t / 1e3 / 65.453))
        plans = [gnncca.GraphPlan(b.edge_index, N).resolve() for b in bs]
        t = timeit(lambda i: gnncca.GraphPlan(bs[i % nset].edge_index, N))
        print("csr (GraphPlan ctor) %8.1f us" % t)
        for path in ("fused", "steps"):
            def run(i):
                b = bs[i % nset]
                engine.forward_inference(model, b.x, b.edge_index, b.edge_attr, plan=plans[i % nset], h_encoded=hs[i % nset],
                                         graph_ptr=b.ptr if path == "fused" else None, caps=b.gnncca_caps)
            t = timeit(run)
            print("MP part via %-6s   %8.1f us   (%.2f G edges/s)" % (path, t, E / t / 1e3))
        engine.CACHE_PLANS = False
        t = timeit(lambda i: model(bs[i % nset]))
        print("full forward         %8.1f us   (%.2f G edges/s)" % (t, E / t / 1e3))
        import cProfile, pstats
        pr = cProfile.Profile()
        pr.enable()
        for i in range(200):
            model(bs[i % nset])
        pr.disable()
        torch.cuda.synchronize()
        pstats.Stats(pr).sort_stats("cumulative").print_stats(22)
        # host cost of one forward (no sync inside): wall time of the Python call
        import time
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(50):
            model(bs[i % nset])
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        print("host time per forward call %8.1f us (includes the end-of-forward flag wait)" % ((t1 - t0) / 50 * 1e6))


if __name__ == "__main__":
    main()
