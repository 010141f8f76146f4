"""Wall time of ``build_graph_batch`` at a BASELINE shape and of its stages launched alone through the C ABI.

    python tools/build_stages.py [--cams 4 --dets 6 --frames 1024] [--reps 200]
"""
import argparse
import importlib
import os
import statistics
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import gnncca  # noqa: E402

_lib = importlib.import_module("gnn-cca_b200._lib")
graph = importlib.import_module("gnn-cca_b200.graph")
syn = importlib.import_module("gnn-cca_b200.synthetic")


def wall(fn, reps, flush=None):
    ts = []
    for _ in range(reps):
        if flush is not None:
            flush.add_(1.0)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        ts.append((time.perf_counter() - t0) * 1e6)
    return statistics.median(ts), min(ts)


def dev_time(fn, reps, flush=None):
    ts = []
    for _ in range(reps):
        if flush is not None:
            flush.add_(1.0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    return statistics.median(ts), min(ts)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cams", type=int, default=4)
    ap.add_argument("--dets", type=int, default=6)
    ap.add_argument("--frames", type=int, default=1024)
    ap.add_argument("--reps", type=int, default=200)
    ap.add_argument("--flush", action="store_true", help="write 256 MB between repetitions (cold L2)")
    ap.add_argument("--profile", action="store_true", help="cProfile of 2000 build_graph_batch calls (host side)")
    ap.add_argument("--trace", action="store_true", help="host time stamps at every C call of build_graph_batch (median over --reps)")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    det = syn.make_detections(a.cams, a.dets, a.frames, seed=5, device=dev)
    flush = torch.zeros(64 << 20, device=dev) if a.flush else None
    for _ in range(5):
        b = gnncca.build_graph_batch(**det)
    print("N = %d nodes, E = %d edges, %d frames" % (b.num_nodes, b.num_edges, a.frames))
    m, lo = wall(lambda: gnncca.build_graph_batch(**det), a.reps, flush)
    print("build_graph_batch wall: median %.1f us min %.1f us" % (m, lo))
    if a.trace:
        lib = _lib.load()
        marks = []
        names = ["gnncca_graph_prepare_all", "gnncca_build_graphs", "gnncca_graph_join"]
        orig = {n: getattr(lib, n) for n in names}
        orig_poll = graph._poll_totals

        def wrap(n, f):
            def g(*args):
                marks.append((n + ":enter", time.perf_counter()))
                r = f(*args)
                marks.append((n + ":exit", time.perf_counter()))
                return r
            return g
        for n in names:
            setattr(lib, n, wrap(n, orig[n]))
        graph._poll_totals = wrap("poll_totals", orig_poll)
        rows = []
        for _ in range(a.reps):
            torch.cuda.synchronize()
            del marks[:]
            t0 = time.perf_counter()
            gnncca.build_graph_batch(**det)
            t1 = time.perf_counter()
            torch.cuda.synchronize()
            t2 = time.perf_counter()
            rows.append([(k, (t - t0) * 1e6) for k, t in marks] + [("return", (t1 - t0) * 1e6), ("synchronized", (t2 - t0) * 1e6)])
        for i, (k, _) in enumerate(rows[0]):
            print("  %-34s %7.1f us" % (k, statistics.median(r[i][1] for r in rows)))
        for n in names:
            setattr(lib, n, orig[n])
        graph._poll_totals = orig_poll
    if a.profile:
        import cProfile
        import pstats
        pr = cProfile.Profile()
        pr.enable()
        for _ in range(2000):
            gnncca.build_graph_batch(**det)
        torch.cuda.synchronize()
        pr.disable()
        pstats.Stats(pr).sort_stats("tottime").print_stats(18)
    m, lo = wall(lambda: gnncca.build_graph_batch(**det, lazy_x=False), a.reps, flush)
    print("build_graph_batch(lazy_x=False) wall: median %.1f us min %.1f us" % (m, lo))
    m, lo = dev_time(lambda: graph.column_norms(det["node_embeds"]), a.reps, flush)
    print("column_norms(node) device: median %.1f us min %.1f us" % (m, lo))
    m, lo = dev_time(lambda: graph.column_norms(det["reid_embeds"]), a.reps, flush)
    print("column_norms(reid) device: median %.1f us min %.1f us" % (m, lo))
    m, lo = dev_time(lambda: graph.column_normalize(det["reid_embeds"]), a.reps, flush)
    print("column_normalize(reid) device: median %.1f us min %.1f us" % (m, lo))
    # the edge fill alone, on the tables of the last build
    lib = _lib.load()
    B, N = a.frames, b.num_nodes
    nbytes = int(lib.gnncca_graph_prepare_scratch_bytes(B, N))
    scratch = torch.empty((nbytes + 7) // 8, dtype=torch.int64, device=dev)
    edge_ptr = torch.empty(B + 1, dtype=torch.int64, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    gp = det["graph_ptr"].to(dev).contiguous()
    cam = det["id_cam"].to(device=dev, dtype=torch.int32).contiguous()

    def counts():
        _lib.check(lib.gnncca_graph_edge_counts(_lib.ptr(gp), _lib.ptr(cam), B, N, _lib.ptr(edge_ptr), _lib.ptr(scratch), scratch.numel() * 8, st), "counts")
    m, lo = dev_time(counts, a.reps, flush)
    print("graph_edge_counts device: median %.1f us min %.1f us" % (m, lo))
    off = (lib.gnncca_graph_tile_total_ptr(_lib.ptr(scratch), B, N) - scratch.data_ptr()) // 8
    n_tiles, E = (int(v) for v in scratch[off:off + 2].tolist())
    reid = det["reid_embeds"].contiguous()
    norm = 1.0 / graph.column_norms(reid)
    ids = det["ids"].to(device=dev, dtype=torch.int64).contiguous()
    ei = torch.empty((2, E), dtype=torch.int64, device=dev)
    ea = torch.empty((E, 4), dtype=torch.float32, device=dev)
    el = torch.empty(E, dtype=torch.float32, device=dev)

    def fill():
        _lib.check(lib.gnncca_build_graphs(_lib.ptr(gp), _lib.ptr(edge_ptr), B, N, E, n_tiles, _lib.ptr(cam), _lib.ptr(ids), _lib.ptr(det["xw"]),
                                           _lib.ptr(det["yw"]), _lib.ptr(det["max_dist"]), _lib.ptr(reid), _lib.ptr(norm), reid.size(1), 0, a.cams * a.dets,
                                           _lib.ptr(scratch), _lib.ptr(ei), _lib.ptr(ea), _lib.ptr(el), st), "fill")
    m, lo = dev_time(fill, a.reps, flush)
    print("build_graphs device, raw reid + norms (%d tiles): median %.1f us min %.1f us" % (n_tiles, m, lo))
    reid = graph.column_normalize(reid)
    norm = None
    m, lo = dev_time(fill, a.reps, flush)
    print("build_graphs device, normalised reid (%d tiles): median %.1f us min %.1f us" % (n_tiles, m, lo))


if __name__ == "__main__":
    main()
