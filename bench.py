#!/usr/bin/env python
"""bench.py -- MPN edges/sec of the GNN-CCA message-passing hot path on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfg2]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one full ``MOTMPNet.forward`` (CSR build + node/edge encoders + 4 message-passing steps + 3
classifier passes) over one batch of synthetic graphs.  The default workload is BASELINE.json configs[1]
(EPFL-Laboratory-shaped: 4 cameras x 6 detections, 1024 frames per GPU).  Under torchrun every rank runs its
own batch of that size (independent per-frame graphs, no data-path collective): weak scaling.

One JSON line is printed by rank 0 (keys: see the driver contract in the task statement / DESIGN.md).
``--impl reference`` times the reference's algorithm on the host cores instead (the CPU restatement in
``oracle/``, pinned against vectors produced by the reference itself -- /root/reference does not exist on the
GPU box and its third-party dependencies are absent from the image).
"""
from __future__ import annotations

import argparse
import copy
import importlib
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (cams, dets, frames per GPU, CONV_TO_M of the dataset the shape comes from, data seed)
    "cfg1": dict(cams=4, dets=9, frames=64, max_dist=50.83, seed=1235, desc="EPFL-Terrace-shaped 4x9, batch 64"),
    "cfg2": dict(cams=4, dets=6, frames=1024, max_dist=44.23, seed=1236, desc="EPFL-Laboratory-shaped 4x6, batch 1024"),
    "cfg3": dict(cams=4, dets=10, frames=64, max_dist=153.54, seed=1237, train=True,
                 desc="EPFL-Basketball-shaped 4x10, batch 64, BCE training step (fwd + loss + bwd + SGD)"),
    "cfg4s": dict(cams=8, dets=64, frames=64, max_dist=50.83, seed=1238, desc="scaled 8x64, 64-graph slice of cfg 4"),
}
METRIC = "MPN edges/sec (4 MP steps)"
UNIT = "edges/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))
    ap.add_argument("--math", default=os.environ.get("GNNCCA_MATH", "tf32"), choices=["fp32", "tf32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="budget of the cpu_baseline leg")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    BAD = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20}
    NOTE = {"sw_power_cap": 0x4}

    def __init__(self, index: int, period: float = 0.005):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons = [], set()
        self.stop_flag = threading.Event()
        self.max_mhz = None
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception as exc:  # noqa: BLE001 -- NVML is optional plumbing; the JSON says when it is missing
            self.err = repr(exc)

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        while not self.stop_flag.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                mask = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                    else int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                for name, bit in {**self.BAD, **self.NOTE}.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:  # noqa: BLE001
                pass
            time.sleep(self.period)

    def result(self):
        if not self.ok:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "note": "nvml unavailable: " + getattr(self, "err", "?")}
        return {"sm_mhz": int(statistics.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def physical_gpu_index(local_index: int) -> int:
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local_index])
        except (ValueError, IndexError):
            return local_index
    return local_index


# ------------------------------------------------------------------------------------------------------------
# algorithmic bytes (DESIGN.md "Measurement")
# ------------------------------------------------------------------------------------------------------------
def forward_bytes(N: int, E: int, f_x: int = 2048) -> int:
    """B_fwd = B_enc + B_mp: encoder N*(4*f_x + 128) + E*(16 + 24); MP part 268 B/edge + 1024 B/node."""
    return N * (4 * f_x + 128) + E * (16 + 24) + E * 268 + N * 1024


def stage_bytes(stage: str, N: int, E: int, f_x: int, hidden: int = 128, fused_encoder: bool = False) -> int:
    if stage == "encoder_node":
        w = 4 * (f_x * hidden + hidden + hidden * 32 + 32)
        if fused_encoder:
            return N * (4 * f_x + 128) + w
        return N * (4 * f_x + 4 * hidden) + N * (4 * hidden + 128) + w     # two kernels: H1 written and re-read
    if stage == "encoder_edge":
        return E * (16 + 24)
    if stage == "csr":
        return E * (16 + 8) + 4 * (N + 1)
    if stage == "mp_fused":      # all steps on chip: edge_attr + int64 edge_index (RAW form) + k logits per edge; encoded node per node
        return E * (16 + 16 + 12) + N * 128
    if stage.startswith("step"):
        k = int(stage[4:])
        return E * 64 + N * 256 + (4 * E if k >= 2 else 0)
    raise KeyError(stage)


# ------------------------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the oracle restatement on the host cores
# ------------------------------------------------------------------------------------------------------------
def cpu_forward_timer(params, arch, sd, x, ei, ea, budget_s: float, warmup: int, steps=None):
    import torch
    from oracle import mpn_oracle
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    spec = mpn_oracle.MPNSpec(params, arch)
    times = []
    with torch.no_grad():
        for _ in range(warmup):
            mpn_oracle.motmpnet_forward(spec, sd, x, ei, ea)
        t_begin = time.perf_counter()
        while True:
            t0 = time.perf_counter()
            mpn_oracle.motmpnet_forward(spec, sd, x, ei, ea)
            times.append(time.perf_counter() - t0)
            if steps is not None:
                if len(times) >= steps:
                    break
            elif time.perf_counter() - t_begin > budget_s and len(times) >= 3:
                break
    return times, cores


def host_inputs(wl, node_dim=2048):
    """The workload's synthetic batch built with the CPU oracle of graph construction (same generator rule as
    gnn-cca_b200/synthetic.py: camera-major rows, every identity in every camera)."""
    import numpy as np
    import torch
    from oracle import graph_oracle
    C, D, B = wl["cams"], wl["dets"], wl["frames"]
    n = C * D
    g = torch.Generator().manual_seed(wl["seed"])
    k = np.arange(n)
    centers = torch.rand((B, D, 2), generator=g, dtype=torch.float64) * torch.tensor([360.0, 288.0], dtype=torch.float64)
    pts = centers[:, None].expand(B, C, D, 2) + 5.0 * torch.randn((B, C, D, 2), generator=g, dtype=torch.float64)
    pts = pts.reshape(B, n, 2).numpy()
    node = torch.relu(torch.randn((B * n, node_dim), generator=g))
    reid = torch.relu(torch.randn((B * n, 256), generator=g))
    frames = [dict(id=(k % D).astype(np.int64), id_cam=(k // D).astype(np.int64), xw=pts[b, :, 0], yw=pts[b, :, 1]) for b in range(B)]
    return graph_oracle.build_batch(frames, reid, node, [wl["max_dist"]] * B)


def run_reference(args):
    import torch
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfgmod = importlib.import_module("gnn-cca_b200.config")
    from oracle import mpn_oracle  # the reference arm and cpu_baseline are the only places bench.py runs oracle/
    wl = WORKLOADS[args.workload]
    params = cfgmod.graph_net_params()
    sd = mpn_oracle.init_state_dict(mpn_oracle.MPNSpec(params, "resnet50"), seed=0)
    b = host_inputs(wl)
    E, N = int(b.edge_index.size(1)), int(b.x.size(0))
    times, cores = cpu_forward_timer(params, "resnet50", sd, b.x, b.edge_index, b.edge_attr, 0.0, args.warmup, steps=args.steps)
    total = sum(times)
    val = E * len(times) / total
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": len(times),
            "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "%s: %s (N=%d nodes, E=%d directed edges per step)" % (args.workload, wl["desc"], N, E),
                       "l2": "n/a (host)"},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": "full %s batch per step, oracle/mpn_oracle.py (torch CPU fp32, %d threads)" % (args.workload, cores)},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------------------
# BASELINE config 3: training step (forward + summed BCE over the classified steps + backward + SGD), train.py:454-494
# ------------------------------------------------------------------------------------------------------------
def cpu_train_timer(params, sd0, x, ei, ea, labels, budget_s, steps=None):
    import torch
    from oracle import mpn_oracle
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    spec = mpn_oracle.MPNSpec(params, "resnet50")
    sd = {k: v.detach().clone().requires_grad_(v.is_floating_point()) for k, v in sd0.items()}
    opt = torch.optim.SGD([v for v in sd.values() if v.requires_grad], lr=1e-3, momentum=0.9, weight_decay=1e-4)
    crit = torch.nn.BCEWithLogitsLoss()
    times, t_begin = [], time.perf_counter()
    while True:
        t0 = time.perf_counter()
        opt.zero_grad()
        out = mpn_oracle.motmpnet_forward(spec, sd, x, ei, ea, training=True)["classified_edges"]
        loss = sum(crit(r.view(-1), labels) for r in out)
        loss.backward()
        opt.step()
        times.append(time.perf_counter() - t0)
        if (steps is not None and len(times) >= steps + 1) or (steps is None and time.perf_counter() - t_begin > budget_s and len(times) >= 4):
            break
    return times[1:], cores


def run_training(args):
    import torch
    import torch.distributed as dist
    wl = WORKLOADS[args.workload]
    world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    cfgmod = importlib.import_module("gnn-cca_b200.config")
    params = cfgmod.graph_net_params()
    if args.impl == "reference":
        if rank != 0:
            return
        from oracle import mpn_oracle
        b = host_inputs(wl)
        sd = mpn_oracle.init_state_dict(mpn_oracle.MPNSpec(params, "resnet50"), seed=0)
        times, cores = cpu_train_timer(params, sd, b.x, b.edge_index, b.edge_attr, b.edge_labels, 0.0, steps=args.steps)
        E = int(b.edge_index.size(1))
        val = E * len(times) / sum(times)
        print(json.dumps({"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": len(times),
                          "warmup": 1, "ms_per_step": 1e3 * sum(times) / len(times), "higher_is_better": True, "scaling": "weak",
                          "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": {"workload": "%s: %s" % (args.workload, wl["desc"])},
                          "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": "full batch per step, oracle + torch autograd"},
                          "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}), flush=True)
        return
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    gnncca = importlib.import_module("gnn-cca_b200")
    syn = importlib.import_module("gnn-cca_b200.synthetic")
    lib = gnncca._lib.load(build_if_missing=False)
    torch.manual_seed(0)
    model = gnncca.MOTMPNet(copy.deepcopy(params), None, "resnet50").to(dev).train()
    opt = torch.optim.SGD(model.parameters(), lr=1e-3, momentum=0.9, weight_decay=1e-4, fused=True)
    batches = [gnncca.build_graph_batch(**syn.make_detections(wl["cams"], wl["dets"], wl["frames"], seed=wl["seed"] + 100 * rank + 7919 * s,
                                                               device=dev, max_dist=wl["max_dist"])) for s in range(8)]
    N, E = batches[0].num_nodes, batches[0].num_edges
    in_bytes = sum(t.numel() * t.element_size() for t in (batches[0].x, batches[0].edge_index, batches[0].edge_attr, batches[0].edge_labels))

    def step_eager(b):
        opt.zero_grad(set_to_none=True)
        out = model(b)["classified_edges"]
        loss = gnncca.loss_stats(out, b.edge_labels).loss      # train.py:80-97 in one pass (gnncca_bce_stats_fwd/bwd)
        loss.backward()
        opt.step()
        return loss

    gstep = gnncca.GraphedTrainStep(model)     # forward + loss + backward captured once per batch shape, replayed afterwards

    def step(b):
        loss = gstep(b)                        # zeroes .grad, runs the same kernels as step_eager, out of one CUDA graph
        opt.step()
        return loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, n):
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        ev0.record()
        for i in range(n):
            fn(batches[i % 8])
        ev1.record()
        barrier()
        return ev0.elapsed_time(ev1)

    for i in range(max(args.warmup, 3)):
        step_eager(batches[i % 8])
    ms_eager = timed(step_eager, args.steps)
    for i in range(max(args.warmup, 3)):
        step(batches[i % 8])
    barrier()
    sampler = ClockSampler(physical_gpu_index(local))
    sampler.start()
    l0 = int(lib.gnncca_launch_count()) + gstep.launches_replayed
    ms = timed(step, args.steps)
    launches = int(lib.gnncca_launch_count()) + gstep.launches_replayed - l0
    t_end = time.perf_counter() + 0.3
    while time.perf_counter() < t_end:
        step(batches[0])
    torch.cuda.synchronize()
    sampler.stop_flag.set(); sampler.join()
    # end to end: inputs from pinned host memory every step, loss read back
    hosts = [[t.cpu().pin_memory() for t in (b.x, b.edge_index, b.edge_attr, b.edge_labels)] for b in batches[:2]]
    e2e_steps = max(3, min(args.steps, 20))
    barrier()
    t0 = time.perf_counter()
    sb = gstep.static_batch(batches[0])          # the graph's own input buffers: host data is copied straight into them
    for i in range(e2e_steps):
        hx, hei, hea, hl = hosts[i & 1]
        sb.x.copy_(hx, non_blocking=True); sb.edge_index.copy_(hei, non_blocking=True)
        sb.edge_attr.copy_(hea, non_blocking=True); sb.edge_labels.copy_(hl, non_blocking=True)
        float(step(sb).item())
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([ms, e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        ms_max, e2e_max = float(t[0]), float(t[1])
        fb = forward_bytes(N, E, int(batches[0].x.size(1)))
        train_bytes = fb + 4 * E * (24 + 24 + 24 + 32) + 3 * N * 8 * 2048          # + saved latents, gradient passes, dW1/x re-reads
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        ach = train_bytes / (ms_max / args.steps * 1e-3) / 1e9
        line = {"metric": METRIC, "value": world * E * args.steps / (ms_max * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "tf32", "data": "synthetic",
                "config": {"workload": "%s: %s per GPU (N=%d, E=%d)" % (args.workload, wl["desc"], N, E), "l2": "8 batches rotate; a batch is 22 MB (fits L2: launch-bound regime)",
                           "step": "forward + loss + backward replayed from one CUDA graph (GraphedTrainStep), then torch SGD; the same step launched eagerly: %.3f ms" % (ms_eager / args.steps),
                           "optimizer": "torch.optim.SGD(fused=True) lr 1e-3 momentum 0.9 wd 1e-4", "parallelism": "graph-sharded x%d, no gradient all-reduce (independent replicas)" % world},
                "clocks": sampler.result(), "gpu_launches": launches,
                "e2e": {"value": world * E * e2e_steps / e2e_max, "unit": UNIT, "h2d_bytes_per_step": in_bytes, "d2h_bytes_per_step": 4, "steps": e2e_steps,
                        "api": "GraphedTrainStep (MOTMPNet.forward + loss_stats + backward) + SGD.step, batch copied from pinned host memory into the graph's input buffers"},
                "roofline": {"bound": "hbm", "kernel": "whole training step (launch-bound at this size)", "achieved": ach, "peak": peak, "unit": "GB/s",
                             "frac": ach / peak, "traffic": None, "algorithmic_bytes_per_launch": train_bytes}}
        if world == 1 and not args.no_cpu_baseline:
            sd = {k: v.detach().cpu() for k, v in model.state_dict().items()}
            b = batches[0]
            times, cores = cpu_train_timer(params, sd, b.x.cpu(), b.edge_index.cpu(), b.edge_attr.cpu(), b.edge_labels.cpu(), args.cpu_seconds)
            line["cpu_baseline"] = {"value": E / statistics.median(times), "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": "%d training steps of the full batch (oracle forward + torch autograd + SGD), median" % len(times)}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------------------
def main():
    args = parse()
    if WORKLOADS[args.workload].get("train"):
        run_training(args)
        return
    if args.impl == "reference":
        run_reference(args)
        return
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path (use --impl reference for the host baseline)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    gnncca = importlib.import_module("gnn-cca_b200")
    cfgmod = importlib.import_module("gnn-cca_b200.config")
    syn = importlib.import_module("gnn-cca_b200.synthetic")
    engine = importlib.import_module("gnn-cca_b200.engine")
    lib = gnncca._lib.load(build_if_missing=False)

    wl = WORKLOADS[args.workload]
    params = cfgmod.graph_net_params()
    torch.manual_seed(0)
    model = gnncca.MOTMPNet(copy.deepcopy(params), None, "resnet50", math=args.math).to(dev).eval()
    # two different input batches per rank, used alternately: together 2 x (x 201 MB + edges) > the 126 MB L2,
    # so no step finds its inputs in cache
    batches = []
    for s in range(2):
        det = syn.make_detections(wl["cams"], wl["dets"], wl["frames"], seed=wl["seed"] + 100 * rank + 7919 * s, device=dev,
                                  max_dist=wl["max_dist"])
        batches.append(gnncca.build_graph_batch(**det))
        del det
    N, E, F = batches[0].num_nodes, batches[0].num_edges, int(batches[0].x.size(1))
    in_bytes = batches[0].x.numel() * 4 + batches[0].edge_index.numel() * 8 + batches[0].edge_attr.numel() * 4

    def step(i, prof=None):
        b = batches[i & 1]
        if prof is None:
            return model(b)["classified_edges"]     # the call a user makes (models/mpn.py:250)
        return engine.forward_inference(model, b.x, b.edge_index, b.edge_attr, prof=prof, graph_ptr=b.ptr, caps=b.gnncca_caps,
                                        trusted=True, edge_ptr=b.edge_ptr)   # same path with CUDA-event marks between the stages

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    with torch.no_grad():
        for i in range(max(args.warmup, 3)):
            step(i)
        barrier()
        sampler = ClockSampler(physical_gpu_index(local))
        sampler.start()
        launches0 = int(lib.gnncca_launch_count())
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        ev0.record()
        for i in range(args.steps):
            step(i)
        ev1.record()
        barrier()
        launches = int(lib.gnncca_launch_count()) - launches0
        ms = ev0.elapsed_time(ev1)
        # keep the clocks sampler alive for a representative amount of load when the timed region was very short
        if ms < 300:
            t_end = time.perf_counter() + 0.3
            i = 0
            while time.perf_counter() < t_end:
                step(i); i += 1
            torch.cuda.synchronize()
        sampler.stop_flag.set()
        sampler.join()

        # per-kernel device times, live: each of the forward's two kernels launched back to back through the C ABI on the
        # same alternating batches (CUDA events on the launching stream around K launches; no host gaps in between)
        import ctypes as C
        _lib = gnncca._lib
        fp = engine.fast_path(model)
        stage_avg = {}
        small_frames = all(engine.frame_capacity(model, N, E, b.num_graphs, b.gnncca_caps) is not None for b in batches)
        if fp is not None and args.math == "tf32" and small_frames:
            stream = torch.cuda.current_stream().cuda_stream
            l1, l2 = model.encoder.node_mlp.linear(0), model.encoder.node_mlp.linear(1)
            le = model.encoder.edge_mlp.linear(0)
            hs = [torch.empty((N, 32), dtype=torch.float32, device=dev) for _ in batches]
            lg = torch.empty((model.num_class_steps, E), dtype=torch.float32, device=dev)
            caps = [engine.frame_capacity(model, N, E, b.num_graphs, b.gnncca_caps) for b in batches]

            def enc(i):
                b = batches[i & 1]
                _lib.check(lib.gnncca_node_encoder(b.x.data_ptr(), l1.weight.data_ptr(), l1.bias.data_ptr(), l2.weight.data_ptr(), l2.bias.data_ptr(),
                                                   hs[i & 1].data_ptr(), N, l1.in_features, l1.out_features, 32, _lib.MATH_TF32, stream), "node_encoder")

            def mp(i):
                b, cap = batches[i & 1], caps[i & 1]
                _lib.check(lib.gnncca_mp_fused_fwd(C.byref(fp.sb.cfg), C.byref(fp.sb.w), le.weight.data_ptr(), le.bias.data_ptr(), fp.fa,
                                                   b.ptr.data_ptr(), b.num_graphs, b.edge_index.data_ptr(), b.edge_ptr.data_ptr(), None, None, None,
                                                   N, E, hs[i & 1].data_ptr(), b.edge_attr.data_ptr(), fp.L, fp.first, cap[0], cap[1], lg.data_ptr(),
                                                   None, None, fp.flags.data_ptr(), stream), "mp_fused_fwd")

            for name, fn in (("encoder_node", enc), ("mp_fused", mp)):
                for i in range(4):
                    fn(i)
                a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda.synchronize()
                a.record()
                for i in range(args.steps):
                    fn(i)
                z.record()
                torch.cuda.synchronize()
                stage_avg[name] = a.elapsed_time(z) / args.steps
        else:   # fp32 math / configurations without the single-call path: CUDA-event marks between the stages of the general pipeline
            profs = []
            for i in range(args.steps):
                pr = engine.StageTimer()
                step(i, prof=pr)
                pr.mark("end")
                profs.append(pr)
            torch.cuda.synchronize()
            stage_ms = {}
            for pr in profs:
                for name, tt in pr.stages():
                    stage_ms.setdefault(name, []).append(tt)
            stage_avg = {k: sum(v) / len(v) for k, v in stage_ms.items()}

    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    value = world * E * args.steps / (ms_max * 1e-3)

    # ---------------------------------------------------------------- end to end through the public host API
    with torch.no_grad():
        hosts = []
        for b in batches:
            hosts.append((b.x.cpu().pin_memory(), b.edge_index.cpu().pin_memory(), b.edge_attr.cpu().pin_memory()))
        ptrs = [b.ptr.cpu().pin_memory() for b in batches]
        outs_host = [torch.empty((E, 1), dtype=torch.float32).pin_memory() for _ in range(model.num_class_steps)]
        e2e_steps = max(3, min(args.steps, 20))
        for i in range(2):
            model.forward_host(*hosts[i & 1], out=outs_host, graph_ptr=ptrs[i & 1], caps=batches[i & 1].gnncca_caps)
        barrier()
        t0 = time.perf_counter()
        for i in range(e2e_steps):
            model.forward_host(*hosts[i & 1], out=outs_host, graph_ptr=ptrs[i & 1], caps=batches[i & 1].gnncca_caps)     # returns after the D2H copy of the logits
        torch.cuda.synchronize()
        e2e_blocking_s = time.perf_counter() - t0
        # the same stream of batches through HostPipeline: batch i+1 is copied in while batch i computes and its logits
        # travel back; every step still moves its inputs host->device and its logits device->host, and every result is
        # waited for (one step later) inside the timed region
        pipe = gnncca.HostPipeline(model)
        outs_pair = [outs_host, [torch.empty((E, 1), dtype=torch.float32).pin_memory() for _ in range(model.num_class_steps)]]
        for i in range(2):
            pipe.submit(*hosts[i & 1], out=outs_pair[i & 1], graph_ptr=ptrs[i & 1], caps=batches[i & 1].gnncca_caps).wait()
        barrier()
        t0 = time.perf_counter()
        pending = None
        for i in range(e2e_steps):
            hd = pipe.submit(*hosts[i & 1], out=outs_pair[i & 1], graph_ptr=ptrs[i & 1], caps=batches[i & 1].gnncca_caps)
            if pending is not None:
                pending.wait()
            pending = hd
        pending.wait()
        torch.cuda.synchronize()
        e2e_s = time.perf_counter() - t0
        # what the link alone allows: the step's input bytes, pinned host -> device, nothing else running
        x_probe = torch.empty_like(batches[0].x)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(3):
            x_probe.copy_(hosts[i & 1][0], non_blocking=True)
        torch.cuda.synchronize()
        h2d_gbs = 3 * x_probe.numel() * 4 / (time.perf_counter() - t0) / 1e9
        del x_probe
    t = torch.tensor([e2e_s, e2e_blocking_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * E * e2e_steps / float(t[0].item())
    e2e_blocking = world * E * e2e_steps / float(t[1].item())
    d2h_bytes = model.num_class_steps * E * 4

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---------------------------------------------------------------- roofline of the dominant kernel
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except OSError:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    fused = bool(getattr(engine, "FUSED_NODE_ENCODER", False)) and args.math == "tf32"
    kern = {k: v for k, v in stage_avg.items() if k not in ("begin", "end")}
    traffic = {}
    prof_file = os.path.join(ROOT, "profiles", "traffic.json")     # dram bytes per launch from the committed ncu --set full capture
    if os.path.exists(prof_file):
        try:
            with open(prof_file) as f:
                traffic = json.load(f).get(args.workload, {})
        except (OSError, ValueError):
            pass
    per_kernel = {}
    for k, ms_k in kern.items():
        try:
            b_k = stage_bytes(k, N, E, F, fused_encoder=fused)
        except KeyError:
            continue
        per_kernel[k] = {"ms": round(ms_k, 5), "algorithmic_bytes": b_k, "achieved": b_k / (ms_k * 1e-3) / 1e9,
                         "frac": b_k / (ms_k * 1e-3) / 1e9 / peak, "traffic": traffic.get(k)}
    # the kernel the HBM roofline is about: the one that moves the forward's bytes (the node encoder reads 8 KB per node,
    # > 90 % of the forward's DRAM traffic).  The message-passing kernel keeps its state on chip and is bound by FP32 /
    # shared-memory issue slots instead; its line is in "kernels".
    dom = max(per_kernel, key=lambda k: per_kernel[k]["algorithmic_bytes"])
    fwd_b = forward_bytes(N, E, F)
    t_step = ms_max / args.steps * 1e-3
    b_min = N * 4 * F + E * (16 + 16 + 12)
    roofline = {"bound": "hbm", "kernel": dom, "achieved": per_kernel[dom]["achieved"], "peak": peak, "unit": "GB/s",
                "frac": per_kernel[dom]["frac"], "traffic": per_kernel[dom]["traffic"], "peak_source": peak_src,
                "algorithmic_bytes_per_launch": per_kernel[dom]["algorithmic_bytes"], "kernel_ms": per_kernel[dom]["ms"],
                "kernels": per_kernel,
                "forward": {"algorithmic_bytes": fwd_b, "achieved": fwd_b / t_step / 1e9, "frac": fwd_b / t_step / 1e9 / peak,
                            "compulsory_bytes": b_min, "frac_of_compulsory": b_min / t_step / 1e9 / peak,
                            "note": "SURVEY 8d contract figure B_fwd (per-step traffic of an unfused pipeline) over the whole-forward time; "
                                    "the fused kernels move only the compulsory bytes, quoted beside it"}}

    # ---------------------------------------------------------------- cpu baseline (the oracle port, rank 0, N=1)
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        sd = {k: v.detach().cpu() for k, v in model.state_dict().items()}
        b = batches[0]
        times, cores = cpu_forward_timer(params, "resnet50", sd, b.x.cpu(), b.edge_index.cpu(), b.edge_attr.cpu(), args.cpu_seconds, 2)
        cpu = {"value": E / statistics.median(times), "unit": UNIT, "cores": cores, "kind": "port",
               "sample": "%d forwards of the full %s batch (E=%d) through oracle/mpn_oracle.py, torch CPU fp32, median" % (len(times), args.workload, E)}

    # graph + edge-attribute construction (train.py:257-361,450 -> build_graph_batch), timed separately: the call
    # includes both column normalisations, the edge count pass, the tiled build kernel and its one small read-back
    det = syn.make_detections(wl["cams"], wl["dets"], wl["frames"], seed=wl["seed"] + 100 * rank + 31, device=dev, max_dist=wl["max_dist"])
    for _ in range(2):
        gnncca.build_graph_batch(**det)
    torch.cuda.synchronize()
    reps = 5
    t0 = time.perf_counter()
    for _ in range(reps):
        gnncca.build_graph_batch(**det)
    torch.cuda.synchronize()
    gb_s = (time.perf_counter() - t0) / reps
    gb_bytes = N * (2048 + 256) * 4 * 3 + E * (16 + 16 + 4 + 2 * 256 * 4 / max(1, wl["dets"]))     # normalise (2 reads + 1 write), outputs, ReID rows per tile
    graph_build = {"value": E / gb_s, "unit": "edges built/s", "ms": gb_s * 1e3,
                   "what": "build_graph_batch: column normalisation of both embeddings + edge_index / edge_attr / labels / ptr of %d frames" % wl["frames"],
                   "approx_bytes": int(gb_bytes), "approx_frac_of_hbm_peak": gb_bytes / gb_s / 1e9 / peak}
    del det

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "tf32" if args.math == "tf32" else "f32", "data": "synthetic",
            "config": {"workload": "%s: %s per GPU (N=%d nodes, E=%d directed edges per step per GPU); MOTMPNet L=4, k=3, resnet50 dims"
                                   % (args.workload, wl["desc"], N, E),
                       "parallelism": "graph-sharded x%d, no collective" % world,
                       "l2": "two input batches used alternately (2 x %.0f MB > 126 MB L2)" % (in_bytes / 1e6),
                       "graphs_per_s": world * wl["frames"] * args.steps / (ms_max * 1e-3)},
            "clocks": sampler.result(), "gpu_launches": launches,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": in_bytes + 8 * (wl["frames"] + 1), "d2h_bytes_per_step": d2h_bytes,
                    "steps": e2e_steps, "api": "gnncca.HostPipeline.submit / wait (pinned host x/edge_index/edge_attr -> pinned host logits, two batches in flight)",
                    "h2d_link_gbs": h2d_gbs, "link_bound_value": world * E / ((in_bytes + 8 * (wl["frames"] + 1)) / (h2d_gbs * 1e9)),
                    "blocking_call_value": e2e_blocking, "blocking_api": "MOTMPNet.forward_host (returns after the logits are on the host)"},
            "roofline": roofline, "graph_build": graph_build}
    if cpu is not None:
        line["cpu_baseline"] = cpu
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
