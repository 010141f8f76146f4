#!/usr/bin/env python
"""bench.py -- MPN edges/sec of the GNN-CCA message-passing hot path on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfgX] [--math tf32|fp32]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one full ``MOTMPNet.forward`` (CSR build + node / edge encoders + 4 message-passing steps + 3 classifier
passes, reference models/mpn.py:250-299) over one batch of synthetic graphs (cfg 3: forward + loss + backward + SGD).

Workloads = the five BASELINE.json configs:
  cfg1  4 x 9  detections, 64 frames per GPU     (weak scaling, frames are independent: no collective)
  cfg2  4 x 6  detections, 1024 frames per GPU   (weak; the 1-GPU headline)
  cfg3  4 x 10 detections, 64 frames, training step (weak, independent replicas)
  cfg4  8 x 64 detections, 4096 frames IN TOTAL, frame-sharded over the ranks (strong scaling, no collective)
  cfg5  ONE graph of 16 x 512 detections (62.9 M directed edges), edges partitioned by source range over the ranks, NCCL
        all-gather of the node features after node updates 1-3 (strong scaling)
  cfg4s a 64-frame slice of cfg 4 per GPU (kernel work)

Line of record when no --workload is given: cfg2 at --gpus 1 (the config BASELINE.json names for 1 x B200), cfg5 at
--gpus N > 1 (the config with a real exchange step); the other configs are attached under "workloads" of the same JSON line.
Every workload entry carries a ``parity`` object: the timed path's logits against the CPU oracle (oracle/), in the same run.

``--impl reference`` times the reference's algorithm on the host cores (the CPU restatement in ``oracle/``, pinned against
vectors produced by the reference itself -- /root/reference does not exist on the GPU box) on a bounded sample of the same
workload; its ``config.workload`` string is byte-identical to our arm's.
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
    "cfg1": dict(kind="frames", cams=4, dets=9, frames=64, max_dist=50.83, seed=1235, scaling="weak",
                 desc="EPFL-Terrace-shaped 4x9, batch 64 per GPU"),
    "cfg2": dict(kind="frames", cams=4, dets=6, frames=1024, max_dist=44.23, seed=1236, scaling="weak",
                 desc="EPFL-Laboratory-shaped 4x6, batch 1024 per GPU"),
    "cfg3": dict(kind="train", cams=4, dets=10, frames=64, max_dist=153.54, seed=1237, scaling="weak",
                 desc="EPFL-Basketball-shaped 4x10, batch 64 per GPU, BCE training step (fwd + loss + bwd + SGD)"),
    "cfg4": dict(kind="frames", cams=8, dets=64, frames=4096, max_dist=50.83, seed=1238, scaling="strong",
                 desc="scaled 8x64, batch 4096 graphs in total, graph-sharded over the GPUs"),
    "cfg4s": dict(kind="frames", cams=8, dets=64, frames=64, max_dist=50.83, seed=1238, scaling="weak",
                  desc="scaled 8x64, 64-graph slice of cfg 4 per GPU"),
    "cfg5": dict(kind="giant", cams=16, dets=512, max_dist=50.83, seed=1239, scaling="strong",
                 desc="single giant graph 16x512 (31.5 M cross-camera pairs = 62.9 M directed edges), edge-partitioned by source range, "
                      "NCCL all-gather of node features per step"),
}
METRIC = "MPN edges/sec (4 MP steps)"
UNIT = "edges/s"
GEN_BLOCK = 64        # frames per generation block of the sharded workloads: block b is seeded seed + b on every rank count


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS))
    ap.add_argument("--math", default=os.environ.get("GNNCCA_MATH", "tf32"), choices=["fp32", "tf32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-attach", action="store_true", help="only the line of record, no attached workloads")
    ap.add_argument("--cpu-seconds", type=float, default=10.0, help="budget of each cpu_baseline leg")
    ap.add_argument("--window-seconds", type=float, default=2.0, help="budget of the repeated timing windows of one workload")
    return ap.parse_args()


def workload_name(name: str, world: int) -> str:
    """The ``config.workload`` string: a function of the workload and the GPU count only, identical in both arms."""
    wl = WORKLOADS[name]
    if wl["kind"] == "giant":
        n = wl["cams"] * wl["dets"]
        return "%s: %s (N=%d nodes, E=%d directed edges per step over %d GPU%s); MOTMPNet L=4, k=3, resnet50 dims" % (
            name, wl["desc"], n, n * (n - wl["dets"]), world, "" if world == 1 else "s")
    n = wl["cams"] * wl["dets"]
    e = n * n - wl["cams"] * wl["dets"] * wl["dets"]
    frames = wl["frames"] if wl["scaling"] == "strong" else wl["frames"] * world
    return "%s: %s (N=%d nodes, E=%d directed edges per step over %d GPU%s); MOTMPNet L=4, k=3, resnet50 dims" % (
        name, wl["desc"], frames * n, frames * e, world, "" if world == 1 else "s")


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

    def finish(self):
        self.stop_flag.set()
        if self.is_alive():
            self.join()
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


def compulsory_bytes(N: int, E: int, f_x: int = 2048) -> int:
    """B_min: x, edge_index and edge_attr read once, three logits written (SURVEY 8d "lower bound")."""
    return N * 4 * f_x + E * (16 + 16 + 12)


def stage_bytes(stage: str, N: int, E: int, f_x: int, hidden: int = 128, fused_encoder: bool = True, L: int = 4, first_cls: int = 2) -> int:
    """Bytes a stage REALLY moves at least (its own inputs and outputs once), used for the per-kernel fractions."""
    if stage == "encoder_node":
        w = 4 * (f_x * hidden + hidden + hidden * 32 + 32)
        if fused_encoder:
            return N * (4 * f_x + 128) + w
        return N * (4 * f_x + 4 * hidden) + N * (4 * hidden + 128) + w     # two kernels: H1 written and re-read
    if stage == "csr":
        return E * (16 + 8) + 4 * (N + 1)
    if stage == "mp_fused":      # all steps on chip: edge_attr + int64 edge_index (RAW form) + k logits per edge; encoded node per node
        return E * (16 + 16 + 12) + N * 128
    if stage.startswith("project"):
        return N * (128 + 32 + 32 + 128)
    if stage.startswith("step"):
        k = int(stage[4:])
        cls = 4 * E if k >= first_cls else 0
        ids = 8 * E                                          # int32 src + dst
        if k == 1:                                           # edge encoder folded in: raw attributes in, e' out
            return E * (16 + 24) + ids + cls + N * (32 + 32 + 128 + 128)
        if k == L:                                           # classify only: e^L is never written, no node update
            return E * 24 + ids + cls + N * (32 + 32)
        return E * (24 + 24) + ids + cls + N * (32 + 32 + 128 + 128)
    raise KeyError(stage)


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except (OSError, ValueError):
        peaks = {}
    if "hbm_gbs" in peaks:
        return float(peaks["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback 6650 GB/s (B200_PROFILING.md)"


def load_traffic(name: str) -> dict:
    """DRAM bytes per launch from the committed ncu --set full capture (profiles/traffic.json)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get(name, {})
    except (OSError, ValueError):
        return {}


def kernel_table(stage_avg: dict, N: int, E: int, F: int, peak: float, traffic: dict, t_step_ms: float):
    """Per-kernel roofline entries + the name of the kernel with the largest share of the step."""
    per = {}
    for k, ms_k in stage_avg.items():
        try:
            b_k = stage_bytes(k, N, E, F)
        except KeyError:
            continue
        if ms_k <= 0:
            continue
        ach = b_k / (ms_k * 1e-3) / 1e9
        per[k] = {"ms": round(ms_k, 5), "share_of_step": round(ms_k / t_step_ms, 4) if t_step_ms > 0 else None, "bytes": b_k,
                  "achieved": ach, "frac": ach / peak, "traffic": traffic.get(k)}
        if k.startswith("step") and 1 < int(k[4:]) < 4:      # a full step (edge update + node update): the SURVEY 8d contract figure beside it
            b_c = E * 64 + N * 256 + 4 * E
            per[k]["contract_bytes"] = b_c
            per[k]["contract_frac"] = b_c / (ms_k * 1e-3) / 1e9 / peak
    dom = max(per, key=lambda k: per[k]["ms"]) if per else None
    return per, dom


# ------------------------------------------------------------------------------------------------------------
# context + timing
# ------------------------------------------------------------------------------------------------------------
class Ctx:
    def __init__(self, args):
        import torch
        import torch.distributed as dist
        self.args, self.torch, self.dist = args, torch, dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a CUDA device: the product has no CPU path (use --impl reference for the host baseline)")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            import datetime
            # a bounded collective timeout: a rank that dies must end the run, not hang the box
            dist.init_process_group("nccl", device_id=self.dev, timeout=datetime.timedelta(seconds=240))
        self.gnncca = importlib.import_module("gnn-cca_b200")
        self.cfgmod = importlib.import_module("gnn-cca_b200.config")
        self.syn = importlib.import_module("gnn-cca_b200.synthetic")
        self.engine = importlib.import_module("gnn-cca_b200.engine")
        self.mg = importlib.import_module("gnn-cca_b200.multigpu")
        self.lib = self.gnncca._lib.load(build_if_missing=False)
        self.params = self.cfgmod.graph_net_params()
        self.peak, self.peak_src = load_peaks()

    def model(self, math=None, train=False):
        self.torch.manual_seed(0)
        m = self.gnncca.MOTMPNet(copy.deepcopy(self.params), None, "resnet50", math=math or self.args.math).to(self.dev)
        return m.train() if train else m.eval()

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, values):
        t = self.torch.tensor(list(values), dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return t.tolist()

    def min_over_ranks(self, values):
        t = self.torch.tensor(list(values), dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MIN)
        return t.tolist()

    def launches(self):
        return int(self.lib.gnncca_launch_count())

    def close(self):
        if self.world > 1:
            self.dist.destroy_process_group()


def tail_steps(ms_per_step: float, seconds: float = 0.25) -> int:
    """Steps that fill about ``seconds`` of device time: a COUNT derived from the rank-agreed step time, never a per-rank clock."""
    return int(max(3, min(5000, seconds * 1e3 / max(ms_per_step, 1e-4))))


def timed_windows(ctx: Ctx, step, steps: int, warmup: int, budget_s: float, extra_launches=lambda: 0):
    """W warm-up steps, then windows of EXACTLY ``steps`` steps, each bracketed by barrier + synchronize on both sides and timed
    with CUDA events on the launching stream; per window the MAX over ranks; the reported time is the MEDIAN window (min / max
    / first beside it).  The number of windows (3..50) is fixed from the first window's duration so that all ranks agree."""
    torch = ctx.torch
    for i in range(max(warmup, 3)):
        step(i)
    ctx.barrier()

    def one(base):
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ctx.barrier()
        l0 = ctx.launches() + extra_launches()
        ev0.record()
        for i in range(steps):
            step(base + i)
        ev1.record()
        ctx.barrier()
        return ev0.elapsed_time(ev1), ctx.launches() + extra_launches() - l0

    first_ms, launches = one(0)
    first_max = ctx.max_over_ranks([first_ms])[0]
    n_more = int(max(2, min(49, budget_s * 1e3 / max(first_max, 1e-3) - 1)))
    local = [first_ms] + [one((w + 1) * steps)[0] for w in range(n_more)]
    per_window = ctx.max_over_ranks(local)
    med = statistics.median(per_window)
    return {"ms_per_step": med / steps, "launches": launches,
            "windows": {"n": len(per_window), "steps_per_window": steps, "median_ms": med, "min_ms": min(per_window), "max_ms": max(per_window),
                        "first_ms": per_window[0]}}


ERR_FORMULA = "max |got - ref| / (1 + |ref| + 0.01 max|ref|)  (the last term is the fp32 reassociation floor: 1e-6 of the tensor's largest magnitude at tol 1e-4)"


def x_in(b):
    """The node-feature matrix a forward on ``b`` will read: the raw embeddings of a lazily normalised batch (graph.Batch.x), else x.
    (Reading ``b.x`` would materialise the normalised copy and change what the timed forward does.)"""
    raw = b.raw_x() if hasattr(b, "raw_x") else None
    return raw[0] if raw is not None else b.x


def x_rows(b, n0, n1):
    """Rows [n0, n1) of the NORMALISED node features (what the reference's MPN is given), without materialising the whole matrix."""
    raw = b.raw_x() if hasattr(b, "raw_x") else None
    return (raw[0][n0:n1] / raw[1]) if raw is not None else b.x[n0:n1]


def rel_err(out, ref):
    """The normalised error of tests/test_gpu_mpn.py::close over a list of tensors (CPU): max |out - ref| / (1 + |ref| + 0.01 max|ref|).
    With sum aggregation and random weights the latents reach 1e13 at degree 7680; an element that is a cancelling sum of such terms
    cannot be reproduced relative to its own size by ANY other fp32 summation order (the reference's atomicAdd path included)."""
    worst = 0.0
    for o, r in zip(out, ref):
        if r.numel():
            r = r.double()
            worst = max(worst, float(((o.double() - r).abs() / (1 + r.abs() + 0.01 * r.abs().max())).max()))
    return worst


def decisions_equal(out, ref, margin: float):
    """Edge decisions sigmoid(logit) >= 0.5 agree on every edge whose reference logit is further than ``margin`` from 0."""
    ok, n_near, n = True, 0, 0
    for o, r in zip(out, ref):
        far = r.abs() > margin
        n_near += int((~far).sum())
        n += r.numel()
        ok = ok and bool(((o >= 0) == (r >= 0))[far].all())
    return ok, n_near, n


# ------------------------------------------------------------------------------------------------------------
# CPU side: the oracle restatement on the host cores (reference arm / cpu_baseline / parity)
# ------------------------------------------------------------------------------------------------------------
def cpu_forward_timer(params, arch, sd, x, ei, ea, budget_s: float, warmup: int, steps=None):
    import torch
    from oracle import mpn_oracle
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    spec = mpn_oracle.MPNSpec(params, arch)
    times, out = [], None
    with torch.no_grad():
        for _ in range(warmup):
            out = mpn_oracle.motmpnet_forward(spec, sd, x, ei, ea)
        t_begin = time.perf_counter()
        while True:
            t0 = time.perf_counter()
            out = mpn_oracle.motmpnet_forward(spec, sd, x, ei, ea)
            times.append(time.perf_counter() - t0)
            if steps is not None:
                if len(times) >= steps:
                    break
            elif time.perf_counter() - t_begin > budget_s and len(times) >= 3:
                break
    return times, cores, out["classified_edges"]


def host_frames(wl, n_frames: int, seed: int, node_dim=2048):
    """``n_frames`` frames of the workload's shape built with the CPU oracle of graph construction (same generator rule as
    gnn-cca_b200/synthetic.py: camera-major rows, every identity in every camera).  Returns (Batch, frames, reid, node)."""
    import numpy as np
    import torch
    from oracle import graph_oracle
    C, D, B = wl["cams"], wl["dets"], n_frames
    n = C * D
    g = torch.Generator().manual_seed(seed)
    k = np.arange(n)
    centers = torch.rand((B, D, 2), generator=g, dtype=torch.float64) * torch.tensor([360.0, 288.0], dtype=torch.float64)
    pts = centers[:, None].expand(B, C, D, 2) + 5.0 * torch.randn((B, C, D, 2), generator=g, dtype=torch.float64)
    pts = pts.reshape(B, n, 2).numpy()
    node = torch.relu(torch.randn((B * n, node_dim), generator=g))
    reid = torch.relu(torch.randn((B * n, 256), generator=g))
    frames = [dict(id=(k % D).astype(np.int64), id_cam=(k // D).astype(np.int64), xw=pts[b, :, 0], yw=pts[b, :, 1]) for b in range(B)]
    return graph_oracle.build_batch(frames, reid, node, [wl["max_dist"]] * B), frames, reid, node


def giant_inputs(wl, node_range, device, full_norm=True):
    """cfg 5 inputs for the sources in ``node_range``: x (all N nodes, column-normalised), this range's slice of the closed-form
    cross-camera enumeration (train.py:292-299) and its edge attributes (a deterministic function of the global edge id, so
    that every partition sees the values of the whole graph)."""
    import torch
    mg = importlib.import_module("gnn-cca_b200.multigpu")
    C, D = wl["cams"], wl["dets"]
    N, deg = C * D, C * D - D
    g = torch.Generator(device=device).manual_seed(wl["seed"])
    x = torch.relu(torch.randn((N, 2048), generator=g, device=device))
    x = x / x.norm(dim=0).clamp_min(1e-12)
    n0, n1 = node_range
    ei = mg.complete_multipartite_slice(C, D, (n0, n1), device)
    eid = torch.arange(n0 * deg, n1 * deg, device=device, dtype=torch.float32)
    ea = torch.stack([torch.frac(eid * 0.61803398875), torch.frac(eid * 0.41421356237), torch.frac(eid * 0.73205080757),
                      torch.frac(eid * 0.23606797750)], dim=1).contiguous()
    return x, ei, ea


REF_SAMPLE = {"cfg4": 2, "cfg4s": 2, "cfg5": 256}     # bounded CPU samples: frames (cfg 4) / source rows (cfg 5) per step


def cpu_sample(name: str, world: int = 1):
    """(x, edge_index, edge_attr, sample description) of the bounded CPU sample of a workload (host tensors)."""
    import torch
    wl = WORKLOADS[name]
    if wl["kind"] == "giant":
        rows = REF_SAMPLE[name]
        x, ei, ea = giant_inputs(wl, (0, rows), torch.device("cpu"))
        return x, ei, ea, "%d source rows of the graph (%d edges) against all %d nodes per step" % (rows, ei.size(1), x.size(0))
    frames = REF_SAMPLE.get(name, wl["frames"])
    b, _, _, _ = host_frames(wl, frames, wl["seed"])
    what = "full %s batch of one GPU" % name if frames == wl["frames"] else "%d frames of the batch" % frames
    return b.x, b.edge_index, b.edge_attr, "%s (%d edges) per step" % (what, b.edge_index.size(1))


def run_reference(args):
    """The reference arm: the oracle port on all host cores, rank 0 only, on a bounded sample of OUR arm's workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfgmod = importlib.import_module("gnn-cca_b200.config")
    from oracle import mpn_oracle  # the reference arm, cpu_baseline and parity are the only places bench.py runs oracle/
    name = args.workload or ("cfg2" if args.gpus == 1 else "cfg5")
    wl = WORKLOADS[name]
    params = cfgmod.graph_net_params()
    sd = mpn_oracle.init_state_dict(mpn_oracle.MPNSpec(params, "resnet50"), seed=0)
    if wl["kind"] == "train":
        b, _, _, _ = host_frames(wl, wl["frames"], wl["seed"])
        times, cores = cpu_train_timer(params, sd, b.x, b.edge_index, b.edge_attr, b.edge_labels, 0.0, steps=args.steps)
        E, sample = int(b.edge_index.size(1)), "full batch per step, oracle forward + torch autograd + SGD"
    else:
        x, ei, ea, sample = cpu_sample(name, args.gpus)
        times, cores, _ = cpu_forward_timer(params, "resnet50", sd, x, ei, ea, 0.0, args.warmup, steps=args.steps)
        E = int(ei.size(1))
    total = sum(times)
    val = E * len(times) / total
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": len(times),
            "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times), "higher_is_better": True, "scaling": wl["scaling"],
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(name, args.gpus)},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": "%s; oracle/mpn_oracle.py (torch CPU fp32, %d threads)" % (sample, cores)},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


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


# ------------------------------------------------------------------------------------------------------------
# frames workloads (cfg 1, 2, 4, 4s): MOTMPNet.forward over a Batch of per-frame graphs
# ------------------------------------------------------------------------------------------------------------
def make_frame_batch(ctx: Ctx, wl, g0: int, g1: int, variant: int = 0):
    """The frames [g0, g1) of a workload as one device Batch.  Generated in blocks of GEN_BLOCK frames, block b seeded with
    seed + b (+ 7919 per variant): every rank count sees the same per-frame inputs.  The embeddings are column-normalised over the
    rank's batch (train.py:257-259: the normalisation couples the graphs of ONE batch, here the rank's)."""
    torch, syn, gnncca = ctx.torch, ctx.syn, ctx.gnncca
    n = wl["cams"] * wl["dets"]
    parts = []
    b0 = g0 // GEN_BLOCK
    b1 = (g1 + GEN_BLOCK - 1) // GEN_BLOCK
    N = (g1 - g0) * n
    node = torch.empty((N, 2048), dtype=torch.float32, device=ctx.dev)
    reid = torch.empty((N, 256), dtype=torch.float32, device=ctx.dev)
    xw = torch.empty(N, dtype=torch.float64, device=ctx.dev)
    yw = torch.empty(N, dtype=torch.float64, device=ctx.dev)
    at = 0
    for b in range(b0, b1):
        det = syn.make_detections(wl["cams"], wl["dets"], GEN_BLOCK, seed=wl["seed"] + b + 7919 * variant, device=ctx.dev, max_dist=wl["max_dist"])
        lo, hi = max(g0, b * GEN_BLOCK) - b * GEN_BLOCK, min(g1, (b + 1) * GEN_BLOCK) - b * GEN_BLOCK
        rows = slice(lo * n, hi * n)
        cnt = (hi - lo) * n
        node[at:at + cnt] = det["node_embeds"][rows]
        reid[at:at + cnt] = det["reid_embeds"][rows]
        xw[at:at + cnt] = det["xw"][rows]
        yw[at:at + cnt] = det["yw"][rows]
        at += cnt
        del det
    k = torch.arange(n, device=ctx.dev)
    B = g1 - g0
    batch = gnncca.build_graph_batch(graph_ptr=torch.arange(0, N + 1, n, device=ctx.dev, dtype=torch.int64),
                                     id_cam=(k // wl["dets"]).repeat(B).to(torch.int32), ids=(k % wl["dets"]).repeat(B).to(torch.int64),
                                     xw=xw, yw=yw, max_dist=torch.full((B,), wl["max_dist"], dtype=torch.float64, device=ctx.dev),
                                     reid_embeds=reid, node_embeds=node)
    del node, reid, parts
    batch.edge_labels = None          # not an input of the forward (3.8 GB at cfg 4)
    return batch


def frames_parity(ctx: Ctx, model, batch, outs, ref_full=None, sample_frames=None):
    """The timed path's logits against the CPU oracle: the whole batch when ``ref_full`` (the oracle's output for it) is given,
    otherwise the frames in ``sample_frames``, each evaluated alone by the oracle (frames are independent given the batch's
    normalised embeddings)."""
    torch = ctx.torch
    from oracle import mpn_oracle
    tol = 1e-3 if model.math == "tf32" else 1e-4
    if ref_full is not None:
        got = [o.detach().cpu() for o in outs]
        err = rel_err(got, ref_full)
        ok, near, n = decisions_equal(got, ref_full, margin=tol * (1 + max(float(r.abs().max()) for r in ref_full)))
        return {"against": "oracle/mpn_oracle.py (CPU fp32) on the whole batch of the timed path", "max_rel_err": err, "err_is": ERR_FORMULA, "tolerance": tol,
                "within_tolerance": err <= tol, "decisions_equal": ok, "edges_checked": n, "edges_inside_margin": near}
    spec = mpn_oracle.MPNSpec(ctx.params, "resnet50")
    sd = {k: v.detach().cpu() for k, v in model.state_dict().items()}
    ptr = batch.ptr[sample_frames + [f + 1 for f in sample_frames]].tolist()
    eptr = batch.edge_ptr[sample_frames + [f + 1 for f in sample_frames]].tolist()
    worst, ok_all, n_all, near_all = 0.0, True, 0, 0
    S = len(sample_frames)
    with torch.no_grad():
        for i in range(S):
            n0, n1, e0, e1 = ptr[i], ptr[S + i], eptr[i], eptr[S + i]
            ref = mpn_oracle.motmpnet_forward(spec, sd, x_rows(batch, n0, n1).cpu(), (batch.edge_index[:, e0:e1] - n0).cpu(), batch.edge_attr[e0:e1].cpu())["classified_edges"]
            got = [o[e0:e1].detach().cpu() for o in outs]
            worst = max(worst, rel_err(got, ref))
            ok, near, n = decisions_equal(got, ref, margin=tol * (1 + max(float(r.abs().max()) for r in ref)))
            ok_all, n_all, near_all = ok_all and ok, n_all + n, near_all + near
    return {"against": "oracle/mpn_oracle.py (CPU fp32) on frames %s of this rank's batch, each evaluated alone" % sample_frames,
            "max_rel_err": worst, "err_is": ERR_FORMULA, "tolerance": tol, "within_tolerance": worst <= tol, "decisions_equal": ok_all, "edges_checked": n_all,
            "edges_inside_margin": near_all}


def run_frames(ctx: Ctx, name: str, steps: int, warmup: int, headline: bool):
    torch, gnncca, engine, args = ctx.torch, ctx.gnncca, ctx.engine, ctx.args
    wl = WORKLOADS[name]
    strong = wl["scaling"] == "strong"
    if strong:
        g0, g1 = ctx.mg.frame_shard(wl["frames"], ctx.rank, ctx.world)
    else:      # weak: every rank has its own batch of the named size, different frames per rank
        g0, g1 = ctx.rank * wl["frames"], (ctx.rank + 1) * wl["frames"]
    model = ctx.model()
    # inputs larger than L2 (126 MB): two batches used alternately where one batch is smaller than that
    n_per, e_per = wl["cams"] * wl["dets"], ctx.syn.edges_per_graph(wl["cams"], wl["dets"])
    one_batch_bytes = (g1 - g0) * (n_per * 8192 + e_per * 32)
    n_batches = 1 if one_batch_bytes > (512 << 20) else max(2, min(16, -(-(300 << 20) // one_batch_bytes)))
    batches = [make_frame_batch(ctx, wl, g0, g1, variant=v) for v in range(n_batches)]
    N, E, F = batches[0].num_nodes, batches[0].num_edges, int(x_in(batches[0]).size(1))
    E_total = E * ctx.world if not strong else wl["frames"] * e_per
    N_total = N * ctx.world if not strong else wl["frames"] * n_per
    in_bytes = N * F * 4 + batches[0].edge_index.numel() * 8 + batches[0].edge_attr.numel() * 4
    keep = {}

    def step(i):
        keep["out"] = model(batches[i % n_batches])["classified_edges"]     # the call a user makes (models/mpn.py:250)

    with torch.no_grad():
        sampler = ClockSampler(physical_gpu_index(ctx.local)) if headline else None
        if sampler:
            for i in range(3):
                step(i)
            ctx.barrier()
            sampler.start()
        tw = timed_windows(ctx, step, steps, warmup, args.window_seconds)
        if sampler:      # a representative amount of load for the sampler when the windows were short (same count on every rank)
            for i in range(tail_steps(tw["ms_per_step"])):
                step(i)
            torch.cuda.synchronize()
        clocks = sampler.finish() if sampler else None
        path = getattr(model.last_plan, "last_path", None)
        # logits of batch 0 from the timed path, for the parity check
        outs0 = [o.clone() for o in model(batches[0])["classified_edges"]]
        torch.cuda.synchronize()

        # ---------------- per-kernel device times, live
        stage_avg = {}
        if path == "fused" and model.math == "tf32":
            stage_avg = fused_kernel_times(ctx, model, batches, steps)
            kb = batches[0]
        else:
            # per-step pipeline: CUDA-event marks between the stages of ONE block-sized forward (blocks of a larger batch repeat it)
            kb = batches[0]
            if path == "blocked":
                blocks = kb._gnncca_blocks[1]
                gb0, gb1, n0, n1, e0, e1 = blocks[0]
                kb = gnncca.Batch(x=x_in(kb)[n0:n1], edge_index=(kb.edge_index[:, e0:e1] - n0).contiguous(), edge_attr=kb.edge_attr[e0:e1],
                                  ptr=(kb.ptr[gb0:gb1 + 1] - n0).contiguous(), num_graphs=gb1 - gb0)
            profs = []
            for i in range(min(steps, 10) + 2):
                pr = engine.StageTimer()
                engine.forward_inference(model, x_in(kb), kb.edge_index, kb.edge_attr, prof=pr, graph_ptr=kb.ptr, trusted=True, _no_block=True)
                pr.mark("end")
                profs.append(pr)
            torch.cuda.synchronize()
            acc = {}
            for pr in profs[2:]:
                for nm, tt in pr.stages():
                    acc.setdefault(nm, []).append(tt)
            stage_avg = {k: sum(v) / len(v) for k, v in acc.items() if k not in ("begin", "end", "encoder_edge")}
    ms_step = tw["ms_per_step"]
    value = E_total / (ms_step * 1e-3)
    kN, kE = kb.num_nodes, kb.num_edges
    per_kernel, dom = kernel_table(stage_avg, kN, kE, F, ctx.peak, load_traffic(name), sum(stage_avg.values()) if path == "blocked" else ms_step)
    fwd_b, b_min = forward_bytes(N, E, F), compulsory_bytes(N, E, F)
    t_rank = ms_step * 1e-3
    roofline = {"bound": "hbm", "kernel": dom, "peak": ctx.peak, "unit": "GB/s", "peak_source": ctx.peak_src,
                "kernel_selection": "largest share of the step's device time", "kernels": per_kernel,
                "kernel_scope": ("one block of %d edges of the frame-blocked forward" % kE) if path == "blocked" else "the whole batch of one rank",
                "forward": {"contract_bytes": fwd_b, "contract_frac": min(fwd_b / t_rank / 1e9 / ctx.peak, 1.0),
                            "compulsory_bytes": b_min, "achieved": b_min / t_rank / 1e9, "frac": b_min / t_rank / 1e9 / ctx.peak,
                            "note": "frac = compulsory bytes of the fused design (x, edge_index, edge_attr once, logits once) over the whole-forward time; "
                                    "contract = SURVEY 8d B_fwd (per-step traffic of an unfused pipeline), capped at 1"}}
    if dom:
        roofline.update(achieved=per_kernel[dom]["achieved"], frac=per_kernel[dom]["frac"], traffic=per_kernel[dom]["traffic"],
                        algorithmic_bytes_per_launch=per_kernel[dom]["bytes"], kernel_ms=per_kernel[dom]["ms"])

    # ---------------- parity + cpu baseline (the same oracle run serves both on the small configs)
    cpu = parity = None
    small = E <= 2_000_000
    if ctx.rank == 0:
        if small:
            sd = {k: v.detach().cpu() for k, v in model.state_dict().items()}
            b = batches[0]
            budget = args.cpu_seconds if (ctx.world == 1 and not args.no_cpu_baseline) else 0.0
            times, cores, ref = cpu_forward_timer(ctx.params, "resnet50", sd, b.x.cpu(), b.edge_index.cpu(), b.edge_attr.cpu(), budget, 1,
                                                  steps=None if budget > 0 else 1)
            parity = frames_parity(ctx, model, b, outs0, ref_full=ref)
            if budget > 0:
                cpu = {"value": E / statistics.median(times), "unit": UNIT, "cores": cores, "kind": "port",
                       "sample": "%d forwards of the full %s batch (E=%d) through oracle/mpn_oracle.py, torch CPU fp32, median" % (len(times), name, E)}
        else:
            B = batches[0].num_graphs
            parity = frames_parity(ctx, model, batches[0], outs0, sample_frames=sorted({0, B // 2, B - 1}))
            if ctx.world == 1 and not args.no_cpu_baseline:
                from oracle import mpn_oracle
                x, ei, ea, sample = cpu_sample(name)
                sd = {k: v.detach().cpu() for k, v in model.state_dict().items()}
                times, cores, _ = cpu_forward_timer(ctx.params, "resnet50", sd, x, ei, ea, args.cpu_seconds, 1)
                cpu = {"value": int(ei.size(1)) / statistics.median(times), "unit": UNIT, "cores": cores, "kind": "port",
                       "sample": "%d forwards of %s, oracle/mpn_oracle.py, torch CPU fp32, median" % (len(times), sample)}
    entry = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": ctx.world, "steps": steps, "warmup": max(warmup, 3),
             "ms_per_step": ms_step, "windows": tw["windows"], "higher_is_better": True, "scaling": wl["scaling"], "vs_baseline": None,
             "dtype": "tf32" if model.math == "tf32" else "f32", "data": "synthetic",
             "config": {"workload": workload_name(name, ctx.world), "per_gpu": {"frames": g1 - g0, "nodes": N, "edges": E},
                        "parallelism": "graph-sharded x%d, no collective" % ctx.world, "device_path": path,
                        "l2": ("%d input batches used alternately (%d x %.0f MB > 126 MB L2)" % (n_batches, n_batches, in_bytes / 1e6)) if n_batches > 1
                              else "one batch of %.1f GB per rank (>> 126 MB L2)" % (in_bytes / 1e9),
                        "graphs_per_s": (wl["frames"] if strong else wl["frames"] * ctx.world) / (ms_step * 1e-3)},
             "gpu_launches": tw["launches"], "roofline": roofline, "parity": parity}
    if clocks is not None:
        entry["clocks"] = clocks
    if cpu is not None:
        entry["cpu_baseline"] = cpu
    if headline and in_bytes > 4e9:
        # a batch of tens of GB does not fit twice (pinned host copy + the pipeline's second device slot): no end-to-end leg here
        entry["e2e"] = {"skipped": "inputs of %.1f GB per rank: the host pipeline keeps two device slots; measured on cfg2" % (in_bytes / 1e9)}
    elif headline:
        entry["e2e"] = frames_e2e(ctx, model, batches, wl, E, min(steps, 20), strong, E_total)
        if args.math == "tf32" and small:      # the price of fp32-exact output, as a number
            m32 = ctx.model(math="fp32")

            def step32(i):
                keep["out32"] = m32(batches[i % n_batches])["classified_edges"]
            with torch.no_grad():
                tw32 = timed_windows(ctx, step32, max(3, steps // 4), 3, 0.5)
                o32 = [o.clone() for o in m32(batches[0])["classified_edges"]]
            entry["value_fp32"] = E_total / (tw32["ms_per_step"] * 1e-3)
            entry["ms_per_step_fp32"] = tw32["ms_per_step"]
            if ctx.rank == 0 and parity is not None:
                entry["parity_fp32"] = frames_parity(ctx, m32, batches[0], o32, ref_full=ref)
        if name in ("cfg1", "cfg2"):
            entry["graph_build"] = graph_build_entry(ctx, wl)
    del batches, model
    torch.cuda.empty_cache()
    return entry


def fused_kernel_times(ctx: Ctx, model, batches, steps: int):
    """Each of the forward's two kernels launched back to back through the C ABI on the alternating batches (CUDA events on
    the launching stream around ``steps`` launches; no host gaps in between)."""
    import ctypes as C
    torch, engine, lib = ctx.torch, ctx.engine, ctx.lib
    _lib = ctx.gnncca._lib
    fp = engine.fast_path(model)
    if fp is None:
        return {}
    N, E = batches[0].num_nodes, batches[0].num_edges
    nb = len(batches)
    stream = torch.cuda.current_stream().cuda_stream
    l1, l2 = model.encoder.node_mlp.linear(0), model.encoder.node_mlp.linear(1)
    le = model.encoder.edge_mlp.linear(0)
    hs = [torch.empty((N, 32), dtype=torch.float32, device=ctx.dev) for _ in batches]
    lg = torch.empty((model.num_class_steps, E), dtype=torch.float32, device=ctx.dev)
    caps = [engine.frame_capacity(model, N, E, b.num_graphs, b.gnncca_caps) for b in batches]

    def enc(i):
        b = batches[i % nb]
        _lib.check(lib.gnncca_node_encoder(x_in(b).data_ptr(), l1.weight.data_ptr(), l1.bias.data_ptr(), l2.weight.data_ptr(), l2.bias.data_ptr(),
                                           hs[i % nb].data_ptr(), N, l1.in_features, l1.out_features, 32, _lib.MATH_TF32, stream), "node_encoder")

    def mp(i):
        b, cap = batches[i % nb], caps[i % nb]
        _lib.check(lib.gnncca_mp_fused_fwd(C.byref(fp.sb.cfg), C.byref(fp.sb.w), le.weight.data_ptr(), le.bias.data_ptr(), fp.fa,
                                           b.ptr.data_ptr(), b.num_graphs, b.edge_index.data_ptr(), b.edge_ptr.data_ptr(), None, None, None,
                                           N, E, hs[i % nb].data_ptr(), b.edge_attr.data_ptr(), fp.L, fp.first, cap[0], cap[1], lg.data_ptr(),
                                           None, None, fp.flags.data_ptr(), stream), "mp_fused_fwd")

    out = {}
    for name, fn in (("encoder_node", enc), ("mp_fused", mp)):
        for i in range(4):
            fn(i)
        a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        for i in range(steps):
            fn(i)
        z.record()
        torch.cuda.synchronize()
        out[name] = a.elapsed_time(z) / steps
    return out


def frames_e2e(ctx: Ctx, model, batches, wl, E, e2e_steps, strong, E_total):
    """The same metric end to end through the public host API: pinned host x / edge_index / edge_attr in, pinned host logits
    out, copies inside the timed region (wall clock, max over ranks)."""
    torch, gnncca = ctx.torch, ctx.gnncca
    e2e_steps = max(3, e2e_steps)
    nb = len(batches)
    with torch.no_grad():
        hosts = [(b.x.cpu().pin_memory(), b.edge_index.cpu().pin_memory(), b.edge_attr.cpu().pin_memory()) for b in batches]
        ptrs = [b.ptr.cpu().pin_memory() for b in batches]
        k = model.num_class_steps
        outs_pair = [[torch.empty((E, 1), dtype=torch.float32).pin_memory() for _ in range(k)] for _ in range(2)]
        for i in range(2):
            model.forward_host(*hosts[i % nb], out=outs_pair[0], graph_ptr=ptrs[i % nb], caps=batches[i % nb].gnncca_caps)
        ctx.barrier()
        t0 = time.perf_counter()
        for i in range(e2e_steps):
            model.forward_host(*hosts[i % nb], out=outs_pair[0], graph_ptr=ptrs[i % nb], caps=batches[i % nb].gnncca_caps)     # returns after the D2H copy of the logits
        torch.cuda.synchronize()
        blocking_s = time.perf_counter() - t0
        # the same stream of batches through HostPipeline: batch i+1 is copied in while batch i computes and its logits travel back;
        # every step still moves its inputs host->device and its logits device->host, and every result is waited for inside the region
        pipe = gnncca.HostPipeline(model)
        for i in range(2):
            pipe.submit(*hosts[i % nb], out=outs_pair[i & 1], graph_ptr=ptrs[i % nb], caps=batches[i % nb].gnncca_caps).wait()
        ctx.barrier()
        t0 = time.perf_counter()
        pending = None
        for i in range(e2e_steps):
            hd = pipe.submit(*hosts[i % nb], out=outs_pair[i & 1], graph_ptr=ptrs[i % nb], caps=batches[i % nb].gnncca_caps)
            if pending is not None:
                pending.wait()
            pending = hd
        pending.wait()
        torch.cuda.synchronize()
        e2e_s = time.perf_counter() - t0
        # what the link alone allows: the step's x bytes, pinned host -> device, nothing else running
        x_probe = torch.empty_like(x_in(batches[0]))
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(3):
            x_probe.copy_(hosts[i % nb][0], non_blocking=True)
        torch.cuda.synchronize()
        h2d_gbs = 3 * x_probe.numel() * 4 / (time.perf_counter() - t0) / 1e9
        del x_probe, hosts
    e2e_max, blocking_max = ctx.max_over_ranks([e2e_s, blocking_s])
    in_bytes = x_in(batches[0]).numel() * 4 + batches[0].edge_index.numel() * 8 + batches[0].edge_attr.numel() * 4 + 8 * (batches[0].num_graphs + 1)
    return {"value": E_total * e2e_steps / e2e_max, "unit": UNIT, "h2d_bytes_per_step": in_bytes, "d2h_bytes_per_step": model.num_class_steps * E * 4,
            "steps": e2e_steps, "api": "gnncca.HostPipeline.submit / wait (pinned host x/edge_index/edge_attr -> pinned host logits, two batches in flight)",
            "h2d_link_gbs": h2d_gbs, "link_bound_value": E_total / (in_bytes / (h2d_gbs * 1e9)),
            "blocking_call_value": E_total * e2e_steps / blocking_max, "blocking_api": "MOTMPNet.forward_host (returns after the logits are on the host)"}


def graph_build_entry(ctx: Ctx, wl):
    """Graph + edge-attribute construction (train.py:257-361,450 -> build_graph_batch), timed separately: the call
    includes both column normalisations, the edge count pass, the tiled build kernel and its one small read-back.  Beside it the
    CPU restatement of the reference loop (oracle/graph_oracle.build_batch) on a bounded sample of the same frames."""
    torch, gnncca, syn = ctx.torch, ctx.gnncca, ctx.syn
    det = syn.make_detections(wl["cams"], wl["dets"], wl["frames"], seed=wl["seed"] + 100 * ctx.rank + 31, device=ctx.dev, max_dist=wl["max_dist"])
    for _ in range(3):
        b = gnncca.build_graph_batch(**det)
    torch.cuda.synchronize()
    N, E = b.num_nodes, b.num_edges
    reps, times = 50, []
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        gnncca.build_graph_batch(**det)
        torch.cuda.synchronize()
        times.append(time.perf_counter() - t0)
    gb_s = statistics.median(times)
    lazy = b.raw_x() is not None
    # bytes the construction really moves: node embeddings read once for the norms (+ the normalised copy written and re-read when it
    # is materialised), ReID embeddings read twice (norms; the fill, which normalises while staging and writes nothing back),
    # detection table, outputs (edge_index 16, edge_attr 16, labels 4 B/edge)
    gb_bytes = N * 2048 * 4 * (1 if lazy else 3) + N * 256 * 4 * 2 + N * 40 + E * (16 + 16 + 4)
    out = {"value": E / gb_s, "unit": "edges built/s", "ms": gb_s * 1e3, "ms_min": min(times) * 1e3,
           "what": "build_graph_batch: column norms of both embeddings + edge_index / edge_attr / labels / ptr of %d frames (wall clock, median of %d, incl. the totals the host waits for)" % (wl["frames"], reps),
           "roofline": {"bound": "hbm", "bytes": gb_bytes, "achieved": gb_bytes / gb_s / 1e9, "peak": ctx.peak, "frac": gb_bytes / gb_s / 1e9 / ctx.peak, "unit": "GB/s"}}
    del det, b
    if ctx.rank == 0 and ctx.world == 1 and not ctx.args.no_cpu_baseline:
        from oracle import graph_oracle
        nf = min(wl["frames"], 64)
        _, frames, reid, node = host_frames(wl, nf, wl["seed"] + 31)
        t0 = time.perf_counter()
        hb = graph_oracle.build_batch(frames, reid, node, [wl["max_dist"]] * nf)
        dt = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": int(hb.edge_index.size(1)) / dt, "unit": "edges built/s", "cores": 1, "kind": "port",
                               "sample": "%d frames (%d edges) through oracle/graph_oracle.build_batch (the reference's per-frame Python loop, train.py:288-361), one pass"
                                         % (nf, hb.edge_index.size(1))}
    return out


# ------------------------------------------------------------------------------------------------------------
# cfg 5: one giant graph, edge-partitioned, NCCL all-gather of the node features per step
# ------------------------------------------------------------------------------------------------------------
def giant_parity(ctx: Ctx, model, wl, x, outs_local, n0, n1, sample_rows: int = 96):
    """(1) this rank's logits from the timed partitioned path against the UNPARTITIONED forward on this GPU, bit for bit;
    (2) rank 0: the unpartitioned forward against the CPU oracle in float64, step by step on sampled source rows -- the inputs
    of step t are the device's outputs of step t-1, so every step's edge update, full-degree segmented reduction and classifier
    is checked against the reference's arithmetic at the size the number is quoted on (an end-to-end CPU evaluation of 62.9 M
    edges takes minutes; tests/test_gpu_giant.py does that on a 3.9 M-edge graph)."""
    torch, engine = ctx.torch, ctx.engine
    C, D = wl["cams"], wl["dets"]
    N, deg = C * D, C * D - D
    with torch.no_grad():
        _, ei_f, ea_f = giant_inputs(wl, (0, N), ctx.dev)
        want_lat = ctx.rank == 0
        res = engine.forward_inference(model, x, ei_f, ea_f, return_latents=want_lat)
        full = res[0] if want_lat else res
        same = all(bool(torch.equal(full[k][n0 * deg:n1 * deg], outs_local[k])) for k in range(len(full)))
    flag = ctx.min_over_ranks([1.0 if same else 0.0])[0] > 0.5
    out = {"bit_identical_to_single_gpu_forward": flag}
    if ctx.rank == 0:
        from oracle import mpn_oracle, mpn_oracle_chunked as moc
        latents = res[1]
        spec = mpn_oracle.MPNSpec(ctx.params, "resnet50")
        sd64 = moc.cast_state_dict({k: v.detach().cpu() for k, v in model.state_dict().items()}, torch.float64)
        g = torch.Generator().manual_seed(7)
        rows = torch.sort(torch.randperm(N, generator=g)[:sample_rows])[0]
        sel = (rows[:, None] * deg + torch.arange(deg)[None, :]).reshape(-1)
        sel_d = sel.to(ctx.dev)
        row, col = ei_f[0, sel_d].cpu(), ei_f[1, sel_d].cpu()
        L, first = model.num_enc_steps, model.num_enc_steps - model.num_class_steps + 1
        tol = 1e-3 if model.math == "tf32" else 1e-4
        worst = {"encoder": 0.0, "edge_latents": 0.0, "node_latents": 0.0, "logits": 0.0}
        per_step = {}
        with torch.no_grad():
            # encoders (all nodes; sampled edges)
            h_ref = mpn_oracle.mlp_forward(sd64, "encoder.node_mlp", spec.enc_node, x.cpu().double())
            worst["encoder"] = rel_err([latents[0][0].cpu()], [h_ref])
            e_ref = mpn_oracle.mlp_forward(sd64, "encoder.edge_mlp", spec.enc_edge, ea_f[sel_d].cpu().double())
            worst["edge_latents"] = rel_err([latents[0][1][sel_d].cpu()], [e_ref])
            h0 = latents[0][0].cpu().double()
            e0 = latents[0][1][sel_d].cpu().double()
            dec_ok, near, n_dec = True, 0, 0
            for t in range(1, L + 1):
                h_prev = latents[t - 1][0].cpu().double()
                e_prev = latents[t - 1][1][sel_d].cpu().double()
                e_new, h_new, lg = moc.motmpnet_step(spec, sd64, h_prev, h0, e_prev, e0, row, col, rows, classify=t >= first)
                ee, he = rel_err([latents[t][1][sel_d].cpu()], [e_new]), rel_err([latents[t][0][rows.to(ctx.dev)].cpu()], [h_new])
                per_step["step%d" % t] = {"edge_latents": ee, "node_latents": he, "max_abs_node_latent": float(h_new.abs().max())}
                worst["edge_latents"], worst["node_latents"] = max(worst["edge_latents"], ee), max(worst["node_latents"], he)
                if lg is not None:
                    got = full[t - first][sel_d].cpu()
                    per_step["step%d" % t]["logits"] = rel_err([got], [lg])
                    worst["logits"] = max(worst["logits"], per_step["step%d" % t]["logits"])
                    ok, nr, nn = decisions_equal([got], [lg], margin=tol * (1 + float(lg.abs().max())))
                    dec_ok, near, n_dec = dec_ok and ok, near + nr, n_dec + nn
        err = max(worst.values())
        out.update({"against": "oracle/mpn_oracle_chunked.motmpnet_step (CPU float64), every step of the forward on %d sampled source rows "
                               "(%d edges, degree %d); inputs of step t = device outputs of step t-1" % (sample_rows, sel.numel(), deg),
                    "max_rel_err": err, "err_is": ERR_FORMULA, "per_quantity": worst, "per_step": per_step, "tolerance": tol, "within_tolerance": err <= tol, "decisions_equal": dec_ok,
                    "edges_checked": n_dec, "edges_inside_margin": near})
        del latents, res
    del ei_f, ea_f, full
    torch.cuda.empty_cache()
    return out


def run_giant(ctx: Ctx, name: str, steps: int, warmup: int, headline: bool):
    torch, engine, mg, args = ctx.torch, ctx.engine, ctx.mg, ctx.args
    wl = WORKLOADS[name]
    C, D = wl["cams"], wl["dets"]
    N, deg = C * D, C * D - D
    E = N * deg
    model = ctx.model()
    rowptr = torch.arange(0, N * deg + 1, deg, dtype=torch.int64)
    ranges = mg.balanced_source_ranges(rowptr, ctx.world)
    n0, n1 = ranges[ctx.rank]
    x, ei, ea = giant_inputs(wl, (n0, n1), ctx.dev)
    E_loc = int(ei.size(1))
    runner = mg.PartitionedForward(model, x, ei, ea, ranges, edge_base=n0 * deg) if hasattr(mg, "PartitionedForward") else None
    keep = {}

    def step(i):
        if runner is not None:
            keep["out"] = runner()
        else:
            keep["out"] = mg.forward_edge_partitioned(model, x, ei, ea, ranges, edge_base=n0 * deg)

    with torch.no_grad():
        sampler = ClockSampler(physical_gpu_index(ctx.local)) if headline else None
        if sampler:
            for i in range(3):
                step(i)
            ctx.barrier()
            sampler.start()
        tw = timed_windows(ctx, step, steps, warmup, args.window_seconds, extra_launches=(lambda: runner.launches_replayed) if runner is not None else (lambda: 0))
        if sampler:
            # More load for the clock sampler.  The step holds NCCL collectives: every rank must run it the SAME number of times (a
            # loop bounded by each rank's own clock leaves the ranks that enqueued more forwards waiting in an all-gather for ever).
            # tw is the max over ranks, identical everywhere.
            for _ in range(tail_steps(tw["ms_per_step"])):
                step(0)
            torch.cuda.synchronize()
            ctx.barrier()
        clocks = sampler.finish() if sampler else None
        outs_local = [o.clone() for o in keep["out"]]
        mg.check_partition_flags(model)
        # ---------------- per-stage device times of the partitioned forward on this rank (CUDA-event marks) + host enqueue time
        profs, host_s = [], []
        for i in range(min(steps, 10) + 2):
            pr = engine.StageTimer()
            torch.cuda.synchronize()
            if ctx.world > 1:
                ctx.dist.barrier()
            t0 = time.perf_counter()
            mg.forward_edge_partitioned(model, x, ei, ea, ranges, edge_base=n0 * deg, prof=pr)
            host_s.append(time.perf_counter() - t0)
            pr.mark("end")
            profs.append(pr)
        torch.cuda.synchronize()
        acc = {}
        for pr in profs[2:]:
            for nm, tt in pr.stages():
                acc.setdefault(nm, []).append(tt)
        stage_local = {k: sum(v) / len(v) for k, v in acc.items() if k not in ("begin", "end")}
    names = sorted(stage_local)
    stage_max = dict(zip(names, ctx.max_over_ranks([stage_local[k] for k in names])))
    ms_step = tw["ms_per_step"]
    value = E / (ms_step * 1e-3)
    step_keys = {k: v for k, v in stage_max.items() if k.startswith("step") or k.startswith("project") or k in ("encoder_node", "csr")}
    per_kernel, dom = kernel_table(step_keys, N, E_loc, 2048, ctx.peak, load_traffic(name), ms_step)
    for k in per_kernel:       # the node-side terms of a partition: its own rows only write, all N rows are gathered
        per_kernel[k]["note"] = "max over ranks; bytes of this rank's partition"
    exch_ms = sum(v for k, v in stage_max.items() if k == "exchange")
    roofline = {"bound": "hbm", "kernel": dom, "peak": ctx.peak, "unit": "GB/s", "peak_source": ctx.peak_src,
                "kernel_selection": "largest share of the step's device time", "kernels": per_kernel,
                "stages_ms_max_over_ranks": {k: round(v, 5) for k, v in stage_max.items()},
                "stages_note": "stage times come from eager launches of the partitioned forward with the REPLICATED encoder (multigpu.forward_edge_partitioned); "
                               "the timed path replays one CUDA graph per rank and, for N > 1, encodes and projects the owned rows only plus one more exchange",
                "exchange": {"ms_per_forward": exch_ms, "bytes_per_rank_per_exchange": (n1 - n0) * 32, "exchanges_per_forward": model.num_enc_steps - 1,
                             "collective": "NCCL all_gather_into_tensor (in place) of the owned rows of the target projection Q (N x 8 fp32 assembled on every rank; the node latents themselves never travel)" if ctx.world > 1 else "none (1 GPU)"},
                "host_enqueue_ms_per_forward": statistics.median(host_s) * 1e3,
                "forward": {"contract_bytes": forward_bytes(N, E), "contract_frac": min(forward_bytes(N, E) / (ms_step * 1e-3) / 1e9 / (ctx.peak * ctx.world), 1.0),
                            "note": "SURVEY 8d B_fwd of the whole graph over the forward time and the AGGREGATE peak of the %d GPUs" % ctx.world}}
    if dom:
        roofline.update(achieved=per_kernel[dom]["achieved"], frac=per_kernel[dom]["frac"], traffic=per_kernel[dom]["traffic"],
                        algorithmic_bytes_per_launch=per_kernel[dom]["bytes"], kernel_ms=per_kernel[dom]["ms"])
    parity = giant_parity(ctx, model, wl, x, outs_local, n0, n1)
    entry = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": ctx.world, "steps": steps, "warmup": max(warmup, 3),
             "ms_per_step": ms_step, "windows": tw["windows"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
             "dtype": "tf32" if model.math == "tf32" else "f32", "data": "synthetic",
             "config": {"workload": workload_name(name, ctx.world), "per_gpu": {"source_rows": n1 - n0, "edges": E_loc},
                        "parallelism": "edge-partitioned by source range x%d; node features all-gathered after node updates 1-3" % ctx.world,
                        "device_path": "cuda-graph replay of the partitioned forward" if (runner is not None and runner.graphed) else "eager launches",
                        "l2": "per-rank edge slice %.1f GB (>> 126 MB L2)" % (E_loc * 32 / 1e9)},
             "gpu_launches": tw["launches"], "roofline": roofline, "parity": parity}
    if clocks is not None:
        entry["clocks"] = clocks
    if ctx.world > 1:
        # the same graph on ONE GPU of this box, in this run: the denominator of the strong-scaling ratio
        with torch.no_grad():
            if ctx.rank == 0:
                xf, eif, eaf = giant_inputs(wl, (0, N), ctx.dev)
                for _ in range(3):
                    engine.forward_inference(model, xf, eif, eaf)
                a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda.synchronize()
                a.record()
                for _ in range(5):
                    engine.forward_inference(model, xf, eif, eaf)
                z.record()
                torch.cuda.synchronize()
                entry["single_gpu"] = {"ms_per_step": a.elapsed_time(z) / 5, "value": E / (a.elapsed_time(z) / 5 * 1e-3),
                                       "what": "the whole graph on rank 0's GPU alone (engine.forward_inference), same run"}
                del xf, eif, eaf
            ctx.barrier()
    if headline:
        entry["e2e"] = giant_e2e(ctx, model, x, ei, ea, ranges, n0 * deg, E, min(steps, 10))
    if ctx.rank == 0 and ctx.world == 1 and not args.no_cpu_baseline:
        xs, eis, eas, sample = cpu_sample(name)
        sd = {k: v.detach().cpu() for k, v in model.state_dict().items()}
        times, cores, _ = cpu_forward_timer(ctx.params, "resnet50", sd, xs, eis, eas, args.cpu_seconds, 1)
        entry["cpu_baseline"] = {"value": int(eis.size(1)) / statistics.median(times), "unit": UNIT, "cores": cores, "kind": "port",
                                 "sample": "%d forwards of %s, oracle/mpn_oracle.py, torch CPU fp32, median" % (len(times), sample)}
    del x, ei, ea, model, runner
    torch.cuda.empty_cache()
    return entry


def giant_e2e(ctx: Ctx, model, x, ei, ea, ranges, edge_base, E, e2e_steps):
    torch, mg = ctx.torch, ctx.mg
    e2e_steps = max(3, e2e_steps)
    with torch.no_grad():
        hx, hei, hea = x.cpu().pin_memory(), ei.cpu().pin_memory(), ea.cpu().pin_memory()
        outs = [torch.empty((ei.size(1), 1), dtype=torch.float32).pin_memory() for _ in range(model.num_class_steps)]
        for _ in range(2):
            mg.forward_edge_partitioned_host(model, hx, hei, hea, ranges, out=outs, edge_base=edge_base)
        ctx.barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            mg.forward_edge_partitioned_host(model, hx, hei, hea, ranges, out=outs, edge_base=edge_base)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    dt_max = ctx.max_over_ranks([dt])[0]
    in_bytes = hx.numel() * 4 + hei.numel() * 8 + hea.numel() * 4
    return {"value": E * e2e_steps / dt_max, "unit": UNIT, "h2d_bytes_per_step": in_bytes, "d2h_bytes_per_step": model.num_class_steps * int(ei.size(1)) * 4,
            "steps": e2e_steps, "bytes_are": "per rank",
            "api": "gnncca.multigpu.forward_edge_partitioned_host (pinned host x / edge slice / attributes -> pinned host logits of the slice)"}


# ------------------------------------------------------------------------------------------------------------
# cfg 3: training step (forward + summed BCE over the classified steps + backward + SGD), train.py:454-494
# ------------------------------------------------------------------------------------------------------------
def run_training(ctx: Ctx, name: str, steps: int, warmup: int, headline: bool):
    torch, gnncca, syn, args = ctx.torch, ctx.gnncca, ctx.syn, ctx.args
    wl = WORKLOADS[name]
    model = ctx.model(train=True)
    opt = torch.optim.SGD(model.parameters(), lr=1e-3, momentum=0.9, weight_decay=1e-4, fused=True)
    batches = [gnncca.build_graph_batch(**syn.make_detections(wl["cams"], wl["dets"], wl["frames"], seed=wl["seed"] + 100 * ctx.rank + 7919 * s,
                                                               device=ctx.dev, max_dist=wl["max_dist"])) for s in range(8)]
    N, E = batches[0].num_nodes, batches[0].num_edges
    in_bytes = sum(t.numel() * t.element_size() for t in (batches[0].x, batches[0].edge_index, batches[0].edge_attr, batches[0].edge_labels))

    def step_eager(i):
        b = batches[i % 8]
        opt.zero_grad(set_to_none=True)
        out = model(b)["classified_edges"]
        loss = gnncca.loss_stats(out, b.edge_labels).loss      # train.py:80-97 in one pass (gnncca_bce_stats_fwd/bwd)
        loss.backward()
        opt.step()
        return loss

    gstep = gnncca.GraphedTrainStep(model)     # forward + loss + backward captured once per batch shape, replayed afterwards

    def step_b(b):
        loss = gstep(b)                        # zeroes .grad, runs the same kernels as step_eager, out of one CUDA graph
        opt.step()
        return loss

    def step(i):
        return step_b(batches[i % 8])

    tw_eager = timed_windows(ctx, step_eager, steps, warmup, 0.5)
    sampler = ClockSampler(physical_gpu_index(ctx.local)) if headline else None
    if sampler:
        for i in range(3):
            step(i)
        ctx.barrier()
        sampler.start()
    tw = timed_windows(ctx, step, steps, warmup, args.window_seconds, extra_launches=lambda: gstep.launches_replayed)
    if sampler:
        for _ in range(tail_steps(tw["ms_per_step"])):
            step(0)
        torch.cuda.synchronize()
    clocks = sampler.finish() if sampler else None
    # parity of the step the bench times: one graphed step from a fresh copy of the weights against the oracle + torch autograd
    parity = None
    if ctx.rank == 0:
        from oracle import mpn_oracle
        m2 = ctx.model(math="fp32", train=True)
        sd0 = {k: v.detach().cpu().clone() for k, v in m2.state_dict().items()}
        b = batches[0]
        for p in m2.parameters():
            p.grad = None
        out = m2(b)["classified_edges"]
        loss = gnncca.loss_stats(out, b.edge_labels).loss
        loss.backward()
        spec = mpn_oracle.MPNSpec(ctx.params, "resnet50")
        sdr = {k: v.clone().requires_grad_(v.is_floating_point()) for k, v in sd0.items()}
        ref = mpn_oracle.motmpnet_forward(spec, sdr, b.x.cpu(), b.edge_index.cpu(), b.edge_attr.cpu(), training=True)["classified_edges"]
        crit = torch.nn.BCEWithLogitsLoss()
        lref = sum(crit(r.view(-1), b.edge_labels.cpu()) for r in ref)
        lref.backward()
        gerr = 0.0
        for nme, p in m2.named_parameters():
            gr = sdr[nme].grad
            gerr = max(gerr, float(((p.grad.cpu() - gr).abs() / (1e-6 + gr.abs().max())).max()))
        parity = {"against": "oracle/mpn_oracle.py + torch autograd (CPU fp32): loss and every parameter gradient of one step on batch 0",
                  "loss_rel_err": abs(float(loss) - float(lref)) / abs(float(lref)), "max_grad_err_rel_to_max": gerr, "tolerance": 2e-4,
                  "within_tolerance": gerr <= 2e-4 and abs(float(loss) - float(lref)) / abs(float(lref)) <= 1e-4}
        del m2
    # end to end: inputs from pinned host memory every step, loss read back
    hosts = [[t.cpu().pin_memory() for t in (b.x, b.edge_index, b.edge_attr, b.edge_labels)] for b in batches[:2]]
    e2e_steps = max(3, min(steps, 20))
    sb = gstep.static_batch(batches[0])          # the graph's own input buffers: host data is copied straight into them
    ctx.barrier()
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        hx, hei, hea, hl = hosts[i & 1]
        sb.x.copy_(hx, non_blocking=True); sb.edge_index.copy_(hei, non_blocking=True)
        sb.edge_attr.copy_(hea, non_blocking=True); sb.edge_labels.copy_(hl, non_blocking=True)
        float(step_b(sb).item())
    e2e_s = time.perf_counter() - t0
    e2e_max = ctx.max_over_ranks([e2e_s])[0]
    ms_step = tw["ms_per_step"]
    fb = forward_bytes(N, E, int(batches[0].x.size(1)))
    train_bytes = fb + 4 * E * (24 + 24 + 24 + 32) + 3 * N * 8 * 2048          # + saved latents, gradient passes, dW1/x re-reads
    ach = train_bytes / (ms_step * 1e-3) / 1e9
    entry = {"metric": METRIC, "value": ctx.world * E / (ms_step * 1e-3), "unit": UNIT, "n_gpus": ctx.world, "steps": steps, "warmup": max(warmup, 3),
             "ms_per_step": ms_step, "windows": tw["windows"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
             "dtype": "tf32", "data": "synthetic",
             "config": {"workload": workload_name(name, ctx.world), "per_gpu": {"frames": wl["frames"], "nodes": N, "edges": E},
                        "l2": "8 batches rotate; a batch is 22 MB (fits L2: launch-bound regime)",
                        "step": "forward + loss + backward replayed from one CUDA graph (GraphedTrainStep), then torch SGD; the same step launched eagerly: %.3f ms"
                                % tw_eager["ms_per_step"],
                        "optimizer": "torch.optim.SGD(fused=True) lr 1e-3 momentum 0.9 wd 1e-4",
                        "parallelism": "graph-sharded x%d, no gradient all-reduce (independent replicas)" % ctx.world},
             "gpu_launches": tw["launches"],
             "e2e": {"value": ctx.world * E * e2e_steps / e2e_max, "unit": UNIT, "h2d_bytes_per_step": in_bytes, "d2h_bytes_per_step": 4, "steps": e2e_steps,
                     "api": "GraphedTrainStep (MOTMPNet.forward + loss_stats + backward) + SGD.step, batch copied from pinned host memory into the graph's input buffers"},
             "roofline": {"bound": "hbm", "kernel": "whole training step (launch-bound at this size)", "achieved": ach, "peak": ctx.peak, "unit": "GB/s",
                          "frac": ach / ctx.peak, "traffic": None, "algorithmic_bytes_per_launch": train_bytes, "peak_source": ctx.peak_src},
             "parity": parity}
    if clocks is not None:
        entry["clocks"] = clocks
    if ctx.rank == 0 and ctx.world == 1 and not args.no_cpu_baseline:
        sd = {k: v.detach().cpu() for k, v in model.state_dict().items()}
        b = batches[0]
        times, cores = cpu_train_timer(ctx.params, sd, b.x.cpu(), b.edge_index.cpu(), b.edge_attr.cpu(), b.edge_labels.cpu(), min(args.cpu_seconds, 8.0))
        entry["cpu_baseline"] = {"value": E / statistics.median(times), "unit": UNIT, "cores": cores, "kind": "port",
                                 "sample": "%d training steps of the full batch (oracle forward + torch autograd + SGD), median" % len(times)}
    del batches, model, gstep
    torch.cuda.empty_cache()
    return entry


RUNNERS = {"frames": run_frames, "giant": run_giant, "train": run_training}


def attach_steps(name: str, steps: int) -> int:
    """Steps per window of an ATTACHED workload (the line of record uses --steps as given)."""
    return {"cfg4": min(steps, 3), "cfg5": min(steps, 10), "cfg4s": min(steps, 10)}.get(name, min(steps, 20))


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
        return
    ctx = Ctx(args)
    head = args.workload or ("cfg2" if ctx.world == 1 else "cfg5")
    line = RUNNERS[WORKLOADS[head]["kind"]](ctx, head, args.steps, args.warmup, True)
    if args.workload is None and not args.no_attach:
        attached = {}
        for name in (("cfg1", "cfg3", "cfg4", "cfg5") if ctx.world == 1 else ("cfg4", "cfg2")):
            try:
                ent = RUNNERS[WORKLOADS[name]["kind"]](ctx, name, attach_steps(name, args.steps), 3, False)
            except Exception as exc:  # noqa: BLE001 -- an attached workload must not take the line of record down with it; the line says so
                import traceback
                ent = {"error": "%s: %s" % (type(exc).__name__, exc), "trace": traceback.format_exc(limit=4)}
                ctx.torch.cuda.empty_cache()
            attached[name] = ent
        line["workloads"] = attached
    if ctx.rank == 0:
        print(json.dumps(line), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
