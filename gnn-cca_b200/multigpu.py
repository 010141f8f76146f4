"""Multi-GPU execution of the hot path on one 8 x B200 box: one process per GPU, ``torch.distributed`` for plumbing.

The reference has no distributed code at all; the two partitionings here follow from the
structure of ``MOTMPNet.forward`` (models/mpn.py:250-299):

* **Frame sharding** (BASELINE configs 1-4).  A PyG ``Batch`` is a disjoint union of per-frame graphs and no
  operation of the forward mixes frames in eval mode, so rank r simply takes a contiguous block of frames.
  No collective on the data path; results are bit-identical to the single-GPU run frame by frame.

* **Edge partition of one giant graph** (config 5).  Edges are source-major, the aggregation index is the source
  (``scatter(..., row)``, mpn.py:99), so rank r owns a contiguous range of SOURCE nodes with all their out-edges:
  the segmented reduction is rank-local and the only exchange is the updated node features (N x 32 fp32, 1 MB at
  N = 8192) after each node update, because the next edge update reads ``x[col]`` for arbitrary targets.  Every
  node's segment lives on one rank, so the summation order -- and therefore every bit of the result -- is the
  same as on one GPU.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

from . import _lib
from .graph import Batch


# ------------------------------------------------------------------------------------------------------------
# frame sharding
# ------------------------------------------------------------------------------------------------------------
def frame_shard(num_frames: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block [g0, g1) of frames for ``rank``: sizes differ by at most one, earlier ranks get the extras."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("rank %d outside world of %d" % (rank, world))
    base, extra = divmod(num_frames, world)
    g0 = rank * base + min(rank, extra)
    return g0, g0 + base + (1 if rank < extra else 0)


def shard_batch(batch: Batch, rank: int, world: int) -> Batch:
    """The sub-batch of ``rank``: its frames' nodes and edges with ``edge_index`` rebased to the first owned node.
    Needs ``batch.ptr`` (frame node offsets) and source-major edges; ``batch.edge_ptr`` is used when present,
    otherwise the edge offsets are found by a binary search in ``edge_index[0]``."""
    ptr = batch.ptr
    B = int(ptr.numel()) - 1
    g0, g1 = frame_shard(B, rank, world)
    sub_ptr = ptr[g0:g1 + 1]
    n0, n1 = int(sub_ptr[0]), int(sub_ptr[-1])
    eptr = getattr(batch, "edge_ptr", None)
    if eptr is not None:
        e0, e1 = int(eptr[g0]), int(eptr[g1])
    else:
        row = batch.edge_index[0].contiguous()
        b = torch.searchsorted(row, torch.tensor([n0, n1], device=row.device, dtype=row.dtype))
        e0, e1 = int(b[0]), int(b[1])
    raw = batch.raw_x() if hasattr(batch, "raw_x") else None
    kw = dict(edge_index=batch.edge_index[:, e0:e1] - n0, edge_attr=batch.edge_attr[e0:e1], ptr=sub_ptr - n0, num_graphs=g1 - g0)
    if raw is not None:      # lazily normalised batch: the shard keeps the raw rows and the norms of the WHOLE batch (train.py:257-259)
        kw.update(x_raw=raw[0][n0:n1], x_norm=raw[1])
    else:
        kw["x"] = batch.x[n0:n1]
    for name in ("edge_labels",):
        if getattr(batch, name, None) is not None:
            kw[name] = getattr(batch, name)[e0:e1]
    for name in ("y", "batch"):
        if getattr(batch, name, None) is not None:
            v = getattr(batch, name)[n0:n1]
            kw[name] = v - g0 if name == "batch" else v
    if eptr is not None:
        kw["edge_ptr"] = eptr[g0:g1 + 1] - e0
    if getattr(batch, "gnncca_caps", None) is not None:
        kw["gnncca_caps"] = batch.gnncca_caps
    if getattr(batch, "gnncca_trusted", False):
        kw["gnncca_trusted"] = True
    return Batch(**kw)


# ------------------------------------------------------------------------------------------------------------
# one giant graph: partition the edges by source-node range
# ------------------------------------------------------------------------------------------------------------
def balanced_source_ranges(rowptr: torch.Tensor, world: int) -> List[Tuple[int, int]]:
    """Cut the node range into ``world`` contiguous pieces with (nearly) equal numbers of out-edges: piece r ends at
    the first node whose cumulative edge count reaches E (r+1) / world.  Pure index arithmetic, bit-exact."""
    rp = rowptr.detach().to("cpu", torch.int64)
    N, E = int(rp.numel()) - 1, int(rp[-1])
    cuts = [0]
    for r in range(1, world):
        target = (E * r + world - 1) // world
        c = int(torch.searchsorted(rp, torch.tensor(target, dtype=torch.int64), right=False))
        cuts.append(min(max(c, cuts[-1]), N))
    cuts.append(N)
    return [(cuts[r], cuts[r + 1]) for r in range(world)]


def exchange_node_rows(h_partial: torch.Tensor, ranges: Sequence[Tuple[int, int]], group=None) -> torch.Tensor:
    """Assemble the full node-feature matrix from per-rank row ranges.  ``h_partial`` is [N, F] with this rank's rows
    valid and all others ZERO.  Equal-sized ranges use one all-gather of the owned rows; ragged ranges one
    all-reduce (x + 0 is exact, so both give the same bits)."""
    if not dist.is_available() or not dist.is_initialized():
        return h_partial
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    if world == 1:
        return h_partial
    sizes = {b - a for a, b in ranges}
    if len(sizes) == 1 and ranges[0][0] == 0 and all(ranges[i][1] == ranges[i + 1][0] for i in range(world - 1)):
        out = torch.empty_like(h_partial)
        a, b = ranges[rank]
        dist.all_gather_into_tensor(out, h_partial[a:b].contiguous(), group=group)
        return out
    out = h_partial.clone()
    dist.all_reduce(out, op=dist.ReduceOp.SUM, group=group)
    return out


def forward_edge_partitioned(model, x: torch.Tensor, edge_index_local: torch.Tensor, edge_attr_local: torch.Tensor,
                             ranges: Sequence[Tuple[int, int]], group=None, edge_base: int = 0, prof=None) -> List[torch.Tensor]:
    """Drive ``edge_partitioned_steps`` with the NCCL exchange of ``exchange_node_rows`` (see there).  ``edge_base``:
    global id of this rank's first edge (the sum of the lower ranks' edge counts).  ``prof``: an ``engine.StageTimer``
    that receives a CUDA-event mark after every stage (csr, encoder, project, step of each step, exchange)."""
    gen = edge_partitioned_steps(model, x, edge_index_local, edge_attr_local, edge_base, prof=prof, owned=ranges[_rank(group)] if _live(group) else None)
    try:
        q_partial = next(gen)
        while True:
            q_full = exchange_node_rows(q_partial, ranges, group)
            if prof is not None:
                prof.mark("exchange")
            q_partial = gen.send(q_full)
    except StopIteration as done:
        return done.value


def _live(group=None) -> bool:
    return dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1


def _rank(group=None) -> int:
    return dist.get_rank(group)


def forward_edge_partitioned_host(model, x: torch.Tensor, edge_index_local: torch.Tensor, edge_attr_local: torch.Tensor,
                                  ranges: Sequence[Tuple[int, int]], out: Optional[List[torch.Tensor]] = None, group=None,
                                  edge_base: int = 0, device: Optional[torch.device] = None) -> List[torch.Tensor]:
    """``forward_edge_partitioned`` for HOST buffers (pinned memory recommended): this rank's node matrix, edge slice and
    attributes are copied to the device, the partitioned forward runs, and the rank's logits are copied back into ``out``
    (host tensors [E_r, 1]; allocated pinned when None).  Returns after the copies have completed."""
    if x.is_cuda or edge_index_local.is_cuda or edge_attr_local.is_cuda:
        raise ValueError("forward_edge_partitioned_host takes host tensors")
    dev = device if device is not None else next(model.parameters()).device
    if dev.type != "cuda":
        raise RuntimeError("gnn-cca_b200: the model must live on a CUDA device (no CPU path)")
    x_d = x.to(dev, non_blocking=True)
    ei_d = edge_index_local.to(dev, non_blocking=True)
    ea_d = edge_attr_local.to(dev, non_blocking=True)
    outs = forward_edge_partitioned(model, x_d, ei_d, ea_d, ranges, group, edge_base)
    if out is None:
        out = [torch.empty(o.shape, dtype=torch.float32, pin_memory=True) for o in outs]
    for o, d in zip(out, outs):
        o.copy_(d, non_blocking=True)
    torch.cuda.current_stream(dev).synchronize()
    return out


def edge_partitioned_steps(model, x: torch.Tensor, edge_index_local: torch.Tensor, edge_attr_local: torch.Tensor, edge_base: int = 0,
                           prof=None, owned: Optional[Tuple[int, int]] = None, buffers: Optional[dict] = None, fold_owned_only: bool = False,
                           shard_encoder: bool = False):
    """Generator form of the edge-partitioned forward: yields this rank's partial TARGET-PROJECTION matrix ``Q`` ([N, 8]: the
    rows of the nodes it owns are valid, all others zero) at every exchange point and expects the assembled matrix back
    through ``send``; returns the rank's logits.  (Tests drive several of these in lock step on one GPU.)

    ``MOTMPNet.forward`` for ONE graph whose edges are partitioned by source-node range across the ranks of
    ``group``.  ``x`` [N, F] is the full node-feature matrix (replicated: the encoder is cheap next to E),
    ``edge_index_local`` [2, E_r] holds GLOBAL node ids of the edges whose source lies in the rank's range
    (source-major), ``edge_attr_local`` their attributes.  Returns this rank's slice of every classified step's
    logits ([E_r, 1] each), identical to the corresponding rows of the single-GPU result.

    What crosses the link.  The edge update of step t+1 reads, of the OTHER ranks' nodes, only ``Q[dst] = W_e[:, dst block] h[dst]``
    (6 floats, padded to 8); ``P[src]`` and ``R[src]`` are needed for the rank's own sources alone.  The fold launch of step t projects
    the freshly aggregated rows of the owned nodes (gnncca_mp_step_fwd_next), so the exchange carries N x 8 floats per step instead
    of the N x 32 node latents, and no projection launch runs between the steps.  ``owned``: this rank's node range; needed (to
    zero the other rows of Q) only with reattach_initial_nodes, where a non-owned row projects the replicated initial latents.
    ``fold_owned_only``: the second launch of every step visits the owned nodes only (the rows of Q outside ``owned`` are then NOT
    written: for exchanges that overwrite them, i.e. the all-gather of equal ranges).
    ``shard_encoder``: the node encoder and the first projection run on the OWNED rows only (the replicated encoder is the largest
    cost that does not shrink with the rank count) and one more exchange of ``Q`` precedes step 1; rows are independent in both
    kernels, so the owned rows carry the same bits as in the replicated run."""
    from . import engine
    from .plan import GraphPlan, _stream
    lib = _lib.load()
    engine.check_supported(model)
    engine._require_cuda(x, "x")
    engine._require_cuda(edge_attr_local, "edge_attr")
    if torch.is_grad_enabled() and any(p.requires_grad for p in model.parameters()):
        raise NotImplementedError("the edge-partitioned path is inference only (use torch.no_grad())")
    dev = x.device
    N, E = int(x.size(0)), int(edge_index_local.size(1))
    math = engine.MATH_MODES[model.math]
    if prof is not None:
        prof.mark("begin")
    # rowptr spans all N nodes; segments outside the owned range are empty.  The order / range flags are NOT read back here (no
    # host synchronisation on the data path): the step kernels take them as their guard and are no-ops on an invalid CSR, and
    # ``check_partition_flags(model)`` reads them whenever the caller wants the verdict.
    plan = GraphPlan(edge_index_local, N)
    guard = _lib.ptr(plan.flags)
    if prof is not None:
        prof.mark("csr")
    buf = buffers if buffers is not None else {}

    def get(name, shape):
        t = buf.get(name)
        if t is None or tuple(t.shape) != tuple(shape):
            t = torch.empty(shape, dtype=torch.float32, device=dev)
            if buffers is not None:
                buffers[name] = t
        return t

    shard = bool(shard_encoder and owned is not None)
    if shard:
        na, nb = int(owned[0]), int(owned[1])
        h = get("h0", (N, 32))                      # rows [na, nb) only
        if nb > na:
            engine.mlp_forward_nograd(model.encoder.node_mlp, x[na:nb], math, out=h[na:nb])
    else:
        h = engine.mlp_forward_nograd(model.encoder.node_mlp, x, math, out=get("h0", (N, 32)))
    if prof is not None:
        prof.mark("encoder_node")
    enc_e = model.encoder.edge_mlp
    ea = edge_attr_local.contiguous()
    L = model.num_enc_steps
    # edge encoder folded into the step kernel when it is the shipped Linear(4, f_edge) + ReLU (gnncca_mp_step_fwd_attr)
    fold = (L > 0 and len(enc_e.layer_specs) == 1 and enc_e.layer_specs[0].relu and enc_e.layer_specs[0].bn is None
            and int(ea.size(1)) == 4 and ea.data_ptr() % 16 == 0)
    e = None if fold else engine.small_mlp_rows(enc_e, ea)
    le = enc_e.linear(0) if fold else None
    h0, e0 = h, e
    sb = engine.StepBinding(model, math)
    sb.cfg.edge_phase = int(edge_base) % int(lib.gnncca_mp_step_tile())      # tile the local slice like the whole graph: bit-identical reductions
    if fold_owned_only and owned is not None:      # the fold launch (aggregate + next projections) visits the owned sources only: the other
        sb.cfg.node_begin, sb.cfg.node_end = int(owned[0]), int(owned[1])     # rows of Q arrive through the exchange, nobody reads their h
    k = model.num_class_steps
    first = L - k + 1
    outs: List[torch.Tensor] = []
    if L == 0:
        return [engine.small_mlp_rows_cls(model, e)]
    P, Q, R = get("P", (N, 8)), get("Q", (N, 8)), get("R", (N, sb.f_node))
    ws_bytes = int(lib.gnncca_mp_step_workspace_bytes(C.byref(sb.cfg), N, E))
    ws = buf.get("ws")
    if ws is None or ws.numel() < ws_bytes:
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        if buffers is not None:
            buffers["ws"] = ws
    h_out = get("h_out", (N, sb.f_node))
    if shard:
        if not fold_owned_only:
            Q.zero_()                               # exchanges that SUM the partial matrices need the other rows at zero
        if nb > na:
            _lib.check(lib.gnncca_node_project(C.byref(sb.cfg), C.byref(sb.w), _lib.ptr(h0[na:nb]), _lib.ptr(h[na:nb]), nb - na, _lib.ptr(P[na:nb]),
                                               _lib.ptr(Q[na:nb]), _lib.ptr(R[na:nb]), _stream()), "node_project")
        if prof is not None:
            prof.mark("project1")
        q_full = yield Q                            # the targets' projections of step 1 from their owners
        if q_full.data_ptr() != Q.data_ptr():
            Q.copy_(q_full)
    else:
        _lib.check(lib.gnncca_node_project(C.byref(sb.cfg), C.byref(sb.w), _lib.ptr(h0), _lib.ptr(h), N, _lib.ptr(P), _lib.ptr(Q), _lib.ptr(R),
                                           _stream()), "node_project")
        if prof is not None:
            prof.mark("project1")
    for step in range(1, L + 1):
        last = step == L
        logits = get("logits%d" % step, (E, 1)) if step >= first else None
        dead_e = last and logits is not None                 # e^L is never read
        use_attr = fold and (step == 1 or model.reattach_initial_edges)
        if use_attr:
            e_out = None if dead_e else (get("e", (E, sb.f_edge)) if step == 1 else e)
        else:
            e_out = None if dead_e else (get("e", (E, sb.f_edge)) if (model.reattach_initial_edges and step == 1) else e)
        if not last:
            _lib.check(lib.gnncca_mp_step_fwd_next(C.byref(sb.cfg), C.byref(sb.w), _lib.ptr(plan.rowptr), _lib.ptr(plan.src), _lib.ptr(plan.dst), N, E,
                                                   _lib.ptr(P), _lib.ptr(Q), _lib.ptr(R), _lib.ptr(ea) if use_attr else 0, 4,
                                                   _lib.ptr(le.weight) if use_attr else 0, _lib.ptr(le.bias) if use_attr else 0,
                                                   _lib.ptr(e0) if (model.reattach_initial_edges and not use_attr) else 0, _lib.ptr(e), _lib.ptr(e_out),
                                                   _lib.ptr(h_out), _lib.ptr(logits), _lib.ptr(h0), _lib.ptr(P), _lib.ptr(Q), _lib.ptr(R), guard,
                                                   _lib.ptr(ws), ws_bytes, _stream()), "mp_step_fwd_next")
        elif use_attr:
            _lib.check(lib.gnncca_mp_step_fwd_attr(C.byref(sb.cfg), C.byref(sb.w), _lib.ptr(plan.rowptr), _lib.ptr(plan.src), _lib.ptr(plan.dst), N, E,
                                                   _lib.ptr(P), _lib.ptr(Q), _lib.ptr(R), _lib.ptr(ea), 4, _lib.ptr(le.weight), _lib.ptr(le.bias),
                                                   _lib.ptr(e), _lib.ptr(e_out), 0, _lib.ptr(logits), guard, _lib.ptr(ws), ws_bytes,
                                                   _stream()), "mp_step_fwd_attr")
        else:
            _lib.check(lib.gnncca_mp_step_fwd(C.byref(sb.cfg), C.byref(sb.w), _lib.ptr(plan.rowptr), _lib.ptr(plan.src), _lib.ptr(plan.dst), N, E,
                                              _lib.ptr(P), _lib.ptr(Q), _lib.ptr(R), _lib.ptr(e0) if model.reattach_initial_edges else 0, _lib.ptr(e),
                                              _lib.ptr(e_out), 0, _lib.ptr(logits), guard, _lib.ptr(ws), ws_bytes, _stream()), "mp_step_fwd")
        if prof is not None:
            prof.mark("step%d" % step)
        e = e_out
        if logits is not None:
            outs.append(logits)
        if not last:
            if model.reattach_initial_nodes and owned is not None:      # non-owned rows project the replicated h0: not this rank's to contribute
                a, b = owned
                Q[:a].zero_()
                Q[b:].zero_()
            q_full = yield Q                                 # the one exchange of the step: N x 8 fp32 over NVLink
            if q_full.data_ptr() != Q.data_ptr():
                Q.copy_(q_full)
    model.__dict__["_gnncca_partition_plan"] = plan        # for check_partition_flags
    return outs


def check_partition_flags(model) -> None:
    """Read back the CSR flags of the model's last edge-partitioned forward (one 16-byte copy, synchronising) and raise when
    the local edge slice was not source-major or held ids outside [0, N): in that case every step kernel of that forward was
    a no-op and its outputs are undefined."""
    plan = model.__dict__.get("_gnncca_partition_plan")
    if plan is None:
        return
    if not plan.source_major():
        raise ValueError("the local edge slice must be source-major")


# On by default for live process groups with equal ranges: equal bit for bit to the replicated form on 4 virtual ranks
# (tests/test_gpu_giant.py) and under NCCL at 2 GPUs (profiles/r02e_cfg5_2gpu_sharded_encoder.json: 2.177 ms vs 2.183 ms, parity object
# identical); the 8-GPU gain (encoder 35 us -> ~9 us + one 11 us exchange per forward) is an estimate from the per-stage table, not yet
# a measurement.  GNNCCA_SHARD_ENCODER=0 restores the replicated encoder.
SHARD_ENCODER = os.environ.get("GNNCCA_SHARD_ENCODER", "1") != "0"


class PartitionedForward:
    """The edge-partitioned forward of ONE rank as a replayable object: all work buffers are allocated once, the launches of one
    forward -- CSR narrowing, node encoder, L steps, and the NCCL all-gathers of the projected rows between them -- are captured
    into ONE CUDA graph on the first call and replayed afterwards (the eager path costs ~20 Python-driven launches and
    allocations per forward, which at 8 GPUs is a third of a 0.8 ms forward).  Inputs are read from the tensors given at
    construction (write new values into them in place); the returned logits are the graph's own buffers, valid until the next call.

        runner = PartitionedForward(model, x, edge_index_local, edge_attr_local, ranges, edge_base=first_edge)
        logits = runner()            # list of [E_r, 1]
    """

    def __init__(self, model, x, edge_index_local, edge_attr_local, ranges, group=None, edge_base: int = 0, use_graph: Optional[bool] = None):
        self.model, self.x, self.ei, self.ea = model, x, edge_index_local, edge_attr_local
        self.ranges, self.group, self.edge_base = list(ranges), group, edge_base
        self.live = _live(group)
        self.rank = _rank(group) if self.live else 0
        sizes = {b - a for a, b in self.ranges}
        self.equal = len(sizes) == 1 and self.ranges[0][0] == 0 and all(self.ranges[i][1] == self.ranges[i + 1][0] for i in range(len(self.ranges) - 1))
        self.buffers: dict = {}
        self.graph = None
        self.outs = None
        self.launches_replayed = 0
        self._launches_per_graph = 0
        if use_graph is None:
            use_graph = os.environ.get("GNNCCA_PARTITION_GRAPH", "1") != "0"
        self.use_graph = use_graph
        self.graphed = False
        self._calls = 0

    def _exchange(self, q: torch.Tensor) -> torch.Tensor:
        if not self.live:
            return q
        a, b = self.ranges[self.rank]
        if self.equal:      # in place: every rank's rows land at their own offset of the same [N, 8] matrix
            dist.all_gather_into_tensor(q, q[a:b], group=self.group)
            return q
        zero = self.buffers.get("qz")
        if zero is None or zero.shape != q.shape:
            zero = self.buffers["qz"] = torch.empty_like(q)
        zero.zero_()
        zero[a:b].copy_(q[a:b])
        dist.all_reduce(zero, op=dist.ReduceOp.SUM, group=self.group)       # x + 0 is exact: the same bits as a gather
        return zero

    def _run(self):
        gen = edge_partitioned_steps(self.model, self.x, self.ei, self.ea, self.edge_base, owned=self.ranges[self.rank], buffers=self.buffers,
                                     fold_owned_only=self.live and self.equal, shard_encoder=self.live and self.equal and SHARD_ENCODER)
        try:
            q = next(gen)
            while True:
                q = gen.send(self._exchange(q))
        except StopIteration as done:
            return done.value

    def __call__(self) -> List[torch.Tensor]:
        self._calls += 1
        if not self.use_graph or self._calls <= 2:      # two eager calls first: buffers allocated, NCCL communicators warmed up
            self.outs = self._run()
            return self.outs
        if self.graph is None:
            lib = _lib.load()
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            n0 = int(lib.gnncca_launch_count())
            with torch.cuda.graph(g):
                self.outs = self._run()
            self._launches_per_graph = int(lib.gnncca_launch_count()) - n0
            self.graph, self.graphed = g, True
        self.graph.replay()
        self.launches_replayed += self._launches_per_graph
        return self.outs


def complete_multipartite_slice(n_cams: int, n_dets: int, node_range: Tuple[int, int], device) -> torch.Tensor:
    """edge_index [2, E_r] (global ids, source-major) of the cross-camera complete graph on n_cams x n_dets
    camera-major detections, restricted to sources in ``node_range`` -- the enumeration rule of train.py:292-299 in
    closed form, generated on the device so that no rank ever holds the whole edge list."""
    a, b = node_range
    n = n_cams * n_dets
    src = torch.arange(a, b, device=device, dtype=torch.int64)
    cam = src // n_dets
    deg = n - n_dets
    k = torch.arange(deg, device=device, dtype=torch.int64)
    # targets of a source in camera c: every node id except [c * D, (c + 1) * D), ascending
    dst = k[None, :] + (k[None, :] >= (cam * n_dets)[:, None]).to(torch.int64) * n_dets
    return torch.stack([src[:, None].expand(-1, deg).reshape(-1), dst.reshape(-1)])
