"""Host-side logic of the multi-GPU paths on CPU: sharding arithmetic, batch slicing, the closed-form giant-graph
enumeration, and the node-row exchange over a world_size-2 ``gloo`` group."""
import importlib
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

mg = importlib.import_module("gnn-cca_b200.multigpu")
graph = importlib.import_module("gnn-cca_b200.graph")
from oracle import graph_oracle  # noqa: E402


def test_frame_shard_partitions_exactly():
    for B in (0, 1, 7, 64, 1024, 4096):
        for world in (1, 2, 3, 4, 8):
            blocks = [mg.frame_shard(B, r, world) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == B
            assert all(blocks[i][1] == blocks[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in blocks]
            assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)
    with pytest.raises(ValueError):
        mg.frame_shard(8, 2, 2)


def _cpu_batch(sizes):
    rng = np.random.default_rng(0)
    frames = []
    for n_cam_counts in sizes:
        cam = np.repeat(np.arange(len(n_cam_counts)), n_cam_counts)
        frames.append(dict(id=rng.integers(0, 5, cam.size), id_cam=cam, xw=rng.random(cam.size), yw=rng.random(cam.size)))
    N = sum(len(f["id"]) for f in frames)
    b = graph_oracle.build_batch(frames, torch.rand(N, 16), torch.rand(N, 8), [50.0] * len(frames))
    return graph.Batch(x=b.x, edge_index=b.edge_index, edge_attr=b.edge_attr, edge_labels=b.edge_labels, y=b.y, batch=b.batch, ptr=b.ptr,
                       num_graphs=len(frames)), frames


def test_shard_batch_reassembles_bit_exactly():
    batch, frames = _cpu_batch([(2, 3), (1, 1), (4, 2, 3), (2, 2), (3, 1), (1, 2, 1), (5, 5)])
    for world in (1, 2, 3):
        parts = [mg.shard_batch(batch, r, world) for r in range(world)]
        off = 0
        xs, eis, eas, labs = [], [], [], []
        for p in parts:
            assert int(p.ptr[0]) == 0 and int(p.ptr[-1]) == p.x.size(0)
            if p.edge_index.numel():
                assert int(p.edge_index.min()) >= 0 and int(p.edge_index.max()) < p.x.size(0)
            xs.append(p.x); eis.append(p.edge_index + off); eas.append(p.edge_attr); labs.append(p.edge_labels)
            off += p.x.size(0)
        assert torch.equal(torch.cat(xs), batch.x) and torch.equal(torch.cat(eis, dim=1), batch.edge_index)
        assert torch.equal(torch.cat(eas), batch.edge_attr) and torch.equal(torch.cat(labs), batch.edge_labels)


def test_balanced_source_ranges():
    rowptr = torch.tensor([0, 10, 10, 25, 40, 41, 80, 100], dtype=torch.int32)
    for world in (1, 2, 4, 7):
        r = mg.balanced_source_ranges(rowptr, world)
        assert r[0][0] == 0 and r[-1][1] == 7 and all(r[i][1] == r[i + 1][0] for i in range(world - 1))
    uniform = torch.arange(0, 8192 * 7680 + 1, 7680, dtype=torch.int64)          # config 5: degree 7680 for every node
    assert mg.balanced_source_ranges(uniform, 8) == [(1024 * i, 1024 * (i + 1)) for i in range(8)]


def test_complete_multipartite_slice_equals_the_reference_enumeration():
    C, Dd = 3, 4
    cam = np.repeat(np.arange(C), Dd)
    full = graph_oracle.enumerate_edges(cam, np.arange(C * Dd))                   # train.py:292-299
    got = torch.cat([mg.complete_multipartite_slice(C, Dd, r, "cpu") for r in [(0, 5), (5, 6), (6, 12)]], dim=1)
    assert got.dtype == torch.int64 and torch.equal(got, full)


def _exchange_worker(rank, world, port, ragged, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        N, F = 10, 4
        ranges = [(0, 3), (3, 10)] if ragged else [(0, 5), (5, 10)]
        full = torch.arange(N * F, dtype=torch.float32).view(N, F) + 0.25
        part = torch.zeros(N, F)
        a, b = ranges[rank]
        part[a:b] = full[a:b]
        out = mg.exchange_node_rows(part, ranges)
        q.put((rank, bool(torch.equal(out, full))))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("ragged", [False, True])
def test_exchange_node_rows_gloo_world2(ragged):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 500) + (50 if ragged else 0)
    procs = [ctx.Process(target=_exchange_worker, args=(r, 2, port, ragged, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res == [(0, True), (1, True)]
