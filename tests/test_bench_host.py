"""Host-side pieces of bench.py and of the chunked oracle that need no GPU."""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from oracle import mpn_oracle as mo, mpn_oracle_chunked as moc  # noqa: E402

cfgmod = importlib.import_module("gnn-cca_b200.config")
mg = importlib.import_module("gnn-cca_b200.multigpu")
engine = importlib.import_module("gnn-cca_b200.engine")


def test_workload_strings_depend_on_workload_and_gpu_count_only():
    for name in bench.WORKLOADS:
        for world in (1, 2, 8):
            a, b = bench.workload_name(name, world), bench.workload_name(name, world)
            assert a == b and a.startswith(name + ":")
    assert "E=442368 " in bench.workload_name("cfg2", 1) and "E=3538944 " in bench.workload_name("cfg2", 8)      # weak: per-GPU batch x N
    assert "E=939524096 " in bench.workload_name("cfg4", 1) and "E=939524096 " in bench.workload_name("cfg4", 8)  # strong: fixed
    assert "E=62914560 " in bench.workload_name("cfg5", 4)


def test_stage_bytes_never_charge_what_a_stage_does_not_move():
    N, E = 32768, 14680064
    full = bench.stage_bytes("step2", N, E, 2048)
    assert bench.stage_bytes("step4", N, E, 2048) < full          # classify only: e^L is not written, no node update
    assert bench.stage_bytes("step1", N, E, 2048) < full          # raw attributes (16 B) instead of latents (24 B) in
    assert full == E * (24 + 24 + 8 + 4) + N * 320


def test_chunked_oracle_equals_oracle():
    """Same functions, same per-segment edge order: bit-identical in the same dtype, for every aggregation and with reattach."""
    for agg, reatt in (("sum", False), ("mean", True), ("max", False)):
        params = cfgmod.graph_net_params(classifier_batchnorm=True)
        params["node_agg_fn"] = agg
        params["reattach_initial_edges"] = params["reattach_initial_nodes"] = reatt
        spec = mo.MPNSpec(params, "bdnet_market")
        sd = mo.init_state_dict(spec, 0)
        C, D = 5, 7
        N = C * D
        ei = mg.complete_multipartite_slice(C, D, (0, N), "cpu")
        g = torch.Generator().manual_seed(1)
        x, ea = torch.relu(torch.randn(N, 512, generator=g)), torch.rand(ei.size(1), 4, generator=g)
        with torch.no_grad():
            a, lat = mo.motmpnet_forward(spec, sd, x, ei, ea, return_latents=True)
            b = moc.motmpnet_forward_chunked(spec, sd, x, ei, ea, max_edges=100)["classified_edges"]
            assert all(torch.equal(p, q) for p, q in zip(a["classified_edges"], b))
            # one step at a time on a subset of the source rows, inputs = the previous step's latents
            rows, deg = torch.tensor([3, 4, 11]), N - D
            sel = torch.cat([torch.arange(r * deg, (r + 1) * deg) for r in rows.tolist()])
            for t in range(1, 5):
                e_new, h_new, lg = moc.motmpnet_step(spec, sd, lat[t - 1][0], lat[0][0], lat[t - 1][1][sel], lat[0][1][sel], ei[0, sel], ei[1, sel],
                                                     rows, classify=t >= 2)
                assert torch.allclose(e_new, lat[t][1][sel], rtol=1e-6, atol=1e-6) and torch.allclose(h_new, lat[t][0][rows], rtol=1e-6, atol=1e-6)
                if lg is not None:
                    assert torch.allclose(lg, a["classified_edges"][t - 2][sel], rtol=1e-5, atol=1e-6)


def test_frame_blocks_cover_the_batch():
    import numpy as np
    for sizes, target in (([5, 0, 9, 9, 1], 10), ([100] * 64, 1600), ([7], 3), ([3, 3, 3], 100)):
        ep = np.concatenate([[0], np.cumsum(np.array(sizes) * 10)]).tolist()
        gp = np.concatenate([[0], np.cumsum(sizes)]).tolist()
        blocks = engine.frame_blocks(gp, ep, target)
        assert blocks[0][0] == 0 and blocks[-1][1] == len(sizes)
        assert all(blocks[i][1] == blocks[i + 1][0] for i in range(len(blocks) - 1))
        assert all(b[1] > b[0] for b in blocks)
        assert all((b[2], b[3], b[4], b[5]) == (gp[b[0]], gp[b[1]], ep[b[0]], ep[b[1]]) for b in blocks)


def test_tail_steps_is_a_count_every_rank_agrees_on():
    """The extra load for the clock sampler is a NUMBER of steps derived from the rank-agreed step time: a loop bounded by each
    rank's own clock around steps that hold NCCL collectives leaves some ranks waiting for peers (the 8-GPU stall of round 2)."""
    import inspect
    assert bench.tail_steps(0.715) == int(0.25 * 1e3 / 0.715)
    assert bench.tail_steps(1e9) == 3 and bench.tail_steps(0.0) == 5000
    for fn in (bench.run_giant, bench.run_frames, bench.run_training):
        src = inspect.getsource(fn)
        assert "while time.perf_counter()" not in src and "tail_steps(" in src
