"""Generate tests/golden/*.npz by RUNNING THE REFERENCE ITSELF (TEST INFRASTRUCTURE ONLY).

Run in the build container only (needs /root/reference):  ``python -m oracle.gen_golden``

What is produced
----------------
``mpn_<case>.npz``   The unmodified ``/root/reference/models/mpn.py`` ``MOTMPNet`` (random init, fixed
                     seed) run on small ragged cross-camera graphs: weights (state_dict), inputs,
                     every classified step's logits; for training cases also the loss of
                     ``train.py:80-97`` (sum of BCE-with-logits over the classified steps) and every
                     parameter gradient, and the BatchNorm buffers after the step.
``graph_<case>.npz`` The reference's own graph-construction loop, i.e. the body of
                     ``inference.validate_GNN_cross_camera_association`` (``inference.py:172-283``),
                     executed UNMODIFIED on CPU: ``torch.Tensor.cuda`` is made the identity, a stub CNN
                     returns prescribed embeddings, and a capturing stand-in for the MPN records the
                     ``Batch`` it is handed (x, edge_index, edge_attr, edge_labels, y, ptr) and stops
                     the loop.  Inputs are real ground-truth frames of EPFL-Terrace / EPFL-Laboratory
                     (parsed by the rules of ``libs/datasets.py:89-134``) and synthetic frames.

The vectors pin ``oracle/mpn_oracle.py`` and ``oracle/graph_oracle.py`` (tests/test_oracle_golden.py);
the CUDA path is then compared with those oracles and with the vectors directly.
"""
from __future__ import annotations

import copy
import json
import os
import sys

import numpy as np
import torch
import yaml

from . import graph_oracle, ref_shim

GOLDEN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
REF = ref_shim.REFERENCE_ROOT


# ----------------------------------------------------------------------------------------------
# small ragged cross-camera graphs
# ----------------------------------------------------------------------------------------------
def ragged_frames(rng: np.random.Generator, n_frames: int, n_cams: int, max_ids: int, p_keep: float = 0.6):
    """Frames with >= 2 cameras, camera-major row order like the reference's DataFrames."""
    frames = []
    while len(frames) < n_frames:
        cam, ids, xw, yw = [], [], [], []
        centers = rng.uniform([0, 0], [360, 288], size=(max_ids, 2))
        for c in range(n_cams):
            for i in range(max_ids):
                if rng.random() < p_keep:
                    cam.append(c); ids.append(i)
                    p = centers[i] + rng.normal(0, 5, size=2)
                    xw.append(p[0]); yw.append(p[1])
        if len(set(cam)) < 2:
            continue
        frames.append(dict(id=np.asarray(ids, np.int64), id_cam=np.asarray(cam, np.int64),
                           xw=np.asarray(xw, np.float64), yw=np.asarray(yw, np.float64),
                           frame=np.full(len(ids), len(frames), np.int64)))
    return frames


def batch_edge_index(frames):
    parts, first = [], 0
    for fr in frames:
        n = len(fr["id"])
        parts.append(graph_oracle.enumerate_edges(fr["id_cam"], np.arange(first, first + n)))
        first += n
    return torch.cat(parts, dim=1), first


# ----------------------------------------------------------------------------------------------
# MPN cases
# ----------------------------------------------------------------------------------------------
def load_params(which: str) -> dict:
    with open(os.path.join(REF, "config", which)) as f:
        return yaml.safe_load(f)["GRAPH_NET_PARAMS"]


def run_mpn_case(name, params, arch, seed, frames_kw, train=False, shuffle_edges=False, isolated=False,
                 randomize_bn=False, edge_in_dim=None):
    mpn = ref_shim.import_reference("models.mpn")
    params = copy.deepcopy(params)
    if edge_in_dim is not None:
        params["encoder_feats_dict"]["edges"]["edge_in_dim"] = edge_in_dim
    ctor_params = copy.deepcopy(params)
    torch.manual_seed(seed)
    model = mpn.MOTMPNet(copy.deepcopy(params), None, arch)
    rng = np.random.default_rng(seed + 100)
    frames = ragged_frames(rng, **frames_kw)
    edge_index, n_nodes = batch_edge_index(frames)
    if isolated:  # one extra node that is nobody's source and nobody's target
        n_nodes += 1
    if shuffle_edges:
        perm = torch.from_numpy(rng.permutation(edge_index.size(1)))
        edge_index = edge_index[:, perm].contiguous()
    f_x = params["encoder_feats_dict"]["nodes"][arch]["node_in_dim"]
    f_a = params["encoder_feats_dict"]["edges"]["edge_in_dim"]
    g = torch.Generator().manual_seed(seed + 200)
    x = torch.relu(torch.randn(n_nodes, f_x, generator=g))
    x = x / x.norm(dim=0).clamp_min(1e-12)
    edge_attr = torch.rand(edge_index.size(1), f_a, generator=g)
    labels = (torch.rand(edge_index.size(1), generator=g) < 0.2).float()
    if randomize_bn:
        for m in model.modules():
            if isinstance(m, torch.nn.BatchNorm1d):
                m.running_mean.copy_(torch.randn(m.num_features, generator=g) * 0.3)
                m.running_var.copy_(torch.rand(m.num_features, generator=g) + 0.5)
                m.weight.data.copy_(torch.rand(m.num_features, generator=g) + 0.5)
                m.bias.data.copy_(torch.randn(m.num_features, generator=g) * 0.2)
    sd0 = {k: v.detach().clone() for k, v in model.state_dict().items()}
    data = graph_oracle.Data(x=x, edge_index=edge_index, edge_attr=edge_attr)
    out = dict(config=json.dumps(ctor_params), arch=arch, train=int(train),
               x=x.numpy(), edge_index=edge_index.numpy(), edge_attr=edge_attr.numpy(), labels=labels.numpy())
    for k, v in sd0.items():
        out["sd/" + k] = v.numpy()
    if train:
        model.train()
        res = model(data)["classified_edges"]
        crit = torch.nn.BCEWithLogitsLoss()
        loss = sum(crit(r.view(-1), labels) for r in res)  # train.py:80-97
        loss.backward()
        out["loss"] = np.asarray(loss.item(), np.float64)
        for k, p in model.named_parameters():
            out["grad/" + k] = p.grad.numpy()
        for k, v in model.state_dict().items():
            if "running_" in k or "num_batches" in k:
                out["sd_after/" + k] = v.detach().numpy()
    else:
        model.eval()
        with torch.no_grad():
            res = model(data)["classified_edges"]
    for i, r in enumerate(res):
        out["logits/%d" % i] = r.detach().numpy()
    out["n_steps"] = np.asarray(len(res))
    np.savez_compressed(os.path.join(GOLDEN, "mpn_%s.npz" % name), **out)
    print("mpn_%s: N=%d E=%d steps=%d" % (name, n_nodes, edge_index.size(1), len(res)))


def gen_mpn():
    tr = load_params("config_training.yaml")      # no BN anywhere
    inf = load_params("config_inference.yaml")    # classifier with BatchNorm
    small = dict(n_frames=3, n_cams=4, max_ids=5)
    run_mpn_case("shipped_resnet50_eval", tr, "resnet50", 0, dict(n_frames=2, n_cams=4, max_ids=6))
    run_mpn_case("shipped_bdnet_train", tr, "bdnet_market", 1, small, train=True)
    run_mpn_case("bncls_eval", inf, "bdnet_market", 2, small, randomize_bn=True)
    run_mpn_case("bncls_train", inf, "bdnet_market", 3, small, train=True, randomize_bn=True)
    p = copy.deepcopy(tr); p["node_agg_fn"] = "mean"
    run_mpn_case("mean_isolated_eval", p, "osnet_market", 4, small, isolated=True)
    run_mpn_case("mean_train", p, "osnet_market", 14, small, train=True)
    p = copy.deepcopy(tr); p["node_agg_fn"] = "max"
    run_mpn_case("max_isolated_eval", p, "osnet_market", 5, small, isolated=True)
    run_mpn_case("max_train", p, "osnet_market", 15, small, train=True)
    p = copy.deepcopy(tr); p["reattach_initial_nodes"] = True; p["reattach_initial_edges"] = True
    run_mpn_case("reattach_eval", p, "bdnet_cuhk", 6, small)
    run_mpn_case("reattach_train", p, "bdnet_cuhk", 16, small, train=True)
    run_mpn_case("shuffled_eval", tr, "bdnet_market", 7, small, shuffle_edges=True, isolated=True)
    run_mpn_case("shuffled_train", tr, "bdnet_market", 17, small, shuffle_edges=True, train=True)
    run_mpn_case("onlydist_eval", tr, "bdnet_market", 8, small, edge_in_dim=2)  # main_training.py:153-158
    p = copy.deepcopy(tr); p["num_enc_steps"] = 2; p["num_class_steps"] = 2
    run_mpn_case("l2k2_eval", p, "bdnet_market", 9, small)
    p = copy.deepcopy(tr); p["num_enc_steps"] = 0; p["num_class_steps"] = 0
    run_mpn_case("l0_eval", p, "bdnet_market", 10, small)


# ----------------------------------------------------------------------------------------------
# graph-construction cases: drive the reference's own loop
# ----------------------------------------------------------------------------------------------
COLS = ("id", "xmin", "ymin", "xmax", "ymax", "frame", "lost", "occluded", "generated", "label")  # datasets.py:21


def parse_gt(dataset: str):
    """libs/datasets.py:53-134 restated (the class itself cannot run on current pandas/numpy)."""
    import pandas as pd
    root = os.path.join(REF, "datasets", dataset)
    cams = sorted(c for c in os.listdir(root) if c[0] != ".")
    parts = []
    for c in cams:
        df = pd.read_csv(os.path.join(root, c, "gt", "gt.txt"), header=None, sep=" ")
        df = df[df.columns[:len(COLS)]]
        df.columns = COLS
        df = df[df["lost"] == 0].copy()
        df["xmin"] = df["xmin"].values + 1
        df["width"] = (df["xmax"] - df["xmin"]).values
        df["height"] = (df["ymax"] - df["ymin"]).values
        df["id_cam"] = int(c[-1:])
        H = np.asarray(pd.read_csv(os.path.join(root, c, "Homography.txt"), header=None, sep="\t"))
        bx = df["xmin"].values + np.round(df["width"].values / 2)
        by = df["ymax"].values
        xw, yw = graph_oracle.homography_image_to_world(bx, by, H)
        df["xw"], df["yw"] = xw, yw
        parts.append(df)
    det = pd.concat(parts)  # == repeated DataFrame.append (datasets.py:122): camera-major row order
    det["frame"] = det["frame"].values.astype(int)
    return det


def valid_frame(det, f) -> bool:  # datasets.py:126-134
    sel = det[det["frame"].values == f]
    if len(np.unique(sel["id_cam"].values)) > 1:
        return bool(np.max(np.bincount(sel["id"].values)) > 1)
    return False


class _Captured(Exception):
    pass


def run_reference_loop(frames_df, max_dist, reid, node, mode):
    """Execute inference.py:172-283 unmodified; return the Batch handed to the MPN."""
    inference = ref_shim.import_reference("inference")
    captured = {}

    class CNN(torch.nn.Module):
        def forward(self, _):
            return node.clone(), reid.clone()

    class Capture(torch.nn.Module):
        def forward(self, batch):
            captured["batch"] = batch
            raise _Captured()

    n_total = sum(len(d) for d in frames_df)
    bboxes = [torch.zeros(len(d), 1) for d in frames_df]
    loader = [[bboxes, [d.copy() for d in frames_df], list(max_dist)]]
    cfg = {"CNN_MODEL": {"arch": "resnet50"},
           "TRAINING": {"ONLY_APPEARANCE": mode == "only_appearance", "ONLY_DIST": mode == "only_dist"}}
    assert n_total == reid.size(0)
    orig_cuda = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self  # the loop hard-codes .cuda(); identity on this CPU box
    try:
        inference.validate_GNN_cross_camera_association(cfg, loader, CNN(), Capture())
    except _Captured:
        pass
    finally:
        torch.Tensor.cuda = orig_cuda
    return captured["batch"]


def save_graph_case(name, frames_df, max_dist, seed, mode="full", reid_dim=256, node_dim=64):
    g = torch.Generator().manual_seed(seed)
    n_total = sum(len(d) for d in frames_df)
    reid = torch.relu(torch.randn(n_total, reid_dim, generator=g))
    node = torch.relu(torch.randn(n_total, node_dim, generator=g))
    b = run_reference_loop(frames_df, max_dist, reid, node, mode)
    out = dict(mode=mode, max_dist=np.asarray(max_dist, np.float64), reid=reid.numpy(), node=node.numpy(),
               graph_ptr=np.cumsum([0] + [len(d) for d in frames_df]).astype(np.int64),
               id=np.concatenate([d["id"].values for d in frames_df]).astype(np.int64),
               id_cam=np.concatenate([d["id_cam"].values for d in frames_df]).astype(np.int64),
               xw=np.concatenate([d["xw"].values for d in frames_df]).astype(np.float64),
               yw=np.concatenate([d["yw"].values for d in frames_df]).astype(np.float64),
               out_x=b.x.numpy(), out_edge_index=b.edge_index.numpy(), out_edge_attr=b.edge_attr.numpy(),
               out_edge_labels=b.edge_labels.numpy(), out_y=b.y.numpy(), out_ptr=b.ptr.numpy(),
               out_batch=b.batch.numpy())
    assert out["out_edge_index"].dtype == np.int64 and out["out_edge_attr"].dtype == np.float32
    np.savez_compressed(os.path.join(GOLDEN, "graph_%s.npz" % name), **out)
    print("graph_%s: B=%d N=%d E=%d mode=%s" % (name, len(frames_df), n_total, b.edge_index.size(1), mode))


def gen_graph():
    import pandas as pd
    for dataset, md, picks in (("EPFL-Terrace", 50.83, None), ("EPFL-Laboratory", 44.23, None)):
        det = parse_gt(dataset)
        fmin, fmax = int(det["frame"].min()), int(det["frame"].max())
        valid = [f for f in range(fmin, fmax + 1, 97) if valid_frame(det, f)]
        # a spread of sizes: smallest, largest and a few in between
        sizes = sorted(valid, key=lambda f: int((det["frame"].values == f).sum()))
        chosen = sorted(set([sizes[0], sizes[len(sizes) // 4], sizes[len(sizes) // 2], sizes[-2], sizes[-1]]))
        frames_df = [det[det["frame"] == f][["id", "id_cam", "xw", "yw", "frame"]] for f in chosen]
        tag = dataset.split("-")[1].lower()
        save_graph_case("%s_real" % tag, frames_df, [md] * len(frames_df), seed=11)
        # also keep the raw table so the "next" row (GT parse + homography) can be pinned later
        print("   frames", chosen, "sizes", [len(d) for d in frames_df])
    rng = np.random.default_rng(5)
    syn = ragged_frames(rng, n_frames=4, n_cams=4, max_ids=6)
    to_df = lambda fr: pd.DataFrame({k: fr[k] for k in ("id", "id_cam", "xw", "yw", "frame")})
    save_graph_case("synthetic_full", [to_df(f) for f in syn], [153.54] * 4, seed=12)
    save_graph_case("synthetic_only_appearance", [to_df(f) for f in syn], [153.54] * 4, seed=13, mode="only_appearance")
    save_graph_case("synthetic_only_dist", [to_df(f) for f in syn], [153.54] * 4, seed=14, mode="only_dist")
    # camera ids not in row order (cameras interleaved): the enumeration is then camera-major, not row-major
    mixed = []
    for fr in ragged_frames(rng, n_frames=3, n_cams=3, max_ids=5):
        perm = rng.permutation(len(fr["id"]))
        mixed.append(to_df({k: fr[k][perm] for k in ("id", "id_cam", "xw", "yw", "frame")}))
    save_graph_case("synthetic_interleaved_cams", mixed, [50.83] * 3, seed=15)
    # full-occupancy Terrace-shaped frames (4 cams x 9), the BASELINE.json cfg-1 shape
    full = ragged_frames(rng, n_frames=2, n_cams=4, max_ids=9, p_keep=1.1)
    save_graph_case("synthetic_terrace_full", [to_df(f) for f in full], [50.83] * 2, seed=16)


# ----------------------------------------------------------------------------------------------
# loss + statistics: the reference's own compute_loss_acc / FocalLoss_binary source, executed as is
# ----------------------------------------------------------------------------------------------
def gen_loss():
    """``loss_<case>.npz``: ``compute_loss_acc`` (train.py:51-208) run UNMODIFIED (its source text compiled with the
    names train.py imports: torch, np, F) under the criteria of main_training.py:258-268, with ``FocalLoss_binary``
    taken the same way from libs/utils.py:264-293; also d loss / d logits by torch autograd."""
    import types
    import torch.nn as nn
    import torch.nn.functional as F

    ns = {"torch": torch, "np": np, "F": F, "nn": nn}
    fn = ref_shim.reference_def("train.py", "compute_loss_acc", ns)
    focal_cls = ref_shim.reference_def("libs/utils.py", "FocalLoss_binary", {"torch": torch, "nn": nn})
    cases = [  # name, criterion name, pos_weight, mode, E, k, positive rate, logit scale, seed
        ("bce_train", "BCE", None, "train", 777, 3, 0.2, 3.0, 1),
        ("bce_weighted_train", "BCE_weighted", 4.7, "train", 1200, 4, 0.15, 4.0, 2),
        ("focal_train", "Focal", None, "train", 513, 2, 0.3, 2.0, 3),
        ("validate", "", None, "validate", 2050, 3, 0.25, 6.0, 4),
        ("no_positives", "BCE", None, "train", 100, 2, 0.0, 3.0, 5),
        ("no_negatives", "", None, "validate", 64, 1, 1.0, 3.0, 6),
        ("all_wrong", "BCE", None, "train", 90, 2, 0.5, 3.0, 7),
        ("extreme_logits", "BCE_weighted", 2.5, "train", 300, 3, 0.4, 60.0, 8),
    ]
    for name, crit, pw, mode, E, k, rate, scale, seed in cases:
        g = torch.Generator().manual_seed(seed)
        labels = (torch.rand(E, generator=g) < rate).float()
        logits = [(torch.randn(E, 1, generator=g) * scale + (labels.view(-1, 1) * 2 - 1) * (0.5 + 0.5 * s)).requires_grad_(True)
                  for s in range(k)]
        if name == "all_wrong":      # every last-step prediction wrong: the `sum_successes == 0` branches of train.py:108,116,124
            with torch.no_grad():
                logits[-1].copy_(-(labels.view(-1, 1) * 2 - 1) * 2.0)
        if crit == "Focal":
            c, cn = focal_cls(reduction="mean"), focal_cls(reduction="none")
        elif crit == "BCE_weighted":
            c, cn = nn.BCEWithLogitsLoss(reduction="mean", pos_weight=torch.tensor(pw)), nn.BCEWithLogitsLoss(reduction="none", pos_weight=torch.tensor(pw))
        elif crit == "BCE":
            c, cn = nn.BCEWithLogitsLoss(reduction="mean"), nn.BCEWithLogitsLoss(reduction="none")
        else:
            c, cn = "", ""
        batch = types.SimpleNamespace(edge_labels=labels)
        loss, p1, p0, pa, l1, l0, probs = fn({"classified_edges": logits}, batch, c, cn, mode)
        grads = torch.autograd.grad(loss, logits)
        out = dict(labels=labels.numpy(), k=np.int64(k), criterion=np.asarray(crit), mode=np.asarray(mode),
                   pos_weight=np.float64(-1.0 if pw is None else pw), loss=loss.detach().numpy(), loss_class1=l1.detach().numpy(),
                   loss_class0=l0.detach().numpy(), precision_class1=np.float64(p1[0]), precision_class0=np.float64(p0[0]),
                   precision_all=np.float64(pa[0]))
        for s in range(k):
            out["logits%d" % s] = logits[s].detach().numpy()
            out["probs%d" % s] = probs[s].detach().numpy()
            out["grad%d" % s] = grads[s].numpy()
        np.savez_compressed(os.path.join(GOLDEN, "loss_%s.npz" % name), **out)
        print("loss_%s: E=%d k=%d loss=%.6f p1=%.3f p0=%.3f p=%.3f" % (name, E, k, float(loss), p1[0], p0[0], pa[0]))


# ----------------------------------------------------------------------------------------------
# edge decisions + pruning + SCC clusters: the reference's utils functions, executed as is with networkx
# ----------------------------------------------------------------------------------------------
def gen_post():
    """``post_<case>.npz``: inference.py:286-312 replayed with the UNMODIFIED ``remove_edges_single_direction`` and
    ``compute_SCC_and_Clusters`` (libs/utils.py:387-404, 295-317; source text compiled with np / nx / torch) on ragged
    cross-camera frame batches with noisy logits (asymmetric decisions, chains, isolated nodes)."""
    import networkx as nx

    ns = {"np": np, "nx": nx, "torch": torch}
    prune_fn = ref_shim.reference_def("libs/utils.py", "remove_edges_single_direction", ns)
    scc_fn = ref_shim.reference_def("libs/utils.py", "compute_SCC_and_Clusters", ns)
    cases = [("small", 5, 3, 4, 1.0, 21), ("noisy", 12, 4, 6, 3.0, 22), ("wide", 6, 4, 10, 2.0, 23), ("confident", 8, 3, 5, 0.2, 24),
             ("all_inactive", 4, 3, 4, -1.0, 25)]
    for name, n_frames, n_cams, max_ids, noise, seed in cases:
        rng = np.random.default_rng(seed)
        frames = ragged_frames(rng, n_frames, n_cams, max_ids)
        src, dst, lab, ptr, eptr = [], [], [], [0], [0]
        for f in frames:                                  # cross-camera edges both ways, camera-ascending like train.py:286-306
            n0, n = ptr[-1], len(f["id"])
            for a in range(n):
                for b in range(n):
                    if f["id_cam"][a] != f["id_cam"][b]:
                        src.append(n0 + a); dst.append(n0 + b); lab.append(float(f["id"][a] == f["id"][b]))
            ptr.append(n0 + n); eptr.append(len(src))
        edge_list = np.asarray([src, dst], np.int64)
        labels = np.asarray(lab, np.float32)
        if noise < 0:
            logits = -np.abs(rng.normal(0, 1, size=len(lab))).astype(np.float32) - 0.1
        else:
            logits = ((labels * 2 - 1) * 2.0 + rng.normal(0, noise, size=len(lab))).astype(np.float32)
        n_nodes = ptr[-1]
        preds = torch.from_numpy(logits)
        predictions = (torch.nn.Sigmoid()(preds) >= 0.5) * 1                                     # inference.py:289-291
        act = [(edge_list[0][pos], edge_list[1][pos]) for pos, p in enumerate(predictions) if p == 1]
        ids_raw, n_raw = scc_fn(nx.DiGraph(act), n_nodes)                                        # inference.py:300-302
        pruned, act_p = prune_fn(act, predictions, edge_list)                                    # inference.py:307-308
        ids_p, n_p = scc_fn(nx.DiGraph(act_p), n_nodes)
        gt = [(edge_list[0][pos], edge_list[1][pos]) for pos, p in enumerate(labels) if p == 1]  # inference.py:297-299
        ids_gt, n_gt = scc_fn(nx.DiGraph(gt), n_nodes)
        np.savez_compressed(os.path.join(GOLDEN, "post_%s.npz" % name), edge_index=edge_list, labels=labels, logits=logits,
                            ptr=np.asarray(ptr, np.int64), edge_ptr=np.asarray(eptr, np.int64), predictions=predictions.numpy().astype(np.int64),
                            ids_raw=ids_raw.astype(np.int64), n_raw=np.int64(n_raw), pruned=pruned.numpy().astype(np.int64),
                            ids_pruned=ids_p.astype(np.int64), n_pruned=np.int64(n_p), ids_gt=ids_gt.astype(np.int64), n_gt=np.int64(n_gt))
        print("post_%s: N=%d E=%d active=%d pruned=%d clusters raw/pruned/gt=%d/%d/%d" % (
            name, n_nodes, len(lab), int(predictions.sum()), int(pruned.sum()), n_raw, n_p, n_gt))


# ----------------------------------------------------------------------------------------------
# detection tables: homography function and valid-frame loop of the reference, from its own source
# ----------------------------------------------------------------------------------------------
def gen_dets():
    """``dets_<sequence>.npz``: a window of the real EPFL ground truth parsed by the rules of datasets.py:89-98; world
    coordinates by the reference's OWN ``apply_homography_image_to_world`` (utils.py:211-225, source compiled as is, called
    per detection as :107 does for scalars) on the foot points of :101-102; valid frames by the reference's OWN loop
    (datasets.py:126-134: the source lines executed against a stand-in ``self`` with the DataFrame, ``np.int`` mapped to
    ``int`` because numpy 2 removed the alias)."""
    import textwrap
    import types
    import pandas as pd

    homog = ref_shim.reference_def("libs/utils.py", "apply_homography_image_to_world", {"np": np})
    with open(os.path.join(REF, "libs", "datasets.py")) as f:
        lines = f.read().splitlines()
    loop_src = textwrap.dedent("\n".join(lines[123:134]))            # datasets.py:124-134
    assert "frames_valid = []" in loop_src and "np.bincount" in loop_src, loop_src
    npx = types.SimpleNamespace(**{k: getattr(np, k) for k in ("min", "max", "unique", "bincount")}, int=int)
    for dataset, f_lo, f_hi in (("EPFL-Terrace", 0, 400), ("EPFL-Laboratory", 1000, 1300)):
        det = parse_gt(dataset)
        det = det[(det["frame"].values >= f_lo) & (det["frame"].values < f_hi)]
        root = os.path.join(REF, "datasets", dataset)
        cams = sorted(c for c in os.listdir(root) if c[0] != ".")
        Hs = np.stack([np.asarray(pd.read_csv(os.path.join(root, c, "Homography.txt"), header=None, sep="\t"), np.float64) for c in cams])
        bx = det["xmin"].values + np.round(det["width"].values / 2)
        by = det["ymax"].values
        xw = np.empty(len(det)); yw = np.empty(len(det))
        for i, (x, y, c) in enumerate(zip(bx, by, det["id_cam"].values)):
            xw[i], yw[i] = homog(float(x), float(y), Hs[int(c)])
        ns = {"np": npx, "self": types.SimpleNamespace(data_det=det.copy())}
        exec(loop_src, ns)
        frames_valid = np.asarray(ns["frames_valid"], np.int64)
        some = frames_valid[:: max(1, len(frames_valid) // 7)][:7]
        rows = np.concatenate([np.nonzero(det["frame"].values == f)[0] for f in some])     # datasets.py:250-251
        name = dataset.split("-")[1].lower()
        np.savez_compressed(os.path.join(GOLDEN, "dets_%s.npz" % name),
                            **{k: det[k].values.astype(np.int64) for k in ("id", "xmin", "ymin", "xmax", "ymax", "frame", "id_cam")},
                            homographies=Hs, xw=xw, yw=yw, frames_valid=frames_valid, some_frames=some, some_rows=rows.astype(np.int64))
        print("dets_%s: M=%d frames %d..%d valid=%d" % (name, len(det), f_lo, f_hi, len(frames_valid)))


def main():
    if not ref_shim.reference_available():
        sys.exit("needs /root/reference (build container only)")
    os.makedirs(GOLDEN, exist_ok=True)
    if "--loss-only" in sys.argv:
        return gen_loss()
    if "--post-only" in sys.argv:
        return gen_post()
    if "--dets-only" in sys.argv:
        return gen_dets()
    gen_mpn()
    gen_graph()
    gen_loss()
    gen_post()
    gen_dets()


if __name__ == "__main__":
    main()
