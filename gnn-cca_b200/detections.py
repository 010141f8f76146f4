"""Detection tables -> device tensors: the part of ``libs/datasets.py`` that feeds graph construction.

The reference keeps a pandas ``DataFrame`` per sequence (``EPFL_dataset.__init__``, datasets.py:53-134), projects every
bounding box's foot point to the ground plane with the camera's homography (utils.py:211-225) in Python, selects the
"valid" frames with a Python loop over ALL frame numbers (datasets.py:126-134) and slices the frame's rows per item
(datasets.py:250-251).  Here the table lives on the device:

    seq = gnncca.detections.prepare_sequence(**gnncca.detections.load_epfl(root, "EPFL-Terrace"), device="cuda")
    seq.frames_valid                                   # == EPFL_dataset.frames_valid
    tab = seq.gather(frames)                           # rows of these frames, frame after frame, camera-major inside a frame
    batch = gnncca.build_graph_batch(max_dist=..., reid_embeds=reid[tab["rows"]], node_embeds=node[tab["rows"]], **tab["graph"])

File parsing (``load_epfl``) is host I/O and stays numpy; everything after it runs in ``libgnncca.so`` (no CPU fallback).
"""
from __future__ import annotations

import os
from typing import Dict, Sequence

import numpy as np
import torch

from . import _lib


def load_epfl(root: str, name: str) -> Dict[str, np.ndarray]:
    """Parse ``<root>/<name>/<camera>/gt/gt.txt`` + ``Homography.txt`` by the rules of datasets.py:53-99 (EPFL branch):
    first 10 columns (id xmin ymin xmax ymax frame lost occluded generated label), rows with lost == 0, xmin += 1,
    id_cam = last character of the camera directory name, cameras in sorted order (camera-major table).  Basketball
    additionally keeps id <= 4 and frame <= 3000 (datasets.py:79-85)."""
    base = os.path.join(root, name)
    cams = sorted(c for c in os.listdir(base) if c[0] != ".")
    cols, Hs = {k: [] for k in ("id", "xmin", "ymin", "xmax", "ymax", "frame", "id_cam")}, []
    for c in cams:
        raw = np.loadtxt(os.path.join(base, c, "gt", "gt.txt"), usecols=range(9), dtype=np.int64, ndmin=2)
        keep = raw[:, 6] == 0                                              # lost == 0
        if "Basketball" in name:
            keep &= (raw[:, 0] <= 4) & (raw[:, 5] <= 3000)
        raw = raw[keep]
        cols["id"].append(raw[:, 0]); cols["xmin"].append(raw[:, 1] + 1); cols["ymin"].append(raw[:, 2])
        cols["xmax"].append(raw[:, 3]); cols["ymax"].append(raw[:, 4]); cols["frame"].append(raw[:, 5])
        cols["id_cam"].append(np.full(raw.shape[0], int(c[-1:]), np.int64))
        Hs.append(np.loadtxt(os.path.join(base, c, "Homography.txt"), dtype=np.float64).reshape(3, 3))
    out = {k: np.concatenate(v) for k, v in cols.items()}
    out["homographies"] = np.stack(Hs)           # indexed by position in the sorted camera list == id_cam for EPFL (c0..c3)
    return out


class DetectionTable:
    """A sequence's detections on the device (camera-major rows as parsed) + the frame index built over them."""

    def __init__(self, cols: Dict[str, torch.Tensor], xw, yw, frame_min: int, frame_ptr, perm, valid):
        self.cols, self.xw, self.yw = cols, xw, yw
        self.frame_min, self.frame_ptr, self.perm, self.valid = frame_min, frame_ptr, perm, valid
        self._frames_valid = None

    @property
    def frames_valid(self) -> torch.Tensor:
        """Frame numbers the reference trains on (datasets.py:126-141), ascending, int64 on the device."""
        if self._frames_valid is None:
            self._frames_valid = torch.nonzero(self.valid, as_tuple=False).view(-1) + self.frame_min
        return self._frames_valid

    def gather(self, frames: Sequence[int]):
        """Rows of ``frames`` (frame numbers), frame after frame, in the table's row order inside a frame (what
        ``data_det[data_det['frame'] == frame]`` returns, datasets.py:250-251).  Returns ``rows`` (indices into the
        table, e.g. to pick the CNN embeddings) and ``graph``: the per-detection arguments of ``build_graph_batch``."""
        dev = self.perm.device
        f = torch.as_tensor(frames, dtype=torch.int64, device=dev).view(-1) - self.frame_min
        if f.numel() and (int(f.min()) < 0 or int(f.max()) >= self.frame_ptr.numel() - 1):
            raise IndexError("frame number outside the sequence")
        a, b = self.frame_ptr[f].to(torch.int64), self.frame_ptr[f + 1].to(torch.int64)
        counts = b - a
        graph_ptr = torch.zeros(f.numel() + 1, dtype=torch.int64, device=dev)
        torch.cumsum(counts, 0, out=graph_ptr[1:])
        n = int(graph_ptr[-1])
        which = torch.repeat_interleave(torch.arange(f.numel(), device=dev), counts, output_size=n)
        pos = a[which] + (torch.arange(n, device=dev) - graph_ptr[:-1][which])
        rows = self.perm[pos].to(torch.int64)
        graph = dict(graph_ptr=graph_ptr, id_cam=self.cols["id_cam"][rows], ids=self.cols["id"][rows], xw=self.xw[rows], yw=self.yw[rows])
        return dict(rows=rows, graph=graph)


def prepare_sequence(id, xmin, ymin, xmax, ymax, frame, id_cam, homographies, device="cuda") -> DetectionTable:
    """Upload a parsed table (``load_epfl``) and build everything the loops need: ground-plane coordinates of every
    detection (gnncca_foot_to_world), the stable frame-sorted order with per-frame offsets (gnncca_csr_sort) and the valid
    frames (gnncca_frame_valid)."""
    lib = _lib.load()
    dev = torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError("gnn-cca_b200 prepares detection tables on CUDA tensors only (no CPU path)")
    t64 = lambda v: torch.as_tensor(np.asarray(v), dtype=torch.int64).to(dev).contiguous()
    cols = dict(id=t64(id), xmin=t64(xmin), ymin=t64(ymin), xmax=t64(xmax), ymax=t64(ymax), frame=t64(frame))
    cols["id_cam"] = torch.as_tensor(np.asarray(id_cam), dtype=torch.int32).to(dev).contiguous()
    M = int(cols["id"].numel())
    H = torch.as_tensor(np.asarray(homographies, dtype=np.float64)).to(dev).contiguous().view(-1, 9)
    stream = torch.cuda.current_stream(dev).cuda_stream
    xw = torch.empty(M, dtype=torch.float64, device=dev)
    yw = torch.empty(M, dtype=torch.float64, device=dev)
    flags = torch.zeros(2, dtype=torch.int32, device=dev)
    fx = [cols[k].to(torch.float64) for k in ("xmin", "xmax", "ymax")]
    _lib.check(lib.gnncca_foot_to_world(fx[0].data_ptr(), fx[1].data_ptr(), fx[2].data_ptr(), cols["id_cam"].data_ptr(), H.data_ptr(),
                                        int(H.size(0)), M, xw.data_ptr(), yw.data_ptr(), flags.data_ptr(), stream), "gnncca_foot_to_world")
    if M == 0:
        z = torch.zeros(1, dtype=torch.int32, device=dev)
        return DetectionTable(cols, xw, yw, 0, z, z[:0], torch.zeros(0, dtype=torch.bool, device=dev))
    fmin, fmax, bad = (int(v) for v in torch.stack([cols["frame"].min(), cols["frame"].max(), flags[0].to(torch.int64)]).tolist())
    if bad:
        raise IndexError("id_cam outside the list of homographies")
    n_frames = fmax - fmin + 1
    key = (cols["frame"] - fmin).to(torch.int32)
    other = torch.zeros(M, dtype=torch.int32, device=dev)
    perm = torch.empty(M, dtype=torch.int32, device=dev)
    frame_ptr = torch.empty(n_frames + 1, dtype=torch.int32, device=dev)
    nbytes = int(lib.gnncca_csr_sort_workspace_bytes(n_frames, M))
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    _lib.check(lib.gnncca_csr_sort(n_frames, M, key.data_ptr(), other.data_ptr(), perm.data_ptr(), frame_ptr.data_ptr(), ws.data_ptr(), nbytes,
                                   stream), "gnncca_csr_sort(frame)")
    valid = torch.empty(n_frames, dtype=torch.uint8, device=dev)
    _lib.check(lib.gnncca_frame_valid(frame_ptr.data_ptr(), perm.data_ptr(), cols["id_cam"].data_ptr(), cols["id"].data_ptr(), n_frames,
                                      valid.data_ptr(), stream), "gnncca_frame_valid")
    return DetectionTable(cols, xw, yw, fmin, frame_ptr, perm, valid.bool())
