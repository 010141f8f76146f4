"""CPU restatement of the reference's detection-table preparation (TEST INFRASTRUCTURE ONLY).

Only ``tests/`` may import this module; the product path (``gnn-cca_b200/detections.py`` -> ``libgnncca.so``) never does.

Restates /root/reference/libs/datasets.py:89-134 (EPFL branch: parse, foot point, homography, valid frames) and :250-251
(rows of a frame) with numpy; the homography is ``graph_oracle.homography_image_to_world`` (utils.py:211-225).  Pinned by
``tests/golden/dets_*.npz``: the parsed real EPFL ground truth with ``apply_homography_image_to_world`` executed from the
reference's own source per detection and the valid-frame loop of datasets.py:126-134 executed from its own source lines
(``oracle/gen_golden.py`` ``gen_dets``).
"""
from __future__ import annotations

import numpy as np

from . import graph_oracle


def foot_points(xmin, xmax, ymax):
    """datasets.py:95,101-102: (xmin + round(width / 2), ymax) with width = xmax - xmin."""
    xmin, xmax, ymax = (np.asarray(v, np.float64) for v in (xmin, xmax, ymax))
    return xmin + np.round((xmax - xmin) / 2), ymax


def world_coordinates(xmin, xmax, ymax, id_cam, homographies):
    bx, by = foot_points(xmin, xmax, ymax)
    xw, yw = np.empty_like(bx), np.empty_like(by)
    for c in np.unique(id_cam):
        m = np.asarray(id_cam) == c
        xw[m], yw[m] = graph_oracle.homography_image_to_world(bx[m], by[m], np.asarray(homographies[int(c)], np.float64))
    return xw, yw


def valid_frames(frame, id_cam, ids):
    """datasets.py:126-134: frames with detections of more than one camera and an identity occurring more than once."""
    frame, id_cam, ids = (np.asarray(v) for v in (frame, id_cam, ids))
    out = []
    for f in range(int(frame.min()), int(frame.max()) + 1):
        sel = frame == f
        if len(np.unique(id_cam[sel])) > 1 and np.max(np.bincount(ids[sel])) > 1:
            out.append(f)
    return np.asarray(out, np.int64)


def frame_rows(frame, f):
    """datasets.py:250-251: the table's rows of frame f, in table order."""
    return np.nonzero(np.asarray(frame) == f)[0]
