"""TEST INFRASTRUCTURE ONLY -- CPU statement of the density grid.

**Parity unpinned**: the reference has no heat-map implementation.  The only upstream primitive is the
object centre of ``wsinsight/insightlib/insight_helpers.py:26-27``, executed here verbatim with numpy; the binning
(floor division by the bin size, first arg-max class) is this repo's definition and is what the kernel is held to.
"""

from __future__ import annotations

import numpy as np


def centers(xywh: np.ndarray):
    xywh = np.asarray(xywh)
    cx = np.rint(xywh[:, 0] + (xywh[:, 2] / 2)).astype(np.int64)      # insight_helpers.py:26
    cy = np.rint(xywh[:, 1] + (xywh[:, 3] / 2)).astype(np.int64)      # insight_helpers.py:27
    return cx, cy


def density_grid(xywh, probs, slide_width: int, slide_height: int, bin_px: int):
    probs = np.asarray(probs, dtype=np.float32)
    if probs.ndim == 1:
        probs = probs[:, None]
    k = probs.shape[1]
    gy, gx = -(-slide_height // bin_px), -(-slide_width // bin_px)
    grid = np.zeros((k, gy, gx), dtype=np.int32)
    cx, cy = centers(xywh)
    cls = probs.argmax(axis=1) if k > 1 else np.zeros(len(probs), dtype=np.int64)
    bx, by = np.floor_divide(cx, bin_px), np.floor_divide(cy, bin_px)
    ok = (cx >= 0) & (cy >= 0) & (bx < gx) & (by < gy)
    np.add.at(grid, (cls[ok], by[ok], bx[ok]), 1)
    return grid, int((~ok).sum())
