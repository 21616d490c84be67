"""Oracle for mask -> polygon -> patch grid (TEST INFRASTRUCTURE ONLY).

Follows the reference:

* ``get_multipolygon_from_binary_arr``  wsinsight/patchlib/patch.py:35-130
  (live ``cv2.findContours(RETR_CCOMP, CHAIN_APPROX_SIMPLE)``; contours scaled in
  float64 exactly as ``c * scale_arr`` at patch.py:73-74; sequential
  union/difference in hierarchy order, patch.py:78-118).
* ``get_patch_coordinates_within_polygon``  wsinsight/patchlib/patch.py:174-242.
* slide-level arithmetic of ``segment_and_patch_one_slide``
  wsinsight/patchlib/pipeline.py:96-112,160-164,357-382.

Shapely/GEOS is not importable here, so polygon algebra + ``contains`` are
restated point-wise: a point's membership is tracked through the same sequence
of ``union`` / ``difference`` operations with an even-odd point-in-polygon test
per contour.  Points lying exactly on a contour edge are reported in a separate
``tie`` mask (the reference decides those in GEOS).
"""

from __future__ import annotations

import cv2 as cv
import numpy as np

# ----------------------------------------------------------------------------
# contours -> ordered list of (add/subtract, polygon) operations
# ----------------------------------------------------------------------------


def contour_ops(mask_u8: np.ndarray, scale=None):
    """Return the reference's sequence of polygon operations.

    Each item is ``(add: bool, vertices: float64 (m,2))`` in the order in which
    ``merge_polygons`` (patch.py:78-118) applies them: contour ``idx`` first, then
    its child chain with ``not add``, then its ``next`` sibling with ``add``.
    Contours with <= 2 points are skipped (patch.py:89).  Returns ``None`` when
    there are no contours (patch.py:61-62).
    """
    contours, hierarchy = cv.findContours(
        np.ascontiguousarray(mask_u8), cv.RETR_CCOMP, cv.CHAIN_APPROX_SIMPLE
    )
    if hierarchy is None:
        return None
    hierarchy = hierarchy.squeeze(0)
    if scale is not None:
        scale_arr = np.array(scale).reshape(1, 1, 2)
        contours = tuple(c * scale_arr for c in contours)
    ops = []
    # iterative version of the recursive merge_polygons (same visiting order)
    stack = [(0, True)]
    while stack:
        idx, add = stack.pop()
        contour = np.squeeze(contours[idx])
        if contour.ndim == 2 and len(contour) > 2:
            ops.append((add, np.asarray(contour, dtype=np.float64)))
        child = hierarchy[idx][2]
        nxt = hierarchy[idx][0]
        # recursion order: child subtree first, then next sibling
        if nxt >= 0:
            stack.append((nxt, add))
        if child >= 0:
            stack.append((child, not add))
    return ops


def _pip(points: np.ndarray, verts: np.ndarray):
    """Even-odd point-in-polygon.  Returns (strictly_inside, on_boundary)."""
    n = len(points)
    inside = np.zeros(n, dtype=bool)
    onb = np.zeros(n, dtype=bool)
    if n == 0:
        return inside, onb
    vx, vy = verts[:, 0], verts[:, 1]
    # bounding-box prefilter
    sel = (
        (points[:, 0] >= vx.min()) & (points[:, 0] <= vx.max())
        & (points[:, 1] >= vy.min()) & (points[:, 1] <= vy.max())
    )
    idx = np.nonzero(sel)[0]
    if idx.size == 0:
        return inside, onb
    x2, y2 = np.roll(vx, -1), np.roll(vy, -1)
    chunk = max(1, int(4_000_000 // max(1, len(vx))))
    for s in range(0, idx.size, chunk):
        ii = idx[s:s + chunk]
        px = points[ii, 0][:, None].astype(np.float64)
        py = points[ii, 1][:, None].astype(np.float64)
        x1 = vx[None, :]
        y1 = vy[None, :]
        xx2 = x2[None, :]
        yy2 = y2[None, :]
        # on-segment test (exact collinearity in float64 + within the segment box)
        cross = (xx2 - x1) * (py - y1) - (px - x1) * (yy2 - y1)
        within = (
            (px >= np.minimum(x1, xx2)) & (px <= np.maximum(x1, xx2))
            & (py >= np.minimum(y1, yy2)) & (py <= np.maximum(y1, yy2))
        )
        onb[ii] = np.any((cross == 0) & within, axis=1)
        # crossing number with half-open rule
        cond = (y1 > py) != (yy2 > py)
        dy = np.where(yy2 == y1, 1.0, yy2 - y1)
        # point is left of the edge at height py  <=>  sign test with cross product
        left = np.where(dy > 0, cross > 0, cross < 0)
        crossings = np.count_nonzero(cond & left, axis=1)
        inside[ii] = (crossings & 1).astype(bool)
    inside &= ~onb
    return inside, onb


def points_in_region(points: np.ndarray, ops):
    """Membership of ``points`` in the region built by ``ops`` (strict interior).

    Returns ``(inside, tie)``; ``tie`` marks points that hit a contour edge in any
    operation (their reference verdict depends on GEOS boundary handling).
    """
    n = len(points)
    inside = np.zeros(n, dtype=bool)
    tie = np.zeros(n, dtype=bool)
    for add, verts in ops:
        pin, onb = _pip(points, verts)
        tie |= onb
        if add:
            inside |= pin
        else:
            inside &= ~(pin | onb)
    return inside, tie


# ----------------------------------------------------------------------------
# grid
# ----------------------------------------------------------------------------

def grid_centroids(slide_width, slide_height, patch_size, half_patch_size, overlap=0.0):
    """patch.py:210-225: step, then product(range(half, W, step), range(half, H, step))."""
    if overlap >= 1:
        raise ValueError(f"overlap must be in (-inf, 1) but got {overlap}")
    step = round((1 - overlap) * patch_size)
    xs = np.arange(half_patch_size, slide_width, step, dtype=np.int64)
    ys = np.arange(half_patch_size, slide_height, step, dtype=np.int64)
    cx = np.repeat(xs, len(ys))
    cy = np.tile(ys, len(xs))
    return np.stack([cx, cy], axis=1), step, (len(xs), len(ys))


def patch_coordinates(mask: np.ndarray, slide_width, slide_height, patch_size,
                      half_patch_size=None, overlap=0.0, return_ties=False):
    """mask (bool, thumbnail) -> (N,2) int64 top-left coords, as pipeline.py:160-168,357-368.

    Returns ``None`` when no contours are found (pipeline.py:165-167).
    """
    ht, wt = mask.shape
    scale = (slide_width / wt, slide_height / ht)
    ops = contour_ops(mask.astype(np.uint8) * 255, scale=scale)
    if ops is None:
        return None
    if half_patch_size is None:
        half_patch_size = round(patch_size / 2)
    cent, step, dims = grid_centroids(slide_width, slide_height, patch_size, half_patch_size, overlap)
    inside, tie = points_in_region(cent, ops)
    coords = cent[inside] - half_patch_size
    if return_ties:
        return coords, cent[tie] - half_patch_size
    return coords


def tile_dim(slide_width, slide_height, patch_size, half_patch_size, overlap=0.0):
    """pipeline.py:370-382: max((centroid-half)/step)+1 per axis, int32."""
    cent, step, dims = grid_centroids(slide_width, slide_height, patch_size, half_patch_size, overlap)
    return (((cent - half_patch_size) / step).max(axis=0).astype(np.int32) + 1)


def slide_params(slide_dims, thumb_size, mpp, patch_size_px, patch_spacing_um_px,
                 min_object_size_um2=200 ** 2, min_hole_size_um2=190 ** 2):
    """Scalar arithmetic of pipeline.py:96-112,359 (python ``round`` = banker's)."""
    patch_size = int(round(patch_size_px * patch_spacing_um_px / mpp))
    thumb_mpp = (mpp * (np.array(slide_dims) / np.array(thumb_size))).mean()
    sq = thumb_mpp ** 2
    return dict(
        patch_size=patch_size,
        half_patch_size=round(patch_size / 2),
        thumb_mpp=float(thumb_mpp),
        min_object_size_px=round(min_object_size_um2 / sq),
        min_hole_size_px=round(min_hole_size_um2 / sq),
    )


# ----------------------------------------------------------------------------
# Raster restatement of the contour/polygon/contains pipeline.
# This is the rule the CUDA grid kernel implements; tests check it against the
# contour oracle above on the CPU so the rule itself is pinned, not assumed.
# ----------------------------------------------------------------------------

def raster_rule(outer: np.ndarray, cent: np.ndarray, slide_width: int, slide_height: int):
    """inside-test of integer centroids against the *outermost-component* mask.

    Exact integer form: nu = cx*Wt, nv = cy*Ht; dual cell (i,j) = (nv // H, nu // W);
    remainders ru = nu % W, rv = nv % H are the in-cell fractions (fu = ru/W, fv = rv/H)
    compared cross-multiplied.  The region is the union of the closed dual cells whose
    four corner pixels are set and of the closed triangles of cells with exactly three
    corners set; a point is kept iff it lies in the *interior* of that region
    (Shapely ``contains`` is strict):

      ru>0, rv>0 : all four corners set, or exactly three set and the point strictly
                   on the far side of the diagonal opposite the missing corner
      ru==0,rv>0 : on the vertical edge a-c: a and c set, at least one of the two
                   right-hand corners and at least one of the two left-hand corners set
      ru>0,rv==0 : same for the horizontal edge a-b
      ru==0,rv==0: at pixel centre (i,j): the pixel and its four 4-neighbours set
    """
    ht, wt = outer.shape
    W, H = int(slide_width), int(slide_height)
    cx = cent[:, 0].astype(np.int64)
    cy = cent[:, 1].astype(np.int64)
    nu = cx * wt
    nv = cy * ht
    j = nu // W
    i = nv // H
    ru = nu - j * W
    rv = nv - i * H
    po = np.pad(outer.astype(bool), 2, constant_values=False)

    def o(ii, jj):
        ok = (ii >= -1) & (ii <= ht) & (jj >= -1) & (jj <= wt)
        return po[np.clip(ii, -2, ht + 1) + 2, np.clip(jj, -2, wt + 1) + 2] & ok

    a, b, c, d = o(i, j), o(i, j + 1), o(i + 1, j), o(i + 1, j + 1)
    cnt = a.astype(int) + b + c + d
    lu = ru * H
    lv = rv * W
    wh = W * H
    gen = cnt == 4
    three = cnt == 3
    gen = gen | (three & ~a & (lu + lv > wh))
    gen = gen | (three & ~b & (lv > lu))
    gen = gen | (three & ~c & (lu > lv))
    gen = gen | (three & ~d & (lu + lv < wh))
    on_v = a & c & (b | d) & (o(i, j - 1) | o(i + 1, j - 1))
    on_h = a & b & (c | d) & (o(i - 1, j) | o(i - 1, j + 1))
    on_p = a & o(i - 1, j) & c & o(i, j - 1) & b
    res = np.where((ru > 0) & (rv > 0), gen,
                   np.where((ru == 0) & (rv > 0), on_v,
                            np.where((ru > 0) & (rv == 0), on_h, on_p)))
    return res
