"""Oracle for the cell-path tile post-processing (TEST INFRASTRUCTURE ONLY).

Follows ``wsinsight/modellib/tilefuse.py`` of the reference:

    _proc_np_hv          :39-103    threshold, label + remove_small_objects, HV min-max normalise, Sobel ksize 21,
                                    Gaussian 3x3, ridge / marker, binary_fill_holes, 5x5 elliptical opening, label +
                                    remove_small_objects, watershed(dist, markers, mask)
    _stitching_worker    :109-174   interior crop, per-instance bbox / class means / largest external contour
    finalize             :280-415   2048-px tiles with 64 px of padding

The OpenCV / scipy calls are executed verbatim.  ``skimage`` is not importable here, so two calls are restated
("parity unpinned" for them): ``remove_small_objects`` on a label image (sizes by ``bincount``, strict ``<``) and
``skimage.segmentation.watershed`` (Meyer flooding over a binary heap keyed by (value, age), 4-connectivity, labels
assigned when a pixel is pushed).  Exact float ties in the flooding order are decided by the heap's insertion order
in skimage; every implementation can differ on them and the tests count them apart.

``trace_external_contour`` restates OpenCV's border follower (``cv2.findContours(RETR_EXTERNAL,
CHAIN_APPROX_SIMPLE)``) for one component and IS pinned against cv2 in tests/test_oracle.py.
"""

from __future__ import annotations

import heapq

import cv2
import numpy as np
from scipy import ndimage as ndi
from scipy.ndimage import binary_fill_holes


def remove_small_objects_labeled(labeled: np.ndarray, min_size: int) -> np.ndarray:
    """skimage.morphology.remove_small_objects on an integer label image: labels with fewer than ``min_size`` pixels
    become 0 (label 0 itself is background and, being counted too, is cleared as well when small -- a no-op)."""
    out = labeled.copy()
    if min_size == 0:
        return out
    sizes = np.bincount(labeled.ravel())
    too_small = sizes < min_size
    out[too_small[labeled]] = 0
    return out


def watershed_restated(image: np.ndarray, markers: np.ndarray, mask: np.ndarray) -> np.ndarray:
    """skimage.segmentation.watershed(image, markers=markers, mask=mask), connectivity 1, no compactness, no lines."""
    h, w = image.shape
    out = np.where(mask, markers, 0).astype(np.int32)
    heap = []
    age = 0
    for idx in np.flatnonzero(out.ravel()):
        heap.append((float(image.flat[idx]), 0, int(idx)))
    heapq.heapify(heap)
    flat = out.ravel()
    img = image.ravel()
    msk = mask.ravel()
    while heap:
        _, _, idx = heapq.heappop(heap)
        y, x = divmod(idx, w)
        for dy, dx in ((-1, 0), (0, -1), (0, 1), (1, 0)):
            yy, xx = y + dy, x + dx
            if yy < 0 or xx < 0 or yy >= h or xx >= w:
                continue
            j = yy * w + xx
            if not msk[j] or flat[j]:
                continue
            age += 1
            flat[j] = flat[idx]
            heapq.heappush(heap, (float(img[j]), age, j))
    return out


def flood_levels(image: np.ndarray, markers: np.ndarray, mask: np.ndarray) -> np.ndarray:
    """Pop level of every pixel of the Meyer flooding: L(m) = image(m) on markers, L(p) = max(image(p), min over
    4-neighbours q of L(q)) elsewhere (inf where unreachable).  The flooding labels p from the neighbour with the
    smallest L; two neighbours with EQUAL L and different labels are an exact tie."""
    h, w = image.shape
    L = np.full((h, w), np.inf)
    m = (markers > 0) & mask
    L[m] = image[m]
    heap = [(float(L[y, x]), y, x) for y, x in zip(*np.nonzero(m))]
    heapq.heapify(heap)
    while heap:
        l, y, x = heapq.heappop(heap)
        if l > L[y, x]:
            continue
        for dy, dx in ((-1, 0), (0, -1), (0, 1), (1, 0)):
            yy, xx = y + dy, x + dx
            if yy < 0 or xx < 0 or yy >= h or xx >= w or not mask[yy, xx] or m[yy, xx]:
                continue
            c = max(float(image[yy, xx]), l)
            if c < L[yy, xx]:
                L[yy, xx] = c
                heapq.heappush(heap, (c, yy, xx))
    return L


def proc_np_hv_stages(np_map: np.ndarray, hv_map: np.ndarray, min_object_size: int) -> dict:
    """``_proc_np_hv`` (tilefuse.py:39-103) with every intermediate kept."""
    H, W = np_map.shape[:2]
    st = {}
    blb_bin = (np_map >= 0.5).astype(np.uint8)
    zero = np.zeros((H, W), dtype=np.int32)
    if blb_bin.sum() == 0:
        return dict(inst=zero, empty=True)
    labeled, num = ndi.label(blb_bin)
    if num > 1:
        labeled = remove_small_objects_labeled(labeled, int(min_object_size))
    blb = (labeled > 0).astype(np.uint8)
    st["blb"] = blb
    if blb.sum() == 0:
        return dict(inst=zero, empty=True, **st)
    h_dir = cv2.normalize(hv_map[:, :, 0], None, alpha=0, beta=1, norm_type=cv2.NORM_MINMAX, dtype=cv2.CV_32F)
    v_dir = cv2.normalize(hv_map[:, :, 1], None, alpha=0, beta=1, norm_type=cv2.NORM_MINMAX, dtype=cv2.CV_32F)
    sobelh = cv2.Sobel(h_dir, cv2.CV_64F, 1, 0, ksize=21)
    sobelv = cv2.Sobel(v_dir, cv2.CV_64F, 0, 1, ksize=21)
    sobelh = 1.0 - cv2.normalize(sobelh, None, alpha=0, beta=1, norm_type=cv2.NORM_MINMAX, dtype=cv2.CV_32F)
    sobelv = 1.0 - cv2.normalize(sobelv, None, alpha=0, beta=1, norm_type=cv2.NORM_MINMAX, dtype=cv2.CV_32F)
    overall = np.maximum(sobelh, sobelv)
    overall = overall - (1.0 - blb)
    overall[overall < 0] = 0
    st["overall"] = overall
    dist = (1.0 - overall) * blb
    dist = -cv2.GaussianBlur(dist, (3, 3), 0)
    st["dist"] = dist
    ridge = (overall >= 0.4).astype(np.uint8)
    marker = blb.astype(np.int16) - ridge.astype(np.int16)
    marker = np.clip(marker, 0, 1).astype(np.uint8)
    st["marker0"] = marker.copy()
    if marker.any():
        marker = binary_fill_holes(marker).astype("uint8")
        kernel = cv2.getStructuringElement(cv2.MORPH_ELLIPSE, (5, 5))
        marker = cv2.morphologyEx(marker, cv2.MORPH_OPEN, kernel)
        st["marker_open"] = marker.copy()
        marker, mnum = ndi.label(marker)
        if mnum > 1:
            marker = remove_small_objects_labeled(marker, int(min_object_size))
    else:
        marker, _ = ndi.label(blb)
    st["marker"] = marker
    inst = watershed_restated(dist, marker, blb.astype(bool)).astype(np.int32)
    return dict(inst=inst, empty=False, **st)


def proc_np_hv(np_map, hv_map, min_object_size):
    return proc_np_hv_stages(np_map, hv_map, min_object_size)["inst"]


def stitching_worker(np_tile, hv_tile, tp_tile, interior_y0, interior_x0, interior_slice, min_object_size, inst=None):
    """``_stitching_worker`` (tilefuse.py:109-174): lists of (1,4) [x, y, w, h], (1,K) class means, (M,2) polygons."""
    pred = proc_np_hv(np_tile, hv_tile, min_object_size) if inst is None else inst
    ys, xs = interior_slice
    labels = pred[ys, xs]
    max_id = int(labels.max()) if labels.size else 0
    if max_id <= 0:
        return [], [], []
    lbl = labels.ravel()
    counts = np.bincount(lbl, minlength=max_id + 1).astype(np.int32)
    counts[0] = 0
    valid_ids = np.nonzero(counts)[0]
    if valid_ids.size == 0:
        return [], [], []
    slices = ndi.find_objects(labels, max_label=max_id)
    K = int(tp_tile.shape[2])
    tp_flat = tp_tile[ys, xs, :].reshape(-1, K).astype(np.float64)
    cls_sums = np.zeros((max_id + 1, K), dtype=np.float64)
    np.add.at(cls_sums, lbl, tp_flat)
    cls_sums[0, :] = 0
    denom = counts.astype(np.float64)
    denom[denom == 0] = 1.0
    cls_means = (cls_sums.T / denom).T.astype(np.float32)
    inst_list, prob_list, poly_list = [], [], []
    for inst_id in valid_ids.tolist():
        sl = slices[inst_id - 1]
        if sl is None:
            continue
        r_sl, c_sl = sl
        rmin, rmax, cmin, cmax = r_sl.start, r_sl.stop, c_sl.start, c_sl.stop
        x, y, w, h = cmin + interior_x0, rmin + interior_y0, cmax - cmin, rmax - rmin
        inst_list.append(np.array([x, y, w, h], dtype=np.int32).reshape(1, -1))
        prob_list.append(cls_means[inst_id].copy().reshape(1, -1))
        local = (labels[rmin:rmax, cmin:cmax] == inst_id).astype(np.uint8)
        cnts, _ = cv2.findContours(local, cv2.RETR_EXTERNAL, cv2.CHAIN_APPROX_SIMPLE)
        if not cnts:
            continue
        cnt = max(cnts, key=cv2.contourArea)
        poly = cnt.squeeze(1).astype(np.int32)
        if poly.ndim != 2 or poly.shape[0] < 3:
            continue
        poly[:, 0] += x
        poly[:, 1] += y
        poly_list.append(poly)
    return inst_list, prob_list, poly_list


def finalize_jobs(H, W, tile_size=2048, padding_size=64):
    """Index-only jobs of ``finalize`` (tilefuse.py:296-321)."""
    jobs = []
    for iy0 in range(0, H, tile_size):
        for ix0 in range(0, W, tile_size):
            iy1, ix1 = min(iy0 + tile_size, H), min(ix0 + tile_size, W)
            py0, py1 = max(0, iy0 - padding_size), min(iy1 + padding_size, H)
            px0, px1 = max(0, ix0 - padding_size), min(ix1 + padding_size, W)
            jobs.append((py0, py1, px0, px1, iy0, ix0, iy0 - py0, iy0 - py0 + (iy1 - iy0), ix0 - px0, ix0 - px0 + (ix1 - ix0)))
    return jobs


def finalize(np_map, hv_map, tp_map, min_object_size, tile_size=2048, padding_size=64):
    """``TileRemapStitcher.finalize`` (tilefuse.py:280-415), tiles in raster order (the reference's order depends on
    its thread pool; compare as sets)."""
    H, W = np_map.shape
    inst_all, prob_all, poly_all = [], [], []
    for (py0, py1, px0, px1, iy0, ix0, a0, a1, b0, b1) in finalize_jobs(H, W, tile_size, padding_size):
        ins, prb, ply = stitching_worker(np.ascontiguousarray(np_map[py0:py1, px0:px1]),
                                         np.ascontiguousarray(hv_map[py0:py1, px0:px1]),
                                         np.ascontiguousarray(tp_map[py0:py1, px0:px1]), iy0, ix0,
                                         (slice(a0, a1), slice(b0, b1)), min_object_size)
        inst_all.extend(ins)
        prob_all.extend(prb)
        poly_all.extend(ply)
    return inst_all, prob_all, poly_all


# ----------------------------------------------------------------------------
# OpenCV border following, restated (what the CUDA contour kernel implements)
# ----------------------------------------------------------------------------
_DX = (1, 1, 0, -1, -1, -1, 0, 1)      # OpenCV chain codes: 0 = E, 1 = NE, 2 = N, 3 = NW, 4 = W, 5 = SW, 6 = S, 7 = SE
_DY = (0, -1, -1, -1, 0, 1, 1, 1)


def trace_external_contour(binary: np.ndarray, start=None) -> np.ndarray:
    """Outer border of the component whose first raster pixel is ``start`` (default: the first non-zero pixel),
    with CHAIN_APPROX_SIMPLE compression: (M, 2) int32 [x, y] -- ``icvFetchContour`` of OpenCV's contours.cpp."""
    h, w = binary.shape
    b = np.pad(binary != 0, 1)

    def at(x, y):
        return b[y + 1, x + 1]

    if start is None:
        ys, xs = np.nonzero(binary)
        if len(ys) == 0:
            return np.zeros((0, 2), np.int32)
        k = np.lexsort((xs, ys))[0]
        start = (int(xs[k]), int(ys[k]))
    x0, y0 = start
    # find the first neighbour clockwise from W
    s_end = s = 4
    while True:
        s = (s - 1) & 7
        if at(x0 + _DX[s], y0 + _DY[s]) or s == s_end:
            break
    pts = []
    if s == s_end:
        return np.array([[x0, y0]], np.int32)      # isolated pixel
    x1, y1 = x0 + _DX[s], y0 + _DY[s]
    x3, y3 = x0, y0
    prev_s = s ^ 4
    px, py = x0, y0
    while True:
        s_end = s
        while True:
            s = (s + 1) & 7
            x4, y4 = x3 + _DX[s], y3 + _DY[s]
            if at(x4, y4):
                break
        if s != prev_s:
            pts.append((px, py))
        prev_s = s
        px, py = px + _DX[s], py + _DY[s]
        if (x4, y4) == (x0, y0) and (x3, y3) == (x1, y1):
            break
        x3, y3 = x4, y4
        s = (s + 4) & 7
    return np.array(pts, np.int32)
