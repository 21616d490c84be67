"""Oracle for tissue segmentation (TEST INFRASTRUCTURE ONLY -- see oracle/__init__.py).

Follows ``wsinsight/patchlib/segment.py:13-97`` of the reference:

    cv.cvtColor(RGB2HSV)[:, :, 1]           segment.py:54,65
    cv.medianBlur(k)                        segment.py:68-77   (ValueError if k<=1 or even)
    cv.threshold(t, 255, THRESH_BINARY) > 0 segment.py:80-83
    skimage binary_closing(ones(c, c))      segment.py:87-89
    skimage remove_small_objects(min_size)  segment.py:92
    skimage remove_small_holes(area_thr)    segment.py:95

Two flavours are provided:

* ``segment_tissue``: the *same third-party calls* where importable (cv2) and
  scipy.ndimage for the three skimage calls (skimage 0.25.2 ``morphology/binary.py``
  and ``morphology/misc.py`` forward to exactly these scipy functions).
* ``*_restated`` helpers: pure-numpy integer restatements of each stage
  (SURVEY.md Appendix A 1-6); the CUDA kernels implement these formulas and the
  tests pin the restatements against the library calls.
"""

from __future__ import annotations

import cv2 as cv
import numpy as np
from scipy import ndimage as ndi

# ----------------------------------------------------------------------------
# Library-call flavour (what the reference executes)
# ----------------------------------------------------------------------------

_FOUR_CONN = ndi.generate_binary_structure(2, 1)


def binary_closing_lib(mask: np.ndarray, k: int, convention: str = "scipy_raw") -> np.ndarray:
    """skimage.morphology.binary_closing(mask, ones((k,k))) (segment.py:87-89) as scipy.ndimage calls.

    skimage ``binary_closing`` = ``binary_dilation`` then ``binary_erosion`` with default ``mode='ignore'``,
    i.e. dilation pads with False and erosion pads with True.  What is NOT pinned (skimage is not importable
    here) is how an EVEN footprint is centred before scipy sees it; the candidates, each as the scipy call it
    would be (they agree for odd k, and for even k everywhere except within k/2 px of the image border):

    ``scipy_raw``         the raw ones((k,k)) goes to scipy (default; == footprint end-padded to odd size)
    ``front_padded``      footprint padded with a zero row/column in FRONT to (k+1, k+1)
                          (skimage ``pad_footprint(fp, pad_end=False)``)
    ``mirrored_erosion``  raw dilation, erosion with the footprint mirrored about its centre
                          (for scipy: ``origin=-1`` on an even footprint)
    """
    fp = np.ones((k, k), dtype=bool)
    if convention == "scipy_raw" or k % 2 == 1:
        tmp = ndi.binary_dilation(mask, structure=fp, border_value=False)
        return ndi.binary_erosion(tmp, structure=fp, border_value=True)
    if convention == "front_padded":
        fpp = np.zeros((k + 1, k + 1), dtype=bool)
        fpp[1:, 1:] = True
        tmp = ndi.binary_dilation(mask, structure=fpp, border_value=False)
        return ndi.binary_erosion(tmp, structure=fpp, border_value=True)
    if convention == "mirrored_erosion":
        tmp = ndi.binary_dilation(mask, structure=fp, border_value=False)
        return ndi.binary_erosion(tmp, structure=fp, border_value=True, origin=-1)
    raise ValueError(convention)


def closing_origins(k: int, convention: str = "scipy_raw"):
    """(dilation origin, erosion origin): window offsets are [-origin .. k-1-origin]."""
    return {"scipy_raw": ((k - 1) // 2, k // 2), "front_padded": (k // 2, (k - 1) // 2),
            "mirrored_erosion": ((k - 1) // 2, (k - 1) // 2)}[convention]


def otsu_threshold_restated(img_u8: np.ndarray) -> int:
    """OpenCV ``getThreshVal_Otsu_8u`` (imgproc/thresh.cpp): the sequential float64 loop the CUDA kernel repeats.
    Pinned against ``cv.threshold(..., THRESH_OTSU)`` in tests/test_oracle.py."""
    h = np.bincount(img_u8.ravel(), minlength=256).astype(np.int64)
    eps = np.float64(np.finfo(np.float32).eps)
    scale = np.float64(1.0) / np.float64(img_u8.size)
    mu = np.float64(0)
    for i in range(256):
        mu += np.float64(i) * np.float64(h[i])
    mu *= scale
    mu1 = q1 = max_sigma = np.float64(0)
    max_val = 0
    for i in range(256):
        p_i = np.float64(h[i]) * scale
        mu1 *= q1
        q1 += p_i
        q2 = np.float64(1.0) - q1
        if min(q1, q2) < eps or max(q1, q2) > np.float64(1.0) - eps:
            continue
        mu1 = (mu1 + np.float64(i) * p_i) / q1
        mu2 = (mu - q1 * mu1) / q2
        sigma = q1 * q2 * (mu1 - mu2) * (mu1 - mu2)
        if sigma > max_sigma:
            max_sigma, max_val = sigma, i
    return int(max_val)


def remove_small_objects_lib(mask: np.ndarray, min_size: int) -> np.ndarray:
    """skimage.morphology.remove_small_objects(bool, min_size) (segment.py:92).

    4-connectivity (``connectivity=1``), strict ``<``; ``min_size == 0`` is a no-op.
    """
    out = mask.copy()
    if min_size == 0:
        return out
    ccs, _ = ndi.label(mask, structure=_FOUR_CONN)
    sizes = np.bincount(ccs.ravel())
    too_small = sizes < min_size
    out[too_small[ccs]] = False
    return out


def remove_small_holes_lib(mask: np.ndarray, area_threshold: int) -> np.ndarray:
    """skimage.morphology.remove_small_holes(bool, area_threshold) (segment.py:95)."""
    inv = np.logical_not(mask)
    inv = remove_small_objects_lib(inv, area_threshold)
    return np.logical_not(inv)


def segment_tissue(
    im_arr: np.ndarray,
    median_filter_size: int = 7,
    binary_threshold=7,
    closing_kernel_size: int = 6,
    min_object_size_px: int = 512,
    min_hole_size_px: int = 1024,
    closing_convention: str = "scipy_raw",
    return_threshold: bool = False,
) -> np.ndarray:
    """Restatement of ``segment_tissue`` (segment.py:13-97), same signature/defaults.

    ``binary_threshold="otsu"`` is the extra mode (not upstream): ``cv.threshold(sat, 0, 255, THRESH_BINARY |
    THRESH_OTSU)`` instead of the fixed threshold."""
    im_arr = np.asarray(im_arr)
    assert im_arr.ndim == 3
    assert im_arr.shape[2] == 3
    sat = cv.cvtColor(np.ascontiguousarray(im_arr), cv.COLOR_RGB2HSV)[:, :, 1]
    if median_filter_size <= 1 or median_filter_size % 2 == 0:
        raise ValueError(
            "median_filter_size must be greater than 1 and odd, but got"
            f" {median_filter_size}"
        )
    sat = cv.medianBlur(np.ascontiguousarray(sat), median_filter_size)
    if isinstance(binary_threshold, str):
        used, thr = cv.threshold(sat, thresh=0, maxval=255, type=cv.THRESH_BINARY | cv.THRESH_OTSU)
    else:
        used, thr = cv.threshold(sat, thresh=binary_threshold, maxval=255, type=cv.THRESH_BINARY)
    mask = thr > 0
    mask = binary_closing_lib(mask, closing_kernel_size, closing_convention)
    mask = remove_small_objects_lib(mask, min_object_size_px)
    mask = remove_small_holes_lib(mask, min_hole_size_px)
    return (mask, int(used)) if return_threshold else mask


def segment_stages(im_arr, median_filter_size=7, binary_threshold=7, closing_kernel_size=6,
                   min_object_size_px=512, min_hole_size_px=1024):
    """Same as ``segment_tissue`` but returns every intermediate (for debugging parity)."""
    sat = cv.cvtColor(np.ascontiguousarray(im_arr), cv.COLOR_RGB2HSV)[:, :, 1]
    med = cv.medianBlur(np.ascontiguousarray(sat), median_filter_size)
    thr = med > binary_threshold
    clo = binary_closing_lib(thr, closing_kernel_size)
    obj = remove_small_objects_lib(clo, min_object_size_px)
    hol = remove_small_holes_lib(obj, min_hole_size_px)
    return dict(sat=sat, med=med, thr=thr, closed=clo, objects=obj, mask=hol)


# ----------------------------------------------------------------------------
# Integer restatements (what the CUDA kernels implement)
# ----------------------------------------------------------------------------

def sdiv_table() -> np.ndarray:
    """OpenCV 8-bit HSV saturation divisor table: sdiv[i] = rint((255<<12)/i), sdiv[0]=0."""
    t = np.zeros(256, dtype=np.int64)
    i = np.arange(1, 256, dtype=np.float64)
    t[1:] = np.rint((255 << 12) / i).astype(np.int64)
    return t


def rgb_to_saturation_restated(rgb: np.ndarray) -> np.ndarray:
    """S = (d*sdiv[V] + 2048) >> 12 with V=max(r,g,b), d=V-min(r,g,b) (Appendix A.1)."""
    rgb = rgb.astype(np.int64)
    v = rgb.max(axis=-1)
    d = v - rgb.min(axis=-1)
    s = (d * sdiv_table()[v] + (1 << 11)) >> 12
    return s.astype(np.uint8)


def median_threshold_restated(sat: np.ndarray, k: int, thresh: int) -> np.ndarray:
    """(medianBlur(sat,k) > thresh) without computing the median.

    median > t  <=>  #{window elements > t} >= (k*k+1)/2, window with replicate
    border (cv2.medianBlur uses BORDER_REPLICATE).  This is the identity the
    CUDA kernel uses (threshold first, then a k x k box count).
    """
    if k <= 1 or k % 2 == 0:
        raise ValueError(f"median_filter_size must be greater than 1 and odd, but got {k}")
    r = k // 2
    b = (sat > thresh).astype(np.int32)
    p = np.pad(b, r, mode="edge")
    # separable box sum
    h, w = b.shape
    rows = np.zeros((h + 2 * r, w), dtype=np.int32)
    for dx in range(k):
        rows += p[:, dx:dx + w]
    cnt = np.zeros((h, w), dtype=np.int32)
    for dy in range(k):
        cnt += rows[dy:dy + h, :]
    return cnt >= (k * k + 1) // 2


def closing_restated(mask: np.ndarray, k: int, convention: str = "scipy_raw") -> np.ndarray:
    """Closing with ones(k,k) as explicit window offsets (Appendix A.4).

    dilation window offsets  [-d .. k-1-d], outside = 0;  erosion window offsets [-e .. k-1-e], outside = 1,
    (d, e) = ``closing_origins(k, convention)``; default (scipy_raw, k=6): [-2..+3] then [-3..+2].
    """
    h, w = mask.shape
    dlo, elo = closing_origins(k, convention)
    dhi, ehi = k - 1 - dlo, k - 1 - elo
    m = mask.astype(bool)
    # dilation (separable OR)
    p = np.pad(m, ((0, 0), (dlo, dhi)), constant_values=False)
    t = np.zeros_like(m)
    for o in range(k):
        t |= p[:, o:o + w]
    p = np.pad(t, ((dlo, dhi), (0, 0)), constant_values=False)
    d = np.zeros_like(m)
    for o in range(k):
        d |= p[o:o + h, :]
    # erosion (separable AND)
    p = np.pad(d, ((0, 0), (elo, ehi)), constant_values=True)
    t = np.ones_like(m)
    for o in range(k):
        t &= p[:, o:o + w]
    p = np.pad(t, ((elo, ehi), (0, 0)), constant_values=True)
    e = np.ones_like(m)
    for o in range(k):
        e &= p[o:o + h, :]
    return e


def outermost_restated(mask: np.ndarray) -> np.ndarray:
    """Pixels of the *outermost* 8-connected foreground components (Appendix A.8).

    A component is outermost iff one of its pixels lies on the image border or is
    4-adjacent to the 4-connected background component that touches the (implicit
    zero) frame around the image.  Islands inside holes are not outermost; the
    reference drops them (``patch.py:78-123``).
    """
    m = mask.astype(bool)
    h, w = m.shape
    pm = np.pad(m, 1, constant_values=False)
    bg_lab, _ = ndi.label(~pm, structure=_FOUR_CONN)
    outside = bg_lab == bg_lab[0, 0]
    fg_lab, n = ndi.label(pm, structure=np.ones((3, 3), dtype=bool))
    touch = np.zeros_like(pm)
    touch[1:, :] |= outside[:-1, :]
    touch[:-1, :] |= outside[1:, :]
    touch[:, 1:] |= outside[:, :-1]
    touch[:, :-1] |= outside[:, 1:]
    ok = np.zeros(n + 1, dtype=bool)
    ok[np.unique(fg_lab[touch & pm])] = True
    ok[0] = False
    return ok[fg_lab][1:-1, 1:-1]
