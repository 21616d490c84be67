"""Oracle for the per-patch data path (TEST INFRASTRUCTURE ONLY).

Follows the reference:

* ``WholeSlideImagePatches._get_patch_from_wsi`` wsinsight/modellib/data.py:270-281
  (``read_region`` -> RGBA, zero padded outside the slide, ``.convert("RGB")``).
* ``make_compose_from_transform_config`` wsinsight/modellib/transforms.py:22-38
  (``Resize`` = PIL antialiased BILINEAR on the PIL image, ``ToTensor``, ``Normalize``).

``preprocess_lib`` executes the very same PIL / torchvision calls.  The
``*_restated`` functions are the fixed-point formulas (Pillow ``Resample.c``:
``precompute_coeffs`` + ``ImagingResampleHorizontal/Vertical_8bpc``,
PRECISION_BITS = 32-8-2 = 22) that the CUDA kernel implements; tests pin them
bit-exact against PIL.
"""

from __future__ import annotations

import numpy as np
import torch
from PIL import Image
from torchvision import transforms as T

PRECISION_BITS = 32 - 8 - 2


def read_region_rgba(slide: np.ndarray, x: int, y: int, w: int, h: int) -> np.ndarray:
    """(H,W,4|3) u8 slide -> (h,w,4) RGBA crop at (x,y), zeros outside the slide.

    tiffslide/openslide ``read_region`` semantics (tests/test_all.py:747-765 of the
    reference: out-of-bounds reads are zero padded).  An RGB slide gets A=255 inside.
    """
    H, W = slide.shape[:2]
    out = np.zeros((h, w, 4), dtype=np.uint8)
    x0, y0 = max(x, 0), max(y, 0)
    x1, y1 = min(x + w, W), min(y + h, H)
    if x1 > x0 and y1 > y0:
        src = slide[y0:y1, x0:x1]
        out[y0 - y:y1 - y, x0 - x:x1 - x, :3] = src[..., :3]
        out[y0 - y:y1 - y, x0 - x:x1 - x, 3] = src[..., 3] if slide.shape[2] == 4 else 255
    return out


def make_compose(size: int | None, mean, std):
    ts = []
    if size is not None:
        ts.append(T.Resize(size=size))
    ts.append(T.ToTensor())
    if mean is not None:
        ts.append(T.Normalize(mean=mean, std=std))
    return T.Compose(ts)


def preprocess_lib(rgba: np.ndarray, size: int | None, mean, std) -> torch.Tensor:
    """RGBA crop -> ``.convert("RGB")`` -> Resize/ToTensor/Normalize -> (3,S,S) f32."""
    im = Image.fromarray(rgba, mode="RGBA").convert("RGB")
    return make_compose(size, mean, std)(im)


def resize_lib_u8(rgb: np.ndarray, size: int) -> np.ndarray:
    """torchvision ``Resize(size)`` on a PIL RGB image -> (S',S',3) u8."""
    im = Image.fromarray(rgb, mode="RGB")
    return np.asarray(T.Resize(size=size)(im))


# ----------------------------------------------------------------------------
# fixed-point restatement
# ----------------------------------------------------------------------------

def resample_taps(in_size: int, out_size: int):
    """Pillow ``precompute_coeffs`` for the BILINEAR (triangle) filter.

    Returns (xmin int32[out], xcount int32[out], coef int32[out, ksize]) with
    coefficients quantised to 22-bit fixed point (round half away from zero).
    """
    scale = in_size / out_size
    filterscale = max(scale, 1.0)
    support = 1.0 * filterscale
    ksize = int(np.ceil(support)) * 2 + 1
    xmin = np.zeros(out_size, dtype=np.int32)
    xcnt = np.zeros(out_size, dtype=np.int32)
    coef = np.zeros((out_size, ksize), dtype=np.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = (xx + 0.5) * scale
        lo = int(center - support + 0.5)
        if lo < 0:
            lo = 0
        hi = int(center + support + 0.5)
        if hi > in_size:
            hi = in_size
        n = hi - lo
        w = np.zeros(ksize, dtype=np.float64)
        ww = 0.0
        for x in range(n):
            v = (x + lo - center + 0.5) * ss
            v = -v if v < 0 else v
            wv = 1.0 - v if v < 1.0 else 0.0
            w[x] = wv
            ww += wv
        if ww != 0.0:
            w[:n] /= ww
        k = np.where(w < 0, (-0.5 + w * (1 << PRECISION_BITS)).astype(np.int64),
                     (0.5 + w * (1 << PRECISION_BITS)).astype(np.int64))
        xmin[xx] = lo
        xcnt[xx] = n
        coef[xx] = k.astype(np.int32)
    return xmin, xcnt, coef


def _clip8(v):
    return np.clip(v >> PRECISION_BITS, 0, 255).astype(np.uint8)


def resize_restated_u8(rgb: np.ndarray, out_w: int, out_h: int) -> np.ndarray:
    """Two-pass Pillow resample (horizontal first into a u8 intermediate, then vertical)."""
    h, w, c = rgb.shape
    src = rgb.astype(np.int64)
    if out_w != w:
        xmin, xcnt, coef = resample_taps(w, out_w)
        tmp = np.empty((h, out_w, c), dtype=np.uint8)
        for xx in range(out_w):
            n = xcnt[xx]
            seg = src[:, xmin[xx]:xmin[xx] + n, :]
            acc = (1 << (PRECISION_BITS - 1)) + np.einsum("hkc,k->hc", seg, coef[xx, :n].astype(np.int64))
            tmp[:, xx, :] = _clip8(acc)
    else:
        tmp = rgb
    if out_h != h:
        ymin, ycnt, coef = resample_taps(h, out_h)
        t64 = tmp.astype(np.int64)
        out = np.empty((out_h, tmp.shape[1], c), dtype=np.uint8)
        for yy in range(out_h):
            n = ycnt[yy]
            seg = t64[ymin[yy]:ymin[yy] + n]
            acc = (1 << (PRECISION_BITS - 1)) + np.einsum("kwc,k->wc", seg, coef[yy, :n].astype(np.int64))
            out[yy] = _clip8(acc)
    else:
        out = tmp
    return out


def normalize_restated(u8_hwc: np.ndarray, mean, std) -> np.ndarray:
    """ToTensor + Normalize in fp32: (x/255 - mean)/std, returns (H,W,3) f32."""
    x = u8_hwc.astype(np.float32) / np.float32(255.0)
    m = np.asarray(mean, dtype=np.float32)
    s = np.asarray(std, dtype=np.float32)
    return (x - m) / s


def preprocess_batch_lib(slide: np.ndarray, coords_xywh: np.ndarray, size, mean, std) -> torch.Tensor:
    """Reference data path for a list of patches -> (B,3,S,S) f32 (default collate)."""
    outs = []
    for (x, y, w, h) in np.asarray(coords_xywh).tolist():
        outs.append(preprocess_lib(read_region_rgba(slide, x, y, w, h), size, mean, std))
    return torch.stack(outs, 0)
