"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the thumbnail the reference segments.

Reference chain: ``patchlib/pipeline.py:102-105`` -> ``slide.get_thumbnail(thumbsize)`` -> tiffslide 2.5.1
``get_thumbnail``: read the best level (level 0 for a single-level slide), composite on the white background,
``Image.thumbnail(size, LANCZOS)`` with Pillow's default ``reducing_gap=2.0`` -- i.e. (Pillow ``Image.py``
``thumbnail``/``resize``, ``Reduce.c``, ``Resample.c``; pillow 11.3.0 pinned upstream, 12.2 importable here):

1. aspect-preserving target size (``round_aspect``),
2. integer box reduction by ``factor = int(side / target / 2) or 1`` per axis: every output pixel is
   ``((sum + n/2) * M(n)) >> 24`` over its (possibly partial, at the right/bottom edge) block of ``n`` pixels with
   ``M(n) = (uint32)((float)(1 << 24) / n)``,
3. Lanczos-3 resample of the reduced image with the fractional source box ``side / factor``: horizontal pass to a
   u8 intermediate, then vertical pass, 22-bit fixed-point coefficients.

Pin: every function here is checked against live Pillow on random images in ``tests/test_oracle.py``.
"""

from __future__ import annotations

import math

import numpy as np

PRECISION_BITS = 32 - 8 - 2


def thumbnail_size(width: int, height: int, size) -> tuple:
    """Pillow ``Image.thumbnail`` target size (aspect preserved); ``(width, height)`` if nothing to do."""
    x, y = math.floor(size[0]), math.floor(size[1])
    if x >= width and y >= height:
        return width, height

    def round_aspect(number, key):
        return max(min(math.floor(number), math.ceil(number), key=key), 1)

    aspect = width / height
    if x / y >= aspect:
        x = round_aspect(y * aspect, key=lambda n: abs(aspect - n / y))
    else:
        y = round_aspect(x / aspect, key=lambda n: 0 if n == 0 else abs(aspect - x / n))
    return x, y


def reduce_factors(width: int, height: int, out_w: int, out_h: int, reducing_gap: float = 2.0) -> tuple:
    return (int(width / out_w / reducing_gap) or 1, int(height / out_h / reducing_gap) or 1)


def _mult(n: int) -> int:
    return int(np.float32(1 << 24) / np.float32(n))


def reduce_u8(img: np.ndarray, fx: int, fy: int) -> np.ndarray:
    """Pillow ``Image.reduce((fx, fy))`` of an (H, W, C) u8 image; output ceil(H/fy) x ceil(W/fx)."""
    h, w, c = img.shape
    oh, ow = -(-h // fy), -(-w // fx)
    pad = np.zeros((oh * fy, ow * fx, c), dtype=np.int64)
    pad[:h, :w] = img
    ss = pad.reshape(oh, fy, ow, fx, c).sum(axis=(1, 3))
    ny = np.minimum(fy, h - np.arange(oh) * fy)
    nx = np.minimum(fx, w - np.arange(ow) * fx)
    n = ny[:, None] * nx[None, :]
    mult = np.vectorize(_mult)(n).astype(np.int64)
    return (((ss + (n // 2)[..., None]) * mult[..., None]) >> 24).astype(np.uint8)


def _lanczos(x: float) -> float:
    def sinc(v):
        if v == 0.0:
            return 1.0
        v = v * math.pi
        return math.sin(v) / v

    if -3.0 <= x < 3.0:
        return sinc(x) * sinc(x / 3.0)
    return 0.0


def lanczos_coeffs(in_size: int, in0: float, in1: float, out_size: int):
    """Pillow ``precompute_coeffs`` + ``normalize_coeffs_8bpc`` for the LANCZOS filter.

    Returns (ksize, lo[out], cnt[out], coef[out][ksize] int32)."""
    # Pillow passes the box as C floats: in0, in1 and their difference are rounded to fp32 before the double math
    f0, f1 = np.float32(in0), np.float32(in1)
    in0 = float(f0)
    scale = float(f1 - f0) / out_size
    filterscale = max(scale, 1.0)
    support = 3.0 * filterscale
    ksize = int(math.ceil(support)) * 2 + 1
    lo = np.zeros(out_size, np.int32)
    cnt = np.zeros(out_size, np.int32)
    coef = np.zeros((out_size, ksize), np.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = in0 + (xx + 0.5) * scale
        xmin = max(int(center - support + 0.5), 0)
        xmax = min(int(center + support + 0.5), in_size) - xmin
        k = [_lanczos((x + xmin - center + 0.5) * ss) for x in range(xmax)]
        ww = sum(k)
        for x in range(xmax):
            v = k[x] / ww if ww != 0.0 else k[x]
            q = v * (1 << PRECISION_BITS)
            coef[xx, x] = int(q - 0.5) if v < 0 else int(q + 0.5)
        lo[xx], cnt[xx] = xmin, xmax
    return ksize, lo, cnt, coef


def _clip8(v):
    return np.clip(v >> PRECISION_BITS, 0, 255)


def resample_lanczos_u8(img: np.ndarray, out_w: int, out_h: int, box) -> np.ndarray:
    """Pillow ``ImagingResample`` (8 bits per channel): horizontal pass over the rows the vertical pass needs,
    u8 intermediate, vertical pass.  ``box`` = (x0, y0, x1, y1) in source pixels (floats)."""
    h, w, c = img.shape
    _, xlo, xcnt, xk = lanczos_coeffs(w, box[0], box[2], out_w)
    _, ylo, ycnt, yk = lanczos_coeffs(h, box[1], box[3], out_h)
    src = img.astype(np.int64)
    tmp = np.zeros((h, out_w, c), np.int64)
    for xx in range(out_w):
        acc = np.full((h, c), 1 << (PRECISION_BITS - 1), np.int64)
        for t in range(xcnt[xx]):
            acc += src[:, xlo[xx] + t] * int(xk[xx, t])
        tmp[:, xx] = _clip8(acc)
    out = np.zeros((out_h, out_w, c), np.uint8)
    for yy in range(out_h):
        acc = np.full((out_w, c), 1 << (PRECISION_BITS - 1), np.int64)
        for t in range(ycnt[yy]):
            acc += tmp[ylo[yy] + t] * int(yk[yy, t])
        out[yy] = _clip8(acc)
    return out


def thumbnail_u8(rgb: np.ndarray, size=(2048, 2048)) -> np.ndarray:
    """(H, W, 3) u8 level-0 image -> the thumbnail ``Image.thumbnail(size, LANCZOS)`` produces."""
    h, w, _ = rgb.shape
    ow, oh = thumbnail_size(w, h, size)
    if (ow, oh) == (w, h):
        return rgb.copy()
    fx, fy = reduce_factors(w, h, ow, oh)
    red = reduce_u8(rgb, fx, fy) if (fx > 1 or fy > 1) else rgb
    return resample_lanczos_u8(red, ow, oh, (0.0, 0.0, w / fx, h / fy))
