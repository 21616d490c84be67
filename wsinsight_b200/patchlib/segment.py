"""Tissue segmentation on the B200: host-side mirror of ``wsinsight/patchlib/segment.py``.

``segment_tissue`` keeps the reference signature and defaults (segment.py:13-21) and the
reference's error behaviour (ValueError for an even / <=1 median size, segment.py:68-72;
assertion on the input shape, segment.py:49-50).  All arithmetic runs in
``wsi_seg_tissue_u8`` / ``wsi_seg_tissue_batch_u8`` (csrc/seg.cu); there is no CPU path.
"""

from __future__ import annotations

import ctypes as C
from typing import Sequence

import numpy as np

from .. import _cabi


def _check_median(median_filter_size: int) -> None:
    if median_filter_size <= 1 or median_filter_size % 2 == 0:
        raise ValueError(
            "median_filter_size must be greater than 1 and odd, but got"
            f" {median_filter_size}"
        )


#: footprint-origin conventions of ``binary_closing(ones(k, k))`` for even k: (dilation origin, erosion origin) as
#: functions of k; window offsets are ``[-origin .. k-1-origin]``.  They agree for odd k and, for even k, everywhere
#: except within k/2 pixels of the thumbnail border.
CLOSING_CONVENTIONS = {
    # scipy.ndimage on the raw even footprint (== a footprint end-padded to odd size): the default
    "scipy_raw": lambda k: ((k - 1) // 2, k // 2),
    # footprint front-padded to (k+1) x (k+1) before scipy sees it (skimage ``pad_footprint(pad_end=False)``)
    "front_padded": lambda k: (k // 2, (k - 1) // 2),
    # raw dilation, erosion with the mirrored footprint
    "mirrored_erosion": lambda k: ((k - 1) // 2, (k - 1) // 2),
}


def closing_origins(closing_kernel_size: int, convention="scipy_raw"):
    if isinstance(convention, str):
        if convention not in CLOSING_CONVENTIONS:
            raise ValueError(f"unknown closing convention {convention!r}; choose from {sorted(CLOSING_CONVENTIONS)}")
        return CLOSING_CONVENTIONS[convention](int(closing_kernel_size))
    d, e = convention
    return int(d), int(e)


def segment_tissue_device(
    thumbs: Sequence,
    median_filter_size: int = 7,
    binary_threshold: int | str = 7,
    closing_kernel_size: int = 6,
    min_object_size_px: int | Sequence[int] = 512,
    min_hole_size_px: int | Sequence[int] = 1024,
    want_outer: bool = True,
    stream=None,
    closing_convention="scipy_raw",
    return_info: bool = False,
):
    """Segment a cohort of device-resident thumbnails in one batched launch sequence.

    ``thumbs``: list of CUDA uint8 tensors (H, W, 3), contiguous.  Returns ``(masks, outers)``:
    lists of CUDA uint8 tensors (H, W) with values 0/1 (views into two packed buffers).
    ``outers`` is the mask restricted to outermost 8-connected components -- the region the
    reference's contour/polygon stage keeps (patch.py:78-123) and the input of the grid filter.

    ``binary_threshold="otsu"`` (extra mode, not upstream): per-slide Otsu threshold on the median-filtered
    saturation, identical to ``cv.threshold(S_med, 0, 255, THRESH_BINARY | THRESH_OTSU)``.
    ``closing_convention``: see :data:`CLOSING_CONVENTIONS`.  ``return_info=True`` appends a CUDA int32 ``(n, 4)``
    tensor ``[threshold used, mask has tissue, 0, 0]`` (no host synchronisation happens here).
    """
    torch = _cabi.require_cuda()
    _check_median(median_filter_size)
    n = len(thumbs)
    if n == 0:
        return [], []
    dev = thumbs[0].device
    for t in thumbs:
        assert t.ndim == 3 and t.shape[2] == 3 and t.dtype == torch.uint8 and t.is_cuda
    Hs = [int(t.shape[0]) for t in thumbs]
    Ws = [int(t.shape[1]) for t in thumbs]
    obj = [int(min_object_size_px)] * n if np.isscalar(min_object_size_px) else [int(v) for v in min_object_size_px]
    hole = [int(min_hole_size_px)] * n if np.isscalar(min_hole_size_px) else [int(v) for v in min_hole_size_px]
    # pack the cohort: rgb offsets 256-byte aligned, masks dense (the ABI requires dense masks)
    rgb_off, mask_off, ro, mo = [], [], 0, 0
    for h, w in zip(Hs, Ws):
        rgb_off.append(ro)
        mask_off.append(mo)
        ro += (h * w * 3 + 255) // 256 * 256
        mo += h * w
    if n == 1 and thumbs[0].is_contiguous():
        rgb = thumbs[0]
    else:
        rgb = torch.empty(ro, dtype=torch.uint8, device=dev)
        for t, o, h, w in zip(thumbs, rgb_off, Hs, Ws):
            rgb[o:o + h * w * 3].view(h, w, 3).copy_(t)
    mask = torch.empty(mo, dtype=torch.uint8, device=dev)
    outer = torch.empty(mo, dtype=torch.uint8, device=dev) if want_outer else None
    lib = _cabi.load()
    Ha, Wa = _cabi.int_array(Hs), _cabi.int_array(Ws)
    ws_bytes = lib.wsi_seg_batch_workspace_bytes(n, Ha, Wa)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    info = torch.empty((n, 4), dtype=torch.int32, device=dev)
    if isinstance(binary_threshold, str):
        if binary_threshold.lower() != "otsu":
            raise ValueError(f"binary_threshold must be an integer or 'otsu', got {binary_threshold!r}")
        thr = -1                                    # WSI_SEG_THRESHOLD_OTSU
    else:
        thr = int(binary_threshold)
        if thr < 0:
            raise ValueError("binary_threshold must be >= 0")
    d_org, e_org = closing_origins(closing_kernel_size, closing_convention)
    with torch.cuda.device(dev):
        _cabi.call(
            "wsi_seg_tissue_batch_ex", rgb.data_ptr(), n, Ha, Wa, _cabi.int_array([3 * w for w in Ws]),
            _cabi.ll_array(rgb_off), _cabi.ll_array(mask_off), int(median_filter_size), thr,
            int(closing_kernel_size), d_org, e_org, _cabi.int_array(obj), _cabi.int_array(hole), mask.data_ptr(),
            outer.data_ptr() if want_outer else None, info.data_ptr(), ws.data_ptr(), ws_bytes,
            _cabi.stream_ptr(stream),
        )
    masks = [mask[o:o + h * w].view(h, w) for o, h, w in zip(mask_off, Hs, Ws)]
    outers = [outer[o:o + h * w].view(h, w) for o, h, w in zip(mask_off, Hs, Ws)] if want_outer else None
    if return_info:
        return masks, outers, info
    return masks, outers


def segment_tissue(
    im_arr,
    median_filter_size: int = 7,
    binary_threshold: int | str = 7,
    closing_kernel_size: int = 6,
    min_object_size_px: int = 512,
    min_hole_size_px: int = 1024,
    closing_convention="scipy_raw",
):
    """Drop-in for ``wsinsight.patchlib.segment.segment_tissue`` (segment.py:13-97).

    ``im_arr``: (H, W, 3) uint8 RGB thumbnail (NumPy).  Returns a boolean (H, W) NumPy array.
    """
    torch = _cabi.require_cuda()
    im_arr = np.asarray(im_arr)
    assert im_arr.ndim == 3
    assert im_arr.shape[2] == 3
    _check_median(median_filter_size)
    t = torch.from_numpy(np.ascontiguousarray(im_arr, dtype=np.uint8)).cuda()
    masks, _ = segment_tissue_device(
        [t], median_filter_size, binary_threshold, closing_kernel_size, min_object_size_px,
        min_hole_size_px, want_outer=False, closing_convention=closing_convention,
    )
    return masks[0].cpu().numpy().astype(bool)
