"""Mask -> tissue region -> in-tissue patch grid on the B200.

Host-side mirror of ``wsinsight/patchlib/patch.py``.  The reference turns the mask into a
Shapely MultiPolygon (``get_multipolygon_from_binary_arr``, patch.py:35-130) and keeps grid
centroids with ``polygon.contains`` (``get_patch_coordinates_within_polygon``, patch.py:174-242).
Here the "polygon" is a :class:`TissueRegion`: the mask restricted to its outermost
8-connected components plus the slide dimensions; the exact integer raster rule of
``wsi_patch_grid`` (csrc/seg.cu) answers the same ``contains`` question without
tracing contours.  Function names, argument order and error behaviour follow the reference.
"""

from __future__ import annotations

import ctypes as C
import dataclasses

import numpy as np

from .. import _cabi


@dataclasses.dataclass
class TissueRegion:
    """Stand-in for the reference's Shapely polygon (device-resident)."""

    outer: object            # CUDA uint8 (Ht, Wt): outermost-component mask
    slide_width: int
    slide_height: int

    @property
    def thumb_size(self):
        return int(self.outer.shape[1]), int(self.outer.shape[0])


def outermost_components(mask_u8, stream=None):
    """Outermost-component restriction of a device mask (CUDA uint8 (H, W), values 0/1)."""
    torch = _cabi.require_cuda()
    lib = _cabi.load()
    h, w = int(mask_u8.shape[0]), int(mask_u8.shape[1])
    outer = torch.empty_like(mask_u8)
    ws_bytes = lib.wsi_seg_workspace_bytes(h, w)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=mask_u8.device)
    with torch.cuda.device(mask_u8.device):
        _cabi.call("wsi_outermost_u8", mask_u8.data_ptr(), h, w, outer.data_ptr(), ws.data_ptr(), ws_bytes,
                   _cabi.stream_ptr(stream))
    return outer


def get_multipolygon_from_binary_arr(arr, scale=None, *, slide_dims=None):
    """Reference name (patch.py:35-130).  ``arr``: (Ht, Wt) uint8/bool mask (NumPy or CUDA tensor).

    Returns ``None`` when the mask has no tissue (patch.py:61-62), else
    ``(TissueRegion, None, None)`` -- contours/hierarchy (used upstream only for the overlay
    JPEG) are not produced on this path.  ``scale`` is (sx, sy) as upstream; integer slide
    dimensions are what the exact arithmetic needs, so pass ``slide_dims=(W, H)`` when known.
    """
    torch = _cabi.require_cuda()
    t = torch.from_numpy(np.ascontiguousarray(arr)).cuda() if isinstance(arr, np.ndarray) else arr
    t = (t != 0).to(torch.uint8).contiguous()
    ht, wt = int(t.shape[0]), int(t.shape[1])
    if slide_dims is None:
        slide_dims = (wt, ht) if scale is None else (int(round(scale[0] * wt)), int(round(scale[1] * ht)))
    if not bool(t.any().item()):
        return None
    return TissueRegion(outermost_components(t), int(slide_dims[0]), int(slide_dims[1])), None, None


def patch_grid_device(region: TissueRegion, patch_size: int, half_patch_size: int, overlap: float = 0.0,
                      stream=None, sync: bool = True):
    """Device-side grid filter.  Returns ``(coords int32 CUDA (N,2), tile_dim (n_i, n_j))``.

    ``sync=False`` skips the device->host read of the patch count: the result is then
    ``(coords_capacity, count_tensor, tile_dim)`` with ``coords_capacity`` sized for every candidate and only its
    first ``count`` rows valid, so a cohort loop can enqueue every slide before it reads any count."""
    torch = _cabi.require_cuda()
    if overlap >= 1:
        raise ValueError(f"overlap must be in (-inf, 1) but got {overlap}")
    step = round((1 - overlap) * patch_size)
    lib = _cabi.load()
    outer = region.outer
    dev = outer.device
    W, H = int(region.slide_width), int(region.slide_height)
    half = int(half_patch_size)
    ni = len(range(half, W, step))
    nj = len(range(half, H, step))
    cap = max(1, ni * nj)
    coords = torch.empty((cap, 2), dtype=torch.int32, device=dev)
    count = torch.zeros(1, dtype=torch.int32, device=dev)
    ws_bytes = max(int(lib.wsi_patch_grid_workspace_bytes(W, H, half, int(step))), 256)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    td = (C.c_int * 2)()
    with torch.cuda.device(dev):
        _cabi.call(
            "wsi_patch_grid", outer.data_ptr(), int(outer.shape[0]), int(outer.shape[1]), W, H, half, int(step),
            coords.data_ptr(), cap, count.data_ptr(), td, ws.data_ptr(), ws_bytes, _cabi.stream_ptr(stream),
        )
    if not sync:
        return coords, count, (int(td[0]), int(td[1]))
    n = int(count.item())
    return coords[:n], (int(td[0]), int(td[1]))


def get_patch_coordinates_within_polygon(
    slide_width: int,
    slide_height: int,
    patch_size: int,
    half_patch_size: int,
    polygon: TissueRegion,
    overlap: float = 0.0,
):
    """Drop-in for patch.py:174-242: (N, 2) int64 top-left coordinates, x-major / y fastest."""
    if overlap >= 1:
        raise ValueError(f"overlap must be in (-inf, 1) but got {overlap}")
    if (int(slide_width), int(slide_height)) != (polygon.slide_width, polygon.slide_height):
        polygon = TissueRegion(polygon.outer, int(slide_width), int(slide_height))
    coords, _ = patch_grid_device(polygon, patch_size, half_patch_size, overlap)
    return coords.cpu().numpy().astype(np.int64)
