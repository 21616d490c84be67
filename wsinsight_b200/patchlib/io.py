"""Patch-file contract between ``patch`` and ``infer`` (mirror of ``wsinsight/patchlib/io.py``).

The reference writes ``results_dir/patches/<stem>.h5`` with ``save_hdf5`` (io.py:51-143) and reads
it back in ``_read_patch_coords`` (``modellib/data.py:22-63``) and ``run_inference.py:207-213``.
``h5py`` is not installable in this image, so the writer emits the reference's HDF5 layout when
``h5py`` imports and an ``.npz`` mirror with the same fields otherwise (same stem, ``.h5.npz``);
:func:`load_patch_file` reads whichever exists.  Host-side formatting only: no kernels here.
"""

from __future__ import annotations

import logging
from pathlib import Path
from typing import List, Optional

import numpy as np

logger = logging.getLogger(__name__)

try:  # pragma: no cover - depends on the image
    import h5py
except Exception:  # noqa: BLE001
    h5py = None


def extract_patches_from_slide(slide, coords, patch_size: int, as_numpy: bool = False):
    """RGB patches ``(N, ps, ps, 3)`` uint8 for top-left ``coords`` (N, 2) at level 0 -- the optional ``/images``
    cache of the patch file (reference ``patchlib/io.py:19-48``: ``read_region`` + ``convert("RGB")`` per patch,
    zero outside the slide).  One gather kernel over the device-resident slide; returns a CUDA tensor, or a numpy
    array with ``as_numpy=True`` (what ``save_hdf5(images=...)`` stores)."""
    from .. import _cabi

    torch = _cabi.require_cuda()
    c = np.asarray(coords.cpu().numpy() if hasattr(coords, "cpu") else coords, dtype=np.int32)
    if c.ndim != 2 or c.shape[1] != 2:
        raise ValueError(f"coords must be (N, 2), got {c.shape}")
    dev = slide.rgba.device
    dc = torch.from_numpy(np.ascontiguousarray(c)).to(dev)
    out = torch.empty((len(c), int(patch_size), int(patch_size), 3), dtype=torch.uint8, device=dev)
    if len(c):
        with torch.cuda.device(dev):
            _cabi.call("wsi_extract_patches_rgb", slide.rgba.data_ptr(), slide.width, slide.height, slide.pitch_bytes,
                       dc.data_ptr(), len(c), int(patch_size), out.data_ptr(), _cabi.stream_ptr())
    return out.cpu().numpy() if as_numpy else out


def square_polygons(coords: np.ndarray, patch_size: int) -> np.ndarray:
    """Per-patch closed 5-vertex squares, vectorised (reference loop: pipeline.py:386-400).

    Returns float32 (5N, 2): [minx,miny],[maxx,miny],[maxx,maxy],[minx,maxy],[minx,miny] with
    max = min + patch_size - 1.
    """
    c = np.asarray(coords, dtype=np.int64)
    mn_x, mn_y = c[:, 0], c[:, 1]
    mx_x, mx_y = mn_x + patch_size - 1, mn_y + patch_size - 1
    poly = np.stack([
        np.stack([mn_x, mn_y], 1), np.stack([mx_x, mn_y], 1), np.stack([mx_x, mx_y], 1),
        np.stack([mn_x, mx_y], 1), np.stack([mn_x, mn_y], 1)], axis=1)
    return poly.reshape(-1, 2).astype(np.float32)


def save_hdf5(
    path,
    coords,
    polygons: Optional[List[np.ndarray]],
    tile_dim,
    patch_size: int,
    patch_spacing_um_px: float,
    compression: Optional[str] = "gzip",
    images=None,
    slide_path: Optional[str] = None,
    slide_mpp: Optional[float] = None,
    slide_width: Optional[float] = None,
    slide_height: Optional[float] = None,
) -> Path:
    """Same signature and validation as io.py:51-143.  ``polygons`` may also be the packed
    (5N, 2) float32 array of :func:`square_polygons`.  Returns the path actually written."""
    coords = np.asarray(coords, dtype=np.int32)
    if coords.ndim != 2:
        raise ValueError(f"coords must have 2 dimensions but got {coords.ndim}")
    if coords.shape[1] != 2:
        raise ValueError(f"length of coords second axis must be 2 but got {coords.shape[1]}")
    if tile_dim is not None:
        tile_dim = np.asarray(tile_dim)
        if tile_dim.shape != (2,):
            raise ValueError(f"tile_dim must be (2,) but got {tile_dim.shape}")
    if images is not None:
        images = np.asarray(images, dtype=np.uint8)
        if images.shape[0] != coords.shape[0]:
            raise ValueError(
                f"images and coords must have same length; got {images.shape[0]} vs {coords.shape[0]}")
    poly_coords = offsets = None
    if polygons is not None and len(polygons) > 0:
        if isinstance(polygons, np.ndarray) and polygons.ndim == 2:
            poly_coords = polygons.astype(np.float32)
            offsets = np.arange(0, len(poly_coords) + 1, 5, dtype=np.int64)
        else:
            lengths = np.array([xy.shape[0] for xy in polygons], dtype=np.int64)
            offsets = np.concatenate(([0], np.cumsum(lengths)))
            poly_coords = np.vstack(polygons).astype(np.float32) if lengths.sum() > 0 else np.zeros((0, 2), np.float32)
    path = Path(path)
    path.parent.mkdir(parents=True, exist_ok=True)
    if h5py is not None:  # pragma: no cover - h5py absent in this image
        with h5py.File(path, "w") as f:
            g_slide = f.create_group("slide")
            if slide_path is not None:
                g_slide.attrs.create("slide_path", slide_path, dtype=h5py.string_dtype(encoding="utf-8"))
            if slide_mpp is not None:
                g_slide.attrs["slide_mpp"] = slide_mpp
            if slide_width is not None:
                g_slide.attrs["slide_width"] = slide_width
            if slide_height is not None:
                g_slide.attrs["slide_height"] = slide_height
            d = f.create_dataset("/coords", data=coords, compression=compression)
            d.attrs["patch_size"] = patch_size
            d.attrs["patch_level"] = 0
            d.attrs["patch_spacing_um_px"] = patch_spacing_um_px
            if tile_dim is not None:
                d.attrs["tile_dim"] = tile_dim
            if images is not None:
                f.create_dataset("/images", data=images, compression=compression, chunks=True)
            if poly_coords is not None:
                g = f.create_group("/polygons")
                dp = g.create_dataset("coords", data=poly_coords, dtype="float32", compression=compression,
                                      shuffle=True, chunks=True)
                g.create_dataset("offsets", data=offsets, dtype="int64")
                g.attrs["layout"] = "ragged_offsets"
                dp.attrs["columns"] = np.array(["x", "y"], dtype="S1")
        return path
    out = Path(str(path) + ".npz")
    fields = dict(coords=coords, patch_size=np.int64(patch_size), patch_level=np.int64(0),
                  patch_spacing_um_px=np.float64(patch_spacing_um_px))
    if tile_dim is not None:
        fields["tile_dim"] = np.asarray(tile_dim, dtype=np.int32)
    if slide_path is not None:
        fields["slide_path"] = np.array(str(slide_path))
    if slide_mpp is not None:
        fields["slide_mpp"] = np.float64(slide_mpp)
    if slide_width is not None:
        fields["slide_width"] = np.int64(slide_width)
    if slide_height is not None:
        fields["slide_height"] = np.int64(slide_height)
    if images is not None:
        fields["images"] = images
    if poly_coords is not None:
        fields["polygons_coords"] = poly_coords
        fields["polygons_offsets"] = offsets
    np.savez_compressed(out, **fields)
    return out


def patch_file_exists(path) -> bool:
    path = Path(path)
    return path.exists() or Path(str(path) + ".npz").exists()


def load_patch_file(path) -> dict:
    """Read a patch file written by :func:`save_hdf5` (either container) into a dict."""
    path = Path(path)
    if path.suffix == ".npz" or (not path.exists() and Path(str(path) + ".npz").exists()):
        p = path if path.suffix == ".npz" else Path(str(path) + ".npz")
        with np.load(p, allow_pickle=False) as z:
            d = {k: z[k] for k in z.files}
        for k in ("patch_size", "patch_level", "slide_width", "slide_height"):
            if k in d:
                d[k] = int(d[k])
        for k in ("patch_spacing_um_px", "slide_mpp"):
            if k in d:
                d[k] = float(d[k])
        if "slide_path" in d:
            d["slide_path"] = str(d["slide_path"])
        return d
    if h5py is None:
        raise ImportError(f"{path} is HDF5 but h5py is not installed")
    with h5py.File(path, "r") as f:  # pragma: no cover
        d = dict(coords=f["/coords"][()])
        for k, v in f["/coords"].attrs.items():
            d[k] = v
        for k, v in f["/slide"].attrs.items():
            d[k] = v
        return d
