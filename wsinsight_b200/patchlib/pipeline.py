"""Per-slide segmentation + patchify orchestration (default branch of the reference only).

Mirror of ``wsinsight/patchlib/pipeline.py``: ``segment_and_patch_one_slide`` (:45-446, default
branch :357-402) and ``segment_and_patch_directory_of_slides`` (:449-508).  The QuPath / StarDist /
HistoQC branches are cell-model workflows outside the hot path and raise ``NotImplementedError``.
Scalar arithmetic (patch size, thumbnail mpp, um^2 -> px thresholds, banker's rounding) follows the
reference line by line; pixel work is one batched kernel sequence over the whole cohort.
"""

from __future__ import annotations

import dataclasses
import logging
from pathlib import Path
from typing import List, Sequence

import numpy as np

from .. import _cabi
from .io import patch_file_exists, save_hdf5, square_polygons
from .patch import TissueRegion, patch_grid_device
from .segment import segment_tissue_device

logger = logging.getLogger(__name__)

PATCHES_DIR = "patches"
MASKS_DIR = "masks"


@dataclasses.dataclass
class SlidePatchResult:
    """What one slide's patchify produced (device-resident until written)."""

    name: str
    patch_size: int
    half_patch_size: int
    min_object_size_px: int
    min_hole_size_px: int
    mask: object = None          # CUDA uint8 (Ht, Wt)
    coords: object = None        # CUDA int32 (N, 2), None if no tissue
    tile_dim: tuple = None
    threshold: int = None        # binary threshold used (the Otsu value in "otsu" mode)


def slide_scalars(slide, patch_size_px, patch_spacing_um_px, thumbsize, min_object_size_um2, min_hole_size_um2):
    """pipeline.py:96-112,359: patch_size, half, thumb_mpp and the px thresholds."""
    from ..wsi import _thumb_size

    mpp = slide.mpp
    patch_size = int(round(patch_size_px * patch_spacing_um_px / mpp))
    if len(thumbsize) != 2:
        raise ValueError(f"Length of 'thumbsize' must be 2 but got {len(thumbsize)}")
    tw, th = _thumb_size(slide.dimensions, tuple(thumbsize))
    thumb_mpp = (mpp * (np.array(slide.dimensions) / np.array((tw, th)))).mean()
    sq = thumb_mpp ** 2
    return dict(patch_size=patch_size, half_patch_size=round(patch_size / 2), thumb_mpp=float(thumb_mpp),
                min_object_size_px=round(min_object_size_um2 / sq), min_hole_size_px=round(min_hole_size_um2 / sq))


def segment_and_patch_slides_device(
    slides: Sequence,
    patch_size_px: int,
    patch_spacing_um_px: float,
    thumbsize=(2048, 2048),
    median_filter_size: int = 7,
    binary_threshold: int = 7,
    closing_kernel_size: int = 6,
    min_object_size_um2: float = 200 ** 2,
    min_hole_size_um2: float = 190 ** 2,
    overlap: float = 0.0,
    thumbs=None,
    closing_convention="scipy_raw",
) -> List[SlidePatchResult]:
    """Cohort-level hot path: thumbnails -> masks -> outermost regions -> in-tissue grids.

    ``thumbs`` (optional) are precomputed CUDA uint8 (Ht, Wt, 3) thumbnails; otherwise
    ``slide.thumbnail_device(thumbsize)`` is used.  Nothing is copied to the host here except the
    per-slide patch count.
    """
    if overlap >= 1:
        raise ValueError(f"overlap must be in (-inf, 1) but got {overlap}")
    sc = [slide_scalars(s, patch_size_px, patch_spacing_um_px, thumbsize, min_object_size_um2, min_hole_size_um2)
          for s in slides]
    if thumbs is None:
        thumbs = [s.thumbnail_device(tuple(thumbsize)) for s in slides]
    masks, outers, info = segment_tissue_device(
        thumbs, median_filter_size, binary_threshold, closing_kernel_size,
        [c["min_object_size_px"] for c in sc], [c["min_hole_size_px"] for c in sc], want_outer=True,
        closing_convention=closing_convention, return_info=True)
    # every slide's grid filter is enqueued before anything is read back: ONE device->host copy for the cohort
    torch = _cabi.require_cuda()
    results, pending = [], []
    for s, c, m, o in zip(slides, sc, masks, outers):
        r = SlidePatchResult(s.name, c["patch_size"], c["half_patch_size"], c["min_object_size_px"],
                             c["min_hole_size_px"], mask=m)
        region = TissueRegion(o, s.width, s.height)
        coords, count, tile_dim = patch_grid_device(region, c["patch_size"], c["half_patch_size"], overlap, sync=False)
        r.tile_dim = tile_dim
        pending.append((coords, count))
        results.append(r)
    counts = torch.cat([cnt for _, cnt in pending]).cpu().tolist() if pending else []
    flags = info.cpu().numpy()
    for r, (coords, _), n, f in zip(results, pending, counts, flags):
        r.threshold = int(f[0])
        # "no contours" in the reference (patch.py:61-62) == empty mask -> None, no files
        r.coords = coords[: int(n)] if int(f[1]) else None
    return results


def draw_contours_on_thumbnail(thumb_rgb: np.ndarray, mask: np.ndarray) -> np.ndarray:
    """Overlay of the reference's mask JPEG (io.py:146-166 + patch.py:57-60): the tissue mask's
    ``cv.findContours(RETR_CCOMP, CHAIN_APPROX_SIMPLE)`` contours drawn 7 px wide on the thumbnail, outer
    boundaries (no parent in the hierarchy) in (0, 255, 255), hole boundaries in (255, 255, 0).
    An exporter, not part of the parity surface of the kernels: host OpenCV on the device-produced mask."""
    import cv2 as cv

    img = np.ascontiguousarray(thumb_rgb).copy()
    contours, hierarchy = cv.findContours(np.ascontiguousarray(mask, dtype=np.uint8) * 255, cv.RETR_CCOMP,
                                          cv.CHAIN_APPROX_SIMPLE)
    if hierarchy is None:
        return img
    parent = hierarchy[0, :, 3]
    outer = [c for c, p in zip(contours, parent) if p < 0]
    holes = [c for c, p in zip(contours, parent) if p >= 0]
    cv.drawContours(img, outer, -1, (0, 255, 255), 7)
    cv.drawContours(img, holes, -1, (255, 255, 0), 7)
    return img


def _write_mask_jpeg(mask_path: Path, thumb_dev, mask_dev) -> None:
    """``masks/<stem>.jpg``: contour overlay, then ``thumbnail((1024, 1024), LANCZOS)`` (pipeline.py:439-444)."""
    from PIL import Image

    img = Image.fromarray(draw_contours_on_thumbnail(thumb_dev.cpu().numpy(), mask_dev.cpu().numpy())).convert("RGB")
    img.thumbnail((1024, 1024), resample=Image.Resampling.LANCZOS)
    mask_path.parent.mkdir(exist_ok=True, parents=True)
    with mask_path.open("wb") as fh:
        img.save(fh)


def segment_and_patch_one_slide(
    slide_path,
    save_dir,
    qupath_detection_dir=None,
    qupath_geojson_detection_dir=None,
    qupath_geojson_annotation_dir=None,
    patch_size_px: int = None,
    patch_spacing_um_px: float = None,
    halo_size_px: int = 0,
    histoqc_dir=None,
    thumbsize=(2048, 2048),
    median_filter_size: int = 7,
    binary_threshold: int = 7,
    closing_kernel_size: int = 6,
    min_object_size_um2: float = 200 ** 2,
    min_hole_size_um2: float = 190 ** 2,
    overlap: float = 0.0,
    object_based: bool = False,
    object_detection=None,
    stardist_normalization_pmin: float = 1.0,
    stardist_normalization_pmax: float = 99.8,
    cache_image_patches: bool = False,
    slide=None,
    closing_convention="scipy_raw",
) -> None:
    """Reference signature (pipeline.py:45-67).  ``slide`` is the opened
    :class:`~wsinsight_b200.wsi.DeviceSlide`; ``slide_path`` names it (stem = slide id)."""
    if qupath_detection_dir or qupath_geojson_detection_dir or qupath_geojson_annotation_dir or histoqc_dir \
            or object_based or object_detection or cache_image_patches:
        raise NotImplementedError("QuPath / HistoQC / object-based / cached-patch branches are outside the B200 "
                                  "hot path (reference pipeline.py:169-356)")
    if slide is None:
        raise ValueError("a DeviceSlide must be supplied (file decoders are out of scope)")
    slide_prefix = Path(str(slide_path)).stem
    save_dir = Path(save_dir)
    patch_path = save_dir / PATCHES_DIR / f"{slide_prefix}.h5"
    mask_path = save_dir / MASKS_DIR / f"{slide_prefix}.jpg"
    if patch_file_exists(patch_path) and mask_path.exists():
        logger.info("Patch output and mask output files already exist")
        return None
    res = segment_and_patch_slides_device(
        [slide], patch_size_px, patch_spacing_um_px, thumbsize, median_filter_size, binary_threshold,
        closing_kernel_size, min_object_size_um2, min_hole_size_um2, overlap,
        closing_convention=closing_convention)[0]
    if res.coords is None:
        logger.warning(f"No tissue was found in slide {slide_path}")
        return None
    coords = res.coords.cpu().numpy()
    logger.info(f"Found {len(coords)} patches within tissue")
    if coords.size > 0:
        save_hdf5(path=patch_path, coords=coords, polygons=square_polygons(coords, res.patch_size),
                  tile_dim=np.asarray(res.tile_dim, dtype=np.int32), patch_size=res.patch_size,
                  patch_spacing_um_px=patch_spacing_um_px, compression="gzip", images=None,
                  slide_path=str(slide_path), slide_mpp=slide.mpp, slide_width=slide.width,
                  slide_height=slide.height)
    else:
        logger.warning(f"No patches found for slide {slide_path}")
    _write_mask_jpeg(mask_path, slide.thumbnail_device(tuple(thumbsize)), res.mask)
    return None


def segment_and_patch_directory_of_slides(
    wsi_dir,
    slide_paths,
    save_dir,
    qupath_detection_dir=None,
    qupath_geojson_detection_dir=None,
    qupath_geojson_annotation_dir=None,
    patch_size_px: int = None,
    patch_spacing_um_px: float = None,
    slides=None,
    **kwargs,
) -> None:
    """Reference loop (pipeline.py:449-508): per-slide try/except, failures logged and skipped."""
    slides = slides or {}
    for i, slide_path in enumerate(slide_paths):
        logger.info(f"Slide {i + 1} of {len(slide_paths)} ({(i + 1) / len(slide_paths):.2%})")
        try:
            segment_and_patch_one_slide(
                slide_path=slide_path, save_dir=save_dir, qupath_detection_dir=qupath_detection_dir,
                qupath_geojson_detection_dir=qupath_geojson_detection_dir,
                qupath_geojson_annotation_dir=qupath_geojson_annotation_dir, patch_size_px=patch_size_px,
                patch_spacing_um_px=patch_spacing_um_px, slide=slides.get(str(slide_path)), **kwargs)
        except Exception as e:  # noqa: BLE001 - logged for operators, as upstream
            logger.error(f"Failed to segment and patch slide\n{slide_path}", exc_info=e)
    return None
