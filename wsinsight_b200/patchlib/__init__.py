"""Tissue segmentation + patchify (mirror of ``wsinsight/patchlib``)."""

from .patch import TissueRegion, get_multipolygon_from_binary_arr, get_patch_coordinates_within_polygon  # noqa: F401
from .pipeline import (  # noqa: F401
    segment_and_patch_directory_of_slides,
    segment_and_patch_one_slide,
    segment_and_patch_slides_device,
)
from .segment import segment_tissue, segment_tissue_device  # noqa: F401
