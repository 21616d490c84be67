"""Macenko stain normalisation on the patch data path.

The reference (``wsinsight/modellib/run_inference.py:232-264``, ``data.py:293-300``) estimates a slide's stain matrix once
from one shuffled batch of 256 patches with HistomicsTK (``rgb_separate_stains_macenko_pca``) and then normalises every
patch with ``deconvolution_based_normalization(patch + EPSILON, W_source=W_est, W_target=W_def)`` before the transform.

Here the once-per-slide estimation is host numpy on a 256-patch sample gathered from the device-resident slide (a 3 x N
SVD, a percentile pick), and the per-patch normalisation -- the part that touches every pixel -- is one kernel over the
gathered patches (``wsi_stain_normalize_rgb``), after which classification runs from the patch stack like the ``/images``
cache path.  HistomicsTK is not installable offline: its algorithm is restated here (and, independently laid out as the
library's call chain, in ``oracle/stain.py``); **parity with HistomicsTK itself is unpinned**.
"""

from __future__ import annotations

import ctypes as C

import numpy as np

from .. import _cabi

EPSILON = 1e-8
I_0 = 255

stain_color_map = {
    "hematoxylin": [0.65, 0.70, 0.29],
    "eosin": [0.07, 0.99, 0.11],
    "dab": [0.27, 0.57, 0.78],
    "null": [0.0, 0.0, 0.0],
}


def default_target_matrix() -> np.ndarray:
    """``W_def`` of run_inference.py:262-264: columns eosin, hematoxylin, null."""
    return np.array([stain_color_map[s] for s in ("eosin", "hematoxylin", "null")], dtype=np.float64).T


def _unit_columns(m: np.ndarray) -> np.ndarray:
    n = np.linalg.norm(m, axis=0)
    return m / np.where(n == 0, 1.0, n)


def _complement(w: np.ndarray) -> np.ndarray:
    third = np.cross(w[:, 0], w[:, 1])
    return np.stack([w[:, 0], w[:, 1], third / np.linalg.norm(third)], axis=1)


def macenko_stain_matrix(pixels_rgb: np.ndarray, I_0: float = I_0, minimum_magnitude: float = 16,
                         min_angle_percentile: float = 0.01, max_angle_percentile: float = 0.99) -> np.ndarray:
    """Stain matrix (3, 3) of a pixel sample (..., 3) by Macenko's PCA method as HistomicsTK implements it: SDA
    transform with white ``I_0``, the plane of the first two left singular vectors, pixels above a magnitude threshold,
    the 1st / 99th percentile "angle" directions, unit columns, third column = normalised cross product."""
    x = np.maximum(np.asarray(pixels_rgb, dtype=np.float64).reshape(-1, 3).T, 1e-10)
    sda = np.maximum(-np.log(x / float(I_0)) * 255.0 / np.log(float(I_0)), 0.0)
    sda = sda[:, np.isfinite(sda).all(axis=0)]
    pcs = np.linalg.svd(sda, full_matrices=False)[0]
    proj = pcs.T[:-1] @ sda
    keep = proj[:, np.linalg.norm(proj, axis=0) > minimum_magnitude]
    unit = _unit_columns(keep)
    angles = (1.0 - unit[1]) * np.sign(unit[0])

    def direction(p):
        i = int(p * angles.size + 0.5)
        return pcs[:, :-1] @ keep[:, np.argpartition(angles, i)[i]]

    w = np.stack([direction(min_angle_percentile), direction(max_angle_percentile)], axis=1)
    return _complement(_unit_columns(w))


def normalization_matrix(W_source: np.ndarray, W_target: np.ndarray) -> np.ndarray:
    """M with ``sda_out = M @ sda_in``: ``W_target @ inv(unit-column, complemented W_source)``."""
    ws = np.asarray(W_source, dtype=np.float64)
    wc = _complement(ws) if np.linalg.norm(ws[:, 2]) <= 1e-16 else ws
    return np.asarray(W_target, dtype=np.float64) @ np.linalg.inv(_unit_columns(wc))


def normalize_patches_device(images, W_source, W_target, out=None):
    """uint8 CUDA tensor (..., 3) of RGB pixels -> stain-normalised uint8 tensor of the same shape (one kernel)."""
    torch = _cabi.require_cuda()
    if images.dtype != torch.uint8 or images.shape[-1] != 3 or not images.is_cuda:
        raise ValueError("images must be a CUDA uint8 tensor with 3 channels last")
    images = images.contiguous()
    out = torch.empty_like(images) if out is None else out
    M = np.ascontiguousarray(normalization_matrix(W_source, W_target), dtype=np.float64)
    with torch.cuda.device(images.device):
        _cabi.call("wsi_stain_normalize_rgb", images.data_ptr(), out.data_ptr(), images.numel() // 3,
                   M.ctypes.data_as(C.POINTER(C.c_double)), _cabi.stream_ptr())
    return out


def estimate_slide_stain_matrix(slide, coords_xywh: np.ndarray, n_patches: int = 256, seed=None) -> np.ndarray:
    """``W_est`` of one slide (run_inference.py:241-261): a random sample of up to 256 patches, concatenated, + EPSILON."""
    from ..patchlib.io import extract_patches_from_slide

    n = len(coords_xywh)
    rng = np.random.default_rng(seed)
    idx = np.sort(rng.choice(n, size=min(n, int(n_patches)), replace=False))
    ps = int(coords_xywh[0, 2])
    px = extract_patches_from_slide(slide, coords_xywh[idx, :2], ps, as_numpy=True)
    return macenko_stain_matrix(px.astype(np.float32) + np.float32(EPSILON), I_0)
