"""Oracle for Macenko stain normalisation (TEST INFRASTRUCTURE ONLY) -- **parity unpinned**.

The reference calls HistomicsTK (``histomicstk``, not importable here and not in the wheelhouse):

    run_inference.py:232-264   W_est = htk...color_deconvolution.rgb_separate_stains_macenko_pca(batch + EPSILON, I_0)
                               W_def = columns of htk...stain_color_map for ['eosin', 'hematoxylin', 'null']
    data.py:293-300            patch = htk...color_normalization.deconvolution_based_normalization(
                                           patch + EPSILON, W_source=W_est, W_target=W_def)  ->  astype(uint8)

The functions below restate HistomicsTK 1.x's published algorithm function by function (``rgb_to_sda``, ``sda_to_rgb``,
``complement_stain_matrix``, ``color_deconvolution``, ``color_convolution``, ``separate_stains_macenko_pca``) as
remembered from its source; no HistomicsTK output was available to pin them, so every test built on this file checks the
CUDA kernel against THIS restatement only.
"""

from __future__ import annotations

import numpy as np

EPSILON = 1e-8
I_0 = 255

STAIN_COLOR_MAP = {
    "hematoxylin": [0.65, 0.70, 0.29],
    "eosin": [0.07, 0.99, 0.11],
    "dab": [0.27, 0.57, 0.78],
    "null": [0.0, 0.0, 0.0],
}


def default_target_matrix() -> np.ndarray:
    """run_inference.py:262-264: stains = ['eosin', 'hematoxylin', 'null'] as columns."""
    return np.array([STAIN_COLOR_MAP[s] for s in ("eosin", "hematoxylin", "null")], dtype=np.float64).T


def rgb_to_sda(im_rgb, I_0, allow_negatives=False):
    is_matrix = im_rgb.ndim == 2
    if is_matrix:
        im_rgb = im_rgb.T
    if I_0 is None:                       # the optical-density form: shift by one, white = 256
        im_rgb = im_rgb.astype(float) + 1
        I_0 = 256
    im_rgb = np.maximum(im_rgb, 1e-10)
    im_sda = -np.log(im_rgb / (1.0 * I_0)) * 255 / np.log(I_0)
    if not allow_negatives:
        im_sda = np.maximum(im_sda, 0)
    return im_sda.T if is_matrix else im_sda


def sda_to_rgb(im_sda, I_0):
    is_matrix = im_sda.ndim == 2
    if is_matrix:
        im_sda = im_sda.T
    od = I_0 is None
    if od:
        I_0 = 256
    im_rgb = I_0 ** (1 - im_sda / 255.0)
    return (im_rgb.T if is_matrix else im_rgb) - od


def _normalize_columns(m):
    n = np.linalg.norm(m, axis=0)
    return m / np.where(n == 0, 1.0, n)


def complement_stain_matrix(w):
    s0, s1 = w[:, 0], w[:, 1]
    s2 = np.cross(s0, s1)
    return np.array([s0, s1, s2 / np.linalg.norm(s2)]).T


def _image_to_matrix(im):
    return im.reshape(-1, im.shape[-1]).T if im.ndim == 3 else im


def color_deconvolution(im_rgb, w, I_0=None):
    wc = complement_stain_matrix(w) if np.linalg.norm(w[:, 2]) <= 1e-16 else w
    wc = _normalize_columns(wc)
    Q = np.linalg.inv(wc)
    m = _image_to_matrix(im_rgb)[:3]
    sda_fwd = rgb_to_sda(m, 255 if I_0 is not None else None, allow_negatives=True)
    sda_deconv = Q @ sda_fwd
    sda_inv = sda_to_rgb(sda_deconv, 255 if I_0 is not None else None)
    return sda_inv.T.reshape(im_rgb.shape), wc


def color_convolution(im_stains, w, I_0=None):
    m = _image_to_matrix(im_stains)
    sda_fwd = rgb_to_sda(m, 255 if I_0 is not None else None, allow_negatives=True)
    sda_conv = w @ sda_fwd
    sda_inv = sda_to_rgb(sda_conv, 255 if I_0 is not None else None)
    return sda_inv.T.reshape(im_stains.shape).clip(0, 255).astype(np.uint8)


def deconvolution_based_normalization(im_src, W_source, W_target):
    stains_float, _ = color_deconvolution(im_src, W_source, I_0=None)
    return color_convolution(stains_float, W_target)


def normalize_patch(patch_u8: np.ndarray, W_est, W_def) -> np.ndarray:
    """data.py:293-300 on one (ps, ps, 3) uint8 patch."""
    arr = patch_u8.astype(np.float32) + EPSILON
    return deconvolution_based_normalization(arr, W_est, W_def).astype(np.uint8)


def _argpercentile(arr, p):
    i = int(p * arr.size + 0.5)
    return np.argpartition(arr, i)[i]


def separate_stains_macenko_pca(im_sda, minimum_magnitude=16, min_angle_percentile=0.01, max_angle_percentile=0.99):
    m = _image_to_matrix(im_sda)
    m = m[:, np.isfinite(m).all(axis=0)]
    pcs = np.linalg.svd(m.astype(float), full_matrices=False)[0]
    proj = pcs.T[:-1] @ m
    filt = proj[:, np.linalg.norm(proj, axis=0) > minimum_magnitude]
    mm = _normalize_columns(filt)
    angles = (1 - mm[1]) * np.sign(mm[0])

    def vec(p):
        return pcs[:, :-1] @ filt[:, _argpercentile(angles, p)]

    w = np.array([vec(min_angle_percentile), vec(max_angle_percentile)]).T
    return complement_stain_matrix(_normalize_columns(w))


def rgb_separate_stains_macenko_pca(im_rgb, I_0):
    return separate_stains_macenko_pca(rgb_to_sda(im_rgb, I_0))
