"""CPU tests: the oracle's restatements against the live third-party calls and golden vectors."""

import cv2
import numpy as np
import pytest
from PIL import Image

from conftest import blob_mask, blob_thumbnail
from oracle import patchgrid, preprocess, segment


def test_rgb_to_saturation_exhaustive_vs_cv2():
    # all 2^24 RGB triples, in 16 slabs of 2^20
    g, b = np.meshgrid(np.arange(256, dtype=np.uint8), np.arange(256, dtype=np.uint8), indexing="ij")
    for r0 in range(0, 256, 16):
        slab = np.empty((16, 256, 256, 3), dtype=np.uint8)
        for k in range(16):
            slab[k, ..., 0] = r0 + k
            slab[k, ..., 1] = g
            slab[k, ..., 2] = b
        img = slab.reshape(16 * 256, 256, 3)
        want = cv2.cvtColor(img, cv2.COLOR_RGB2HSV)[..., 1]
        got = segment.rgb_to_saturation_restated(img)
        assert np.array_equal(want, got)


@pytest.mark.parametrize("k,t", [(7, 7), (3, 7), (5, 20), (9, 3)])
def test_median_threshold_identity_vs_cv2(k, t):
    rng = np.random.default_rng(k * 100 + t)
    sat = rng.integers(0, 30, size=(97, 131), dtype=np.uint8)
    want = cv2.medianBlur(sat, k) > t
    got = segment.median_threshold_restated(sat, k, t)
    assert np.array_equal(want, got)


def test_median_size_validation():
    with pytest.raises(ValueError, match="median_filter_size must be greater than 1 and odd"):
        segment.segment_tissue(np.zeros((8, 8, 3), np.uint8), median_filter_size=4)
    with pytest.raises(ValueError):
        segment.median_threshold_restated(np.zeros((8, 8), np.uint8), 1, 7)


@pytest.mark.parametrize("k", [1, 2, 3, 5, 6, 7, 8])
def test_closing_restated_vs_scipy(k):
    for seed in range(3):
        m = blob_mask(83, 117, seed, density=0.4, smooth=2.0) ^ (np.random.default_rng(seed).random((83, 117)) < 0.05)
        want = segment.binary_closing_lib(m, k)
        got = segment.closing_restated(m, k)
        assert np.array_equal(want, got)
        assert np.all(got >= m)                                         # extensive
        assert np.array_equal(segment.closing_restated(got, k), got)   # idempotent


def test_segment_tissue_stages_consistent():
    img = blob_thumbnail(160, 200, 3)
    st = segment.segment_stages(img, 7, 7, 6, 40, 30)
    assert np.array_equal(st["thr"], segment.median_threshold_restated(st["sat"], 7, 7))
    assert np.array_equal(st["closed"], segment.closing_restated(st["thr"], 6))
    assert np.array_equal(st["mask"], segment.segment_tissue(img, 7, 7, 6, 40, 30))
    # removers: strict '<' and 4-connectivity
    m = np.zeros((10, 10), bool)
    m[1:3, 1:3] = True          # area 4
    m[5, 5] = m[6, 6] = True    # two diagonal pixels = two 4-connected objects of area 1
    assert segment.remove_small_objects_lib(m, 4)[1, 1]
    assert not segment.remove_small_objects_lib(m, 5)[1, 1]
    assert not segment.remove_small_objects_lib(m, 2)[5, 5]


def test_golden_purple_coordinates(golden_purple):
    """Reference golden CSVs (tests/reference/*/purple.csv): uniform slide => full grid, in order."""
    for name, g in golden_purple.items():
        W, H, ps = g["slide_width"], g["slide_height"], g["patch_size"]
        thumb = np.empty((2048, 2048, 3), np.uint8)
        thumb[:] = g["rgb"]
        mask = segment.segment_tissue(thumb, 7, 7, 6, 100, 100)
        assert mask.all()
        coords = patchgrid.patch_coordinates(mask, W, H, ps)
        assert coords.tolist() == g["coords"], name
        # raster rule gives the same answer
        outer = segment.outermost_restated(mask)
        cent, step, dims = patchgrid.grid_centroids(W, H, ps, round(ps / 2))
        keep = patchgrid.raster_rule(outer, cent, W, H)
        assert (cent[keep] - round(ps / 2)).tolist() == g["coords"], name


def test_island_in_hole_is_dropped():
    """the reference's sequential union/difference drops islands inside holes."""
    m = np.zeros((64, 64), bool)
    m[4:60, 4:60] = True
    m[16:48, 16:48] = False     # hole
    m[28:36, 28:36] = True      # island in the hole
    coords = patchgrid.patch_coordinates(m, 640, 640, 20)
    cx = coords + 10
    island = (cx[:, 0] > 290) & (cx[:, 0] < 350) & (cx[:, 1] > 290) & (cx[:, 1] < 350)
    assert not island.any()
    outer = segment.outermost_restated(m)
    assert not outer[30, 30] and outer[10, 10]


@pytest.mark.parametrize("seed", range(6))
def test_raster_rule_matches_contour_oracle(seed):
    rng = np.random.default_rng(seed)
    ht, wt = int(rng.integers(40, 90)), int(rng.integers(40, 90))
    mask = blob_mask(ht, wt, seed, density=0.45, smooth=3.0)
    if seed % 2:
        mask = segment.segment_tissue(np.where(mask[..., None], np.uint8([200, 80, 180]), np.uint8(245)).astype(np.uint8),
                                      3, 7, 6, 12, 10)
    W = wt * int(rng.integers(3, 9)) + int(rng.integers(0, 5))
    H = ht * int(rng.integers(3, 9)) + int(rng.integers(0, 5))
    ps = int(rng.integers(3, 9))
    got_all = patchgrid.patch_coordinates(mask, W, H, ps, return_ties=True)
    if got_all is None:
        assert not mask.any()
        return
    want, ties = got_all
    outer = segment.outermost_restated(mask)
    half = round(ps / 2)
    cent, step, dims = patchgrid.grid_centroids(W, H, ps, half)
    keep = patchgrid.raster_rule(outer, cent, W, H)
    got = cent[keep] - half
    tie_set = {tuple(t) for t in ties.tolist()}
    a = {tuple(t) for t in want.tolist()} - tie_set
    b = {tuple(t) for t in got.tolist()} - tie_set
    assert a == b
    # order: x-major, y fastest
    assert got.tolist() == sorted(got.tolist())


def test_slide_params_config2():
    """BASELINE config 2: 100k x 80k @ 0.5 um/px, ResNet34 350 px @ 0.25 (SURVEY section 8 sizes)."""
    p = patchgrid.slide_params((100000, 80000), (2048, 1638), 0.5, 350, 0.25)
    assert p["patch_size"] == 175 and p["half_patch_size"] == 88
    assert p["min_object_size_px"] == 67 and p["min_hole_size_px"] == 61
    cent, step, dims = patchgrid.grid_centroids(100000, 80000, 175, 88)
    assert dims == (571, 457) and len(cent) == 260947


@pytest.mark.parametrize("ps,S", [(350, 224), (175, 224), (700, 224), (100, 299), (200, 299), (224, 224), (37, 64)])
def test_resize_restated_vs_pil(ps, S):
    rng = np.random.default_rng(ps + S)
    rgb = rng.integers(0, 256, size=(ps, ps, 3), dtype=np.uint8)
    want = preprocess.resize_lib_u8(rgb, S)
    got = preprocess.resize_restated_u8(rgb, S, S)
    assert np.array_equal(want, got)


def test_preprocess_lib_normalize_and_padding():
    slide = np.random.default_rng(0).integers(0, 256, size=(300, 400, 4), dtype=np.uint8)
    crop = preprocess.read_region_rgba(slide, 350, 250, 100, 100)
    assert crop[60:, :, :].max() == 0 and crop[:, 60:, :].max() == 0      # zero padded out of bounds
    mean, std = [0.7238, 0.5716, 0.6779], [0.112, 0.1459, 0.1089]
    t = preprocess.preprocess_lib(crop, 64, mean, std).numpy()
    u8 = preprocess.resize_restated_u8(crop[..., :3], 64, 64)
    want = preprocess.normalize_restated(u8, mean, std).transpose(2, 0, 1)
    assert np.array_equal(t, want)


def test_otsu_restatement_vs_cv2():
    """The float64 loop the CUDA kernel repeats (oracle.segment.otsu_threshold_restated) against cv.threshold(OTSU)."""
    import cv2
    from oracle import segment
    rng = np.random.default_rng(0)
    for t in range(60):
        h, w = rng.integers(5, 200, 2)
        kind = t % 4
        if kind == 0:
            img = rng.integers(0, 256, (h, w)).astype(np.uint8)
        elif kind == 1:
            img = np.clip(rng.normal(rng.integers(0, 255), rng.integers(1, 60), (h, w)), 0, 255).astype(np.uint8)
        elif kind == 2:
            img = np.where(rng.random((h, w)) < rng.random(), np.clip(rng.normal(30, 10, (h, w)), 0, 255),
                           np.clip(rng.normal(180, 30, (h, w)), 0, 255)).astype(np.uint8)
        else:
            img = np.full((h, w), rng.integers(0, 256), np.uint8)
        t_cv, _ = cv2.threshold(img, 0, 255, cv2.THRESH_BINARY | cv2.THRESH_OTSU)
        assert int(t_cv) == segment.otsu_threshold_restated(img)


def test_closing_conventions_restated_vs_scipy():
    """Every candidate footprint-origin convention: explicit window offsets == the scipy call it stands for; the
    front-padded candidate differs from the default only within k/2 px of the border."""
    from conftest import blob_mask
    from oracle import segment
    for conv in ("scipy_raw", "front_padded", "mirrored_erosion"):
        for k in (2, 3, 4, 6, 7, 8):
            for t in range(3):
                m = blob_mask(80, 90, t + k, density=0.5, smooth=2.0)
                m[0:5, 10:30] = True
                m[:, 0] |= (t % 2 == 1)
                assert np.array_equal(segment.binary_closing_lib(m, k, conv), segment.closing_restated(m, k, conv))
    m = blob_mask(80, 90, 3, density=0.5, smooth=2.0)
    d = np.argwhere(segment.binary_closing_lib(m, 6, "scipy_raw") != segment.binary_closing_lib(m, 6, "front_padded"))
    assert len(d) > 0 and all(y < 3 or x < 3 or y >= 77 or x >= 87 for y, x in d.tolist())
