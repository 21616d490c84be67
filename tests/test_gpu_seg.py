"""GPU parity: tissue segmentation + patch grid through the C ABI vs the CPU oracle (bit-exact)."""

import numpy as np
import pytest
import torch

from conftest import blob_mask, blob_thumbnail
from oracle import patchgrid, segment

pytestmark = pytest.mark.gpu


def _seg(img, **kw):
    from wsinsight_b200.patchlib import segment_tissue
    return segment_tissue(img, **kw)


def _report(name, got, want):
    bad = np.argwhere(got != want)
    return f"{name}: {len(bad)} mismatching pixels of {got.size}; first {bad[:5].tolist()}"


@pytest.mark.parametrize("shape,seed", [((64, 64), 0), ((97, 131), 1), ((256, 300), 2), ((511, 257), 3), ((1024, 819), 4)])
def test_stage_median_threshold(shape, seed):
    img = blob_thumbnail(*shape, seed)
    st = segment.segment_stages(img, 7, 7, 1, 0, 0)
    got = _seg(img, median_filter_size=7, binary_threshold=7, closing_kernel_size=1, min_object_size_px=0,
               min_hole_size_px=0)
    assert np.array_equal(got, st["thr"]), _report("median/threshold", got, st["thr"])


@pytest.mark.parametrize("k", [3, 5, 9])
def test_stage_median_other_sizes(k):
    img = blob_thumbnail(120, 150, 10 + k)
    want = segment.segment_stages(img, k, 12, 1, 0, 0)["thr"]
    got = _seg(img, median_filter_size=k, binary_threshold=12, closing_kernel_size=1, min_object_size_px=0,
               min_hole_size_px=0)
    assert np.array_equal(got, want), _report(f"median {k}", got, want)


@pytest.mark.parametrize("k", [2, 3, 6, 7, 8])
def test_stage_closing(k):
    img = blob_thumbnail(200, 260, 20 + k, smooth=3)
    want = segment.segment_stages(img, 7, 7, k, 0, 0)["closed"]
    got = _seg(img, closing_kernel_size=k, min_object_size_px=0, min_hole_size_px=0)
    assert np.array_equal(got, want), _report(f"closing {k}", got, want)


@pytest.mark.parametrize("shape,seed,obj,hole", [((200, 260), 30, 50, 0), ((200, 260), 31, 0, 60), ((300, 333), 32, 120, 90),
                                                 ((1638, 2048), 33, 67, 61), ((2048, 2048), 34, 512, 1024)])
def test_full_mask(shape, seed, obj, hole):
    img = blob_thumbnail(*shape, seed, smooth=5 if shape[0] < 1000 else 12)
    want = segment.segment_tissue(img, 7, 7, 6, obj, hole)
    got = _seg(img, min_object_size_px=obj, min_hole_size_px=hole)
    assert got.dtype == bool and np.array_equal(got, want), _report("mask", got, want)


def _border_touching_thumbnail(h, w, seed):
    """tissue touching all four thumbnail borders (where the closing conventions differ) + interior blobs"""
    img = blob_thumbnail(h, w, seed, smooth=3)
    rng = np.random.default_rng(seed)
    fg = lambda n, m: np.stack([rng.integers(150, 230, (n, m)), rng.integers(60, 150, (n, m)),  # noqa: E731
                                rng.integers(140, 210, (n, m))], -1).astype(np.uint8)
    img[:9, w // 5: w // 2] = fg(9, w // 2 - w // 5)                 # top
    img[h - 8:, w // 3: 3 * w // 4] = fg(8, 3 * w // 4 - w // 3)     # bottom
    img[h // 4: h // 2, :10] = fg(h // 2 - h // 4, 10)               # left
    img[h // 3: 2 * h // 3, w - 7:] = fg(2 * h // 3 - h // 3, 7)     # right
    img[:6, :6] = fg(6, 6)                                           # corner
    return img


@pytest.mark.parametrize("convention", ["scipy_raw", "front_padded", "mirrored_erosion"])
@pytest.mark.parametrize("k", [6, 4, 7])
def test_closing_conventions_vs_scipy_forms(convention, k):
    """Each candidate convention of the even-footprint closing against the scipy call it stands for, on a thumbnail
    with tissue on all four borders; the candidates differ from the default only within k/2 px of the border."""
    img = _border_touching_thumbnail(150, 190, 70 + k)
    thr = segment.segment_stages(img, 7, 7, 1, 0, 0)["thr"]
    want = segment.binary_closing_lib(thr, k, convention)
    assert np.array_equal(want, segment.closing_restated(thr, k, convention))
    got = _seg(img, closing_kernel_size=k, min_object_size_px=0, min_hole_size_px=0, closing_convention=convention)
    assert np.array_equal(got, want), _report(f"closing {k} {convention}", got, want)
    base = segment.binary_closing_lib(thr, k, "scipy_raw")
    if convention == "front_padded":
        d = np.argwhere(want != base)
        b = k // 2
        assert all(y < b or x < b or y >= 150 - b or x >= 190 - b for y, x in d.tolist())
        if k % 2 == 0:
            assert len(d) > 0          # the test image does exercise the difference
    # full pipeline under the convention
    want_full = segment.segment_tissue(img, 7, 7, k, 40, 30, closing_convention=convention)
    got_full = _seg(img, closing_kernel_size=k, min_object_size_px=40, min_hole_size_px=30,
                    closing_convention=convention)
    assert np.array_equal(got_full, want_full), _report("mask", got_full, want_full)


@pytest.mark.parametrize("shape,seed", [((120, 150), 0), ((300, 333), 1), ((1638, 2048), 2), ((64, 64), 3)])
def test_otsu_mode_vs_cv2(shape, seed):
    """Extra mode (not upstream): threshold chosen by cv.threshold(..., THRESH_OTSU) on the median-filtered S."""
    from wsinsight_b200.patchlib import segment_tissue_device
    img = blob_thumbnail(*shape, 80 + seed, smooth=5 if shape[0] < 1000 else 12)
    want, t_want = segment.segment_tissue(img, 7, "otsu", 6, 30, 20, return_threshold=True)
    masks, _, info = segment_tissue_device([torch.from_numpy(img).cuda()], 7, "otsu", 6, 30, 20, return_info=True)
    assert int(info[0, 0].item()) == t_want, (int(info[0, 0].item()), t_want)
    got = masks[0].cpu().numpy().astype(bool)
    assert np.array_equal(got, want), _report("otsu mask", got, want)
    # the exact median itself (other window sizes too)
    for k in (3, 5):
        w2, t2 = segment.segment_tissue(img, k, "otsu", 1, 0, 0, return_threshold=True)
        m2, _, i2 = segment_tissue_device([torch.from_numpy(img).cuda()], k, "otsu", 1, 0, 0, return_info=True)
        assert int(i2[0, 0].item()) == t2 and np.array_equal(m2[0].cpu().numpy().astype(bool), w2)


def test_otsu_cohort_and_flags():
    from wsinsight_b200.patchlib import segment_tissue_device
    imgs = [blob_thumbnail(200, 260, 90), np.full((40, 50, 3), 245, np.uint8), blob_thumbnail(97, 131, 91, smooth=4)]
    masks, outers, info = segment_tissue_device([torch.from_numpy(i).cuda() for i in imgs], 7, "otsu", 6, 20, 20,
                                                return_info=True)
    info = info.cpu().numpy()
    for i, img in enumerate(imgs):
        want, t = segment.segment_tissue(img, 7, "otsu", 6, 20, 20, return_threshold=True)
        assert info[i, 0] == t and np.array_equal(masks[i].cpu().numpy().astype(bool), want)
        assert bool(info[i, 1]) == bool(want.any())


def test_legacy_and_run_based_labelling_agree(monkeypatch):
    """A/B switch WSI_B200_SEG_LEGACY=1 (per-pixel unions, k*k-tap morphology) vs the default kernels."""
    from wsinsight_b200.patchlib import segment_tissue_device
    imgs = [blob_thumbnail(301, 277, 95, smooth=3), _border_touching_thumbnail(130, 170, 96)]
    new = segment_tissue_device([torch.from_numpy(i).cuda() for i in imgs], 7, 7, 6, [25, 30], [25, 20])
    monkeypatch.setenv("WSI_B200_SEG_LEGACY", "1")
    old = segment_tissue_device([torch.from_numpy(i).cuda() for i in imgs], 7, 7, 6, [25, 30], [25, 20])
    for a, b in zip(new[0] + new[1], old[0] + old[1]):
        assert torch.equal(a, b)


def test_edge_cases_empty_full_tiny():
    for img in (np.full((40, 50, 3), 245, np.uint8), np.tile(np.uint8([160, 32, 240]), (40, 50, 1)),
                np.tile(np.uint8([160, 32, 240]), (1, 1, 1)), blob_thumbnail(7, 300, 5), blob_thumbnail(300, 5, 6)):
        want = segment.segment_tissue(img, 7, 7, 6, 10, 10)
        got = _seg(img, min_object_size_px=10, min_hole_size_px=10)
        assert np.array_equal(got, want), _report(str(img.shape), got, want)


def test_median_size_error_message():
    with pytest.raises(ValueError, match="median_filter_size must be greater than 1 and odd, but got 4"):
        _seg(np.zeros((8, 8, 3), np.uint8), median_filter_size=4)
    from wsinsight_b200 import _cabi
    lib = _cabi.load()
    t = torch.zeros(64, dtype=torch.uint8, device="cuda")
    rc = lib.wsi_seg_tissue_u8(t.data_ptr(), 4, 4, 12, 4, 7, 6, 0, 0, t.data_ptr(), None, t.data_ptr(), 64, None)
    assert rc == -2 and "median_filter_size must be greater than 1 and odd, but got 4" in _cabi.last_error()


def test_cohort_batch_matches_single_slides():
    """BASELINE config 4 in miniature: ragged cohort, per-slide thresholds, one launch sequence."""
    from wsinsight_b200.patchlib import segment_tissue_device
    shapes = [(200, 260), (130, 97), (256, 256), (64, 300), (301, 77)]
    imgs = [blob_thumbnail(h, w, 40 + i, smooth=4) for i, (h, w) in enumerate(shapes)]
    obj = [30, 0, 80, 10, 55]
    hole = [20, 40, 0, 10, 35]
    masks, outers = segment_tissue_device([torch.from_numpy(i).cuda() for i in imgs], 7, 7, 6, obj, hole)
    for img, o, h, m, ou in zip(imgs, obj, hole, masks, outers):
        want = segment.segment_tissue(img, 7, 7, 6, o, h)
        assert np.array_equal(m.cpu().numpy().astype(bool), want), _report("batch mask", m.cpu().numpy().astype(bool), want)
        wo = segment.outermost_restated(want)
        assert np.array_equal(ou.cpu().numpy().astype(bool), wo), _report("batch outer", ou.cpu().numpy().astype(bool), wo)


@pytest.mark.parametrize("seed", range(4))
def test_outermost_components(seed):
    from wsinsight_b200.patchlib.patch import outermost_components
    m = blob_mask(150 + 31 * seed, 190, 50 + seed, density=0.5, smooth=2.5)
    m[60:100, 60:100] = True
    m[70:90, 70:90] = False
    m[78:82, 78:82] = True      # island in a hole
    got = outermost_components(torch.from_numpy(m.astype(np.uint8)).cuda()).cpu().numpy().astype(bool)
    want = segment.outermost_restated(m)
    assert np.array_equal(got, want), _report("outer", got, want)


def _grid(mask, W, H, ps, overlap=0.0):
    from wsinsight_b200.patchlib import get_multipolygon_from_binary_arr, get_patch_coordinates_within_polygon
    res = get_multipolygon_from_binary_arr(mask.astype(np.uint8) * 255, slide_dims=(W, H))
    if res is None:
        return None
    return get_patch_coordinates_within_polygon(W, H, ps, round(ps / 2), res[0], overlap)


def test_grid_golden_purple(golden_purple):
    for name, g in golden_purple.items():
        img = np.empty((2048, 2048, 3), np.uint8)
        img[:] = g["rgb"]
        mask = _seg(img, min_object_size_px=100, min_hole_size_px=100)
        assert mask.all()
        got = _grid(mask, g["slide_width"], g["slide_height"], g["patch_size"])
        assert got.tolist() == g["coords"], name


@pytest.mark.parametrize("seed", range(6))
def test_grid_vs_contour_oracle(seed):
    rng = np.random.default_rng(100 + seed)
    ht, wt = int(rng.integers(60, 200)), int(rng.integers(60, 200))
    mask = segment.segment_tissue(blob_thumbnail(ht, wt, 60 + seed, smooth=4), 7, 7, 6, 20, 15)
    W = wt * int(rng.integers(8, 40)) + int(rng.integers(0, 7))
    H = ht * int(rng.integers(8, 40)) + int(rng.integers(0, 7))
    ps = int(rng.integers(9, 40))
    overlap = [0.0, 0.0, 0.5, 0.25, -0.5, 0.0][seed]
    got = _grid(mask, W, H, ps, overlap)
    ref = patchgrid.patch_coordinates(mask, W, H, ps, round(ps / 2), overlap, return_ties=True)
    if ref is None:
        assert got is None
        return
    want, ties = ref
    tie = {tuple(t) for t in ties.tolist()}
    a = [t for t in map(tuple, want.tolist()) if t not in tie]
    b = [t for t in map(tuple, got.tolist()) if t not in tie]
    assert a == b, f"{len(a)} vs {len(b)}; ties {len(tie)}"
    assert got.dtype == np.int64 and got.tolist() == sorted(got.tolist())


def test_grid_self_touching_necks_vs_contour_oracle():
    """Contours that touch themselves (1-px necks, diagonal contacts, a hole reaching a neck): the shapes for which the
    reference repairs polygons with Shapely's buffer(0) (patch.py:95-97).  Oracle: live cv2 contours + even-odd
    point-in-polygon in the reference's add / subtract order; GEOS' own repair is NOT pinned (shapely is absent)."""
    m = np.zeros((60, 80), bool)
    m[5:25, 5:25] = True
    m[14, 25:40] = True                      # 1-px horizontal neck
    m[5:25, 40:60] = True
    m[25:40, 50] = True                      # 1-px vertical neck
    m[40:55, 40:70] = True
    m[45:50, 45:50] = False                  # hole
    m[30:38, 10:18] = True
    m[38:46, 18:26] = True                   # diagonal contact of two squares (8-connected, one contour)
    m[12:16, 12:16] = False
    m[16:20, 16:20] = False                  # two holes touching diagonally
    W, H, ps = 80 * 23 + 3, 60 * 19 + 1, 17
    got = _grid(m, W, H, ps)
    want, ties = patchgrid.patch_coordinates(m, W, H, ps, round(ps / 2), 0.0, return_ties=True)
    tie = {tuple(t) for t in ties.tolist()}
    a = [t for t in map(tuple, want.tolist()) if t not in tie]
    b = [t for t in map(tuple, got.tolist()) if t not in tie]
    assert a == b and len(a) > 100, f"{len(a)} vs {len(b)}; ties {len(tie)}"


def test_grid_errors_and_empty():
    from wsinsight_b200.patchlib import get_multipolygon_from_binary_arr, get_patch_coordinates_within_polygon
    assert get_multipolygon_from_binary_arr(np.zeros((32, 32), np.uint8), slide_dims=(320, 320)) is None
    reg = get_multipolygon_from_binary_arr(np.ones((32, 32), np.uint8), slide_dims=(320, 320))[0]
    with pytest.raises(ValueError, match="overlap must be in"):
        get_patch_coordinates_within_polygon(320, 320, 16, 8, reg, overlap=1.0)
    c = get_patch_coordinates_within_polygon(320, 320, 1000, 500, reg)
    assert c.shape == (0, 2)


def test_full_size_config2_properties():
    """BASELINE config 2 sizes (100k x 80k @0.5, patch 175): properties that need no CPU oracle run."""
    from wsinsight_b200.patchlib.patch import TissueRegion, patch_grid_device
    img = blob_thumbnail(1638, 2048, 77, smooth=25)
    mask = _seg(img, min_object_size_px=67, min_hole_size_px=61)
    want = segment.segment_tissue(img, 7, 7, 6, 67, 61)
    assert np.array_equal(mask, want)
    outer = torch.from_numpy(segment.outermost_restated(want).astype(np.uint8)).cuda()
    coords, td = patch_grid_device(TissueRegion(outer, 100000, 80000), 175, 88)
    c = coords.cpu().numpy().astype(np.int64)
    assert td == (571, 457)
    assert len(c) > 1000 and (c % 175 == 0).all()
    lin = c[:, 0] * 100000 + c[:, 1]
    assert (np.diff(lin) > 0).all()                       # strictly x-major / y-fastest sorted
    # every kept centroid sits in a 2x2 thumbnail cell with at least 3 tissue corners
    u = ((c[:, 0] + 88) * 2048) // 100000
    v = ((c[:, 1] + 88) * 1638) // 80000
    o = np.pad(outer.cpu().numpy(), ((0, 1), (0, 1)))
    assert ((o[v, u].astype(int) + o[v, u + 1] + o[v + 1, u] + o[v + 1, u + 1]) >= 3).all()
    # and agrees with the CPU raster rule on the full 260 947-candidate grid
    cent, step, dims = patchgrid.grid_centroids(100000, 80000, 175, 88)
    keep = patchgrid.raster_rule(outer.cpu().numpy().astype(bool), cent, 100000, 80000)
    assert np.array_equal(cent[keep] - 88, c)
