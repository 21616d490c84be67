"""N3 / a13: device finalize of the cell path (csrc/cellpost.cu) vs the oracle (oracle/cellpost.py), which executes the
reference's OpenCV / scipy calls and restates remove_small_objects and the watershed (skimage is not importable:
"parity unpinned" for those two; everything else is pinned against the live library)."""

import numpy as np
import pytest

from conftest import blob_mask, synth_nuclei
from oracle import cellpost


# ---------------------------------------------------------------- CPU: the oracle's own restatements
def test_contour_restatement_vs_cv2():
    """trace_external_contour (what the CUDA tracer implements) == cv2.findContours(RETR_EXTERNAL, CHAIN_APPROX_SIMPLE)."""
    import cv2
    from scipy import ndimage as ndi

    n = 0
    for seed in range(12):
        m = blob_mask(60, 70, seed, density=0.45, smooth=2.0 + seed % 3)
        lab, k = ndi.label(m, structure=np.ones((3, 3)))
        for i in range(1, k + 1):
            c = (lab == i).astype(np.uint8)
            cnts, _ = cv2.findContours(c, cv2.RETR_EXTERNAL, cv2.CHAIN_APPROX_SIMPLE)
            assert len(cnts) == 1
            assert np.array_equal(cnts[0].squeeze(1), cellpost.trace_external_contour(c))
            n += 1
    for c in (np.ones((1, 1), np.uint8), np.ones((1, 5), np.uint8), np.ones((4, 1), np.uint8), np.eye(5, dtype=np.uint8),
              np.array([[0, 1, 0], [1, 1, 1], [0, 1, 0]], np.uint8), np.ones((2, 2), np.uint8)):
        cnts, _ = cv2.findContours(c.copy(), cv2.RETR_EXTERNAL, cv2.CHAIN_APPROX_SIMPLE)
        assert np.array_equal(cnts[0].reshape(-1, 2), cellpost.trace_external_contour(c))
    assert n > 50


def _two_phase_watershed(image, markers, mask):
    """The formulation the CUDA kernels use: flood levels, then (hops, label) from the minimal-level neighbours."""
    L = cellpost.flood_levels(image, markers, mask)
    h, w = image.shape
    BIG = np.int64(1) << 60
    m = (markers > 0) & mask
    key = np.where(m, markers.astype(np.int64), BIG)
    Lp = np.pad(L, 1, constant_values=np.inf)
    mL = np.minimum.reduce([Lp[:-2, 1:-1], Lp[1:-1, :-2], Lp[1:-1, 2:], Lp[2:, 1:-1]])
    while True:
        kp = np.pad(key, 1, constant_values=BIG)
        best = key.copy()
        for sl in ((slice(0, -2), slice(1, -1)), (slice(1, -1), slice(0, -2)), (slice(1, -1), slice(2, None)),
                   (slice(2, None), slice(1, -1))):
            qL, qk = Lp[sl], kp[sl]
            cand = np.where((qL == mL) & np.isfinite(qL) & (qk < BIG), qk + (np.int64(1) << 32), BIG)
            best = np.where(mask & ~m, np.minimum(best, cand), best)
        if (best == key).all():
            break
        key = best
    return np.where(key < BIG, key & 0xffffffff, 0).astype(np.int32)


@pytest.mark.parametrize("seed", range(3))
def test_parallel_watershed_formulation_equals_heap_flooding(seed):
    npm, hv, tp = synth_nuclei(120, 150, seed)
    st = cellpost.proc_np_hv_stages(npm, hv, 10)
    assert not st["empty"] and st["marker"].max() > 5
    got = _two_phase_watershed(st["dist"], st["marker"], st["blb"].astype(bool))
    assert np.array_equal(got, st["inst"])


def test_oracle_finalize_jobs_and_lists():
    npm, hv, tp = synth_nuclei(150, 170, 5)
    ins, prb, ply = cellpost.finalize(npm, hv, tp, 10, tile_size=64, padding_size=16)
    assert len(ins) == len(prb) >= len(ply) > 10
    assert all(b.shape == (1, 4) for b in ins) and all(p.shape == (1, tp.shape[2]) for p in prb)
    assert len(cellpost.finalize_jobs(150, 170, 64, 16)) == 9


# ---------------------------------------------------------------- GPU
def _canon(boxes, probs):
    o = np.lexsort((boxes[:, 3], boxes[:, 2], boxes[:, 1], boxes[:, 0]))
    return boxes[o], probs[o], o


@pytest.mark.gpu
@pytest.mark.parametrize("h,w,seed,K,min_size", [(128, 160, 0, 5, 10), (200, 230, 1, 3, 20), (97, 61, 2, 6, 5),
                                                  (300, 280, 3, 4, 10)])
def test_single_tile_stages_vs_oracle(h, w, seed, K, min_size):
    """One padded tile == the whole image: instance map, boxes, class means and polygons against the oracle."""
    import torch

    from wsinsight_b200.modellib.tilefuse import TileRemapStitcher

    npm, hv, tp = synth_nuclei(h, w, seed, K=K)
    st = TileRemapStitcher(K, w, h, 32, 0, 1.0, 1.0, min_object_size=min_size, canvas_tile=4096)
    t = st._tile(0, 0)
    t[0].copy_(torch.from_numpy(npm)); t[1].copy_(torch.from_numpy(hv)); t[2].copy_(torch.from_numpy(tp))
    boxes, probs, pxy, poff, maps = st.finalize_arrays(tile_size=4096, padding_size=64, return_inst_maps=True)
    want = cellpost.proc_np_hv_stages(npm, hv, min_size)
    inst = want["inst"]
    # instance ids are dense in raster order of the marker's first pixel on both sides up to removed labels: compare as
    # partitions
    got = maps[0]
    assert got.shape == inst.shape
    assert np.array_equal(got > 0, inst > 0), f"foreground differs on {int(((got > 0) != (inst > 0)).sum())} px"
    pairs = np.unique(np.stack([got[inst > 0], inst[inst > 0]], 1), axis=0)
    n_diff_px = 0
    if len(pairs) != len(np.unique(pairs[:, 0])) or len(pairs) != len(np.unique(pairs[:, 1])):
        # not a bijection: count the pixels outside the majority mapping (exact float ties of the flooding order)
        from collections import Counter
        maj = {}
        for a, b in zip(got[inst > 0].tolist(), inst[inst > 0].tolist()):
            maj.setdefault(b, Counter())[a] += 1
        n_diff_px = sum(sum(c.values()) - max(c.values()) for c in maj.values())
    assert n_diff_px <= 2e-4 * (inst > 0).sum(), f"{n_diff_px} px labelled differently"
    ins, prb, ply = cellpost.stitching_worker(npm, hv, tp, 0, 0, (slice(0, h), slice(0, w)), min_size, inst=inst)
    wb, wp = np.concatenate(ins), np.concatenate(prb)
    if n_diff_px == 0:
        assert len(boxes) == len(wb)
        assert np.array_equal(boxes, wb)                        # same order: label order == ndi.label order
        assert np.abs(probs - wp).max() <= 1e-6
        polys = [pxy[poff[i]:poff[i + 1]] for i in range(len(boxes)) if poff[i + 1] > poff[i]]
        assert len(polys) == len(ply)
        for a, b in zip(polys, ply):
            assert np.array_equal(a, b)


@pytest.mark.gpu
@pytest.mark.parametrize("tile,pad,canvas_tile", [(64, 16, 50), (100, 32, 2048), (2048, 64, 128)])
def test_finalize_tiled_vs_oracle(tile, pad, canvas_tile):
    """finalize() over several jobs and sparse canvas tiles: the same instance records as the oracle's finalize."""
    import torch

    from wsinsight_b200.modellib.tilefuse import TileRemapStitcher

    h, w, K = 210, 260, 5
    npm, hv, tp = synth_nuclei(h, w, 11, K=K)
    npm[:, 130:200] = 0.0                      # an empty band: some canvas tiles / jobs hold nothing
    st = TileRemapStitcher(K, w, h, 32, 0, 1.0, 1.0, min_object_size=10, canvas_tile=canvas_tile)
    T = canvas_tile
    for ty in range(st.tiles_y):
        for tx in range(st.tiles_x):
            sl = (slice(ty * T, min(h, ty * T + T)), slice(tx * T, min(w, tx * T + T)))
            if npm[sl].max() < 0.5 and canvas_tile < 2048:
                continue                       # never touched: stays unallocated
            t = st._tile(ty, tx)
            t[0].copy_(torch.from_numpy(npm[sl])); t[1].copy_(torch.from_numpy(hv[sl])); t[2].copy_(torch.from_numpy(tp[sl]))
    dense = [x.cpu().numpy() for x in st.window(0, h, 0, w)]
    ins, prb, ply = st.finalize(tile_size=tile, padding_size=pad)
    wi, wp, wl = cellpost.finalize(dense[0], dense[1], dense[2], 10, tile_size=tile, padding_size=pad)
    assert len(ins) == len(prb) and all(b.shape == (1, 4) and b.dtype == np.int32 for b in ins)
    gb, gp, _ = _canon(np.concatenate(ins), np.concatenate(prb))
    wb, wpp, _ = _canon(np.concatenate(wi), np.concatenate(wp))
    same = len(gb) == len(wb) and np.array_equal(gb, wb)
    if not same:
        # exact float ties may move a few boundary pixels: at least 99 % of the boxes must be identical
        sa, sb = {tuple(r) for r in gb.tolist()}, {tuple(r) for r in wb.tolist()}
        assert len(sa & sb) >= 0.99 * max(len(sa), len(sb)), (len(sa), len(sb), len(sa & sb))
    else:
        assert np.abs(gp - wpp).max() <= 1e-6
        assert sorted(map(lambda a: a.tobytes(), ply)) == sorted(map(lambda a: a.tobytes(), wl))
    # release=True frees canvas tiles behind the sweep and gives the same records
    b2 = st.finalize_arrays(tile_size=tile, padding_size=pad, release=True)[0]
    assert np.array_equal(_canon(b2, np.zeros((len(b2), 1)))[0], gb) and st.allocated_tiles() == 0


@pytest.mark.gpu
def test_finalize_empty_and_single_blob():
    import torch

    from wsinsight_b200.modellib.tilefuse import TileRemapStitcher

    st = TileRemapStitcher(3, 80, 70, 32, 0, 1.0, 1.0)
    assert st.finalize() == ([], [], [])                       # nothing accumulated
    t = st._tile(0, 0)
    assert st.finalize() == ([], [], [])                       # canvases all zero
    npm = np.zeros((70, 80), np.float32)
    npm[20:40, 30:55] = 0.9                                    # one blob, no HV structure: falls back to label(blb)
    hv = np.zeros((70, 80, 2), np.float32)
    tp = np.full((70, 80, 3), 1 / 3, np.float32)
    t[0].copy_(torch.from_numpy(npm)); t[1].copy_(torch.from_numpy(hv)); t[2].copy_(torch.from_numpy(tp))
    ins, prb, ply = st.finalize()
    wi, wp, wl = cellpost.finalize(npm, hv, tp, 20)
    assert len(ins) == len(wi) and all(np.array_equal(a, b) for a, b in zip(ins, wi))
    assert all(np.array_equal(a, b) for a, b in zip(ply, wl)) and np.abs(np.concatenate(prb) - np.concatenate(wp)).max() < 1e-6


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["tiny", "all_foreground", "k16_min0", "single_row"])
def test_finalize_edge_cases(case):
    """Degenerate tiles: smaller than the Sobel kernel (reflect-101 folds more than once), one blob covering everything,
    the maximum class count with min_object_size = 0, a one-pixel-high canvas."""
    import torch

    from wsinsight_b200.modellib.tilefuse import TileRemapStitcher

    rng = np.random.default_rng(7)
    if case == "tiny":
        h, w, K, ms = 5, 7, 3, 2
        npm = (rng.random((h, w)) > 0.4).astype(np.float32) * 0.9
        hv = rng.normal(0, 0.3, (h, w, 2)).astype(np.float32)
    elif case == "all_foreground":
        h, w, K, ms = 40, 50, 2, 10
        npm = np.full((h, w), 0.8, np.float32)
        yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
        hv = np.stack([(xx % 13) / 6.5 - 1, (yy % 11) / 5.5 - 1], -1).astype(np.float32)
    elif case == "k16_min0":
        h, w, K, ms = 90, 110, 16, 0
        npm, hv, _ = synth_nuclei(h, w, 21, K=4)
    else:
        h, w, K, ms = 1, 64, 3, 3
        npm = (rng.random((h, w)) > 0.5).astype(np.float32)
        hv = rng.normal(0, 0.3, (h, w, 2)).astype(np.float32)
    tp = rng.random((h, w, K)).astype(np.float32)
    tp /= tp.sum(-1, keepdims=True)
    st = TileRemapStitcher(K, w, h, 16, 0, 1.0, 1.0, min_object_size=ms)
    t = st._tile(0, 0)
    t[0].copy_(torch.from_numpy(npm)); t[1].copy_(torch.from_numpy(hv)); t[2].copy_(torch.from_numpy(tp))
    boxes, probs, pxy, poff, maps = st.finalize_arrays(return_inst_maps=True)
    want = cellpost.proc_np_hv_stages(npm, hv, ms)["inst"]
    got = maps[0]
    assert np.array_equal(got > 0, want > 0)
    same_partition = len(np.unique(np.stack([got[want > 0], want[want > 0]], 1), axis=0)) == len(np.unique(want[want > 0]))
    wi, wp, wl = cellpost.stitching_worker(npm, hv, tp, 0, 0, (slice(0, h), slice(0, w)), ms, inst=want)
    if same_partition and len(wi):
        assert np.array_equal(boxes, np.concatenate(wi)) and np.abs(probs - np.concatenate(wp)).max() <= 1e-6
        polys = [pxy[poff[i]:poff[i + 1]] for i in range(len(boxes)) if poff[i + 1] > poff[i]]
        assert len(polys) == len(wl) and all(np.array_equal(a, b) for a, b in zip(polys, wl))
    else:
        assert len(boxes) == len(wi) or not same_partition
