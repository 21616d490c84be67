"""a13 / N3: fused tile-stitcher batch step vs the reference's torch operations (tilefuse.py:219-278)."""

import numpy as np
import pytest
import torch

from oracle import tilefuse as otf


def _inputs(B, K, h, seed, W, H, S, halo):
    g = torch.Generator().manual_seed(seed)
    np_l = torch.randn((B, 2, h, h), generator=g) * 3
    hv = torch.randn((B, 2, h, h), generator=g)
    tp_l = torch.randn((B, K, h, h), generator=g) * 4
    # non-overlapping interiors on an S grid, some hanging over every canvas edge, one fully outside
    cols = (W + S - 1) // S + 1
    idx = torch.randperm((cols + 1) * ((H + S - 1) // S + 2), generator=g)[:B]
    xs = (idx % (cols + 1) - 1) * S - halo + 3
    ys = (idx // (cols + 1) - 1) * S - halo + 5
    coords = torch.stack([xs, ys, torch.full_like(xs, S + 2 * halo), torch.full_like(xs, S + 2 * halo)], 1)
    coords[0, 0], coords[0, 1] = W + 50, 0                      # entirely outside: skipped
    coords[1, 0], coords[1, 1] = (W // S // 2) * S - halo + 3, -halo + 5     # at least one tile inside the canvas (same grid)
    for i in range(2, B):                                       # keep the interiors disjoint (the reference's never overlap)
        if coords[i, 0] == coords[1, 0] and coords[i, 1] == coords[1, 1]:
            coords[i, 0] = W + 50
    return np_l, hv, tp_l, coords


def test_oracle_shapes_and_sums():
    np_l, hv, tp_l, coords = _inputs(3, 4, 20, 0, 100, 90, 32, 4)
    a, b, c = otf.accumulate(np_l, hv, tp_l, coords, 32, 4, 0.5, 100, 90)
    assert a.shape == (90, 100) and b.shape == (90, 100, 2) and c.shape == (90, 100, 4)
    s = c.sum(-1)
    assert np.all((np.abs(s - 1) < 1e-5) | (s == 0))             # renormalised where a tile was pasted


@pytest.mark.gpu
@pytest.mark.parametrize("B,K,h,S,halo,W,H,alpha", [(9, 6, 164, 256, 46, 700, 600, 256 / 164), (5, 2, 40, 40, 0, 150, 170, 1.0),
                                                   (7, 5, 64, 37, 5, 120, 90, 0.37), (3, 16, 10, 64, 8, 100, 100, 2.0)])
@pytest.mark.parametrize("canvas_tile", [2048, 64])
def test_stitcher_matches_reference_ops(B, K, h, S, halo, W, H, alpha, canvas_tile):
    from wsinsight_b200.modellib.tilefuse import TileRemapStitcher

    np_l, hv, tp_l, coords = _inputs(B, K, h, B + K + h, W, H, S, halo)
    # canvas_tile = 64: the sparse canvases are many small tiles, every paste straddles several of them
    st = TileRemapStitcher(K, W, H, S, halo, slide_mpp=1.0, model_mpp=alpha, canvas_tile=canvas_tile)
    st.accumulate_batch_torch({"np": np_l.cuda(), "hv": hv.cuda(), "tp": tp_l.cuda()}, coords)
    a, b, c = otf.accumulate(np_l, hv, tp_l, coords, S, halo, alpha, W, H)
    for got, want, name in ((st.np_map, a, "np"), (st.hv_map, b, "hv"), (st.tp_map, c, "tp")):
        d = np.abs(got.cpu().numpy() - want).max()
        assert d <= 2e-6 * max(1.0, np.abs(want).max()), f"{name}: max abs diff {d}"
    assert (st.np_map != 0).any()
    # second batch through the reference's alternative keys accumulates into the same canvases
    np2, hv2, tp2, coords2 = _inputs(B, K, h, 99, W, H, S, halo)
    st.accumulate_batch_torch({"nuclei_binary_map": np2.cuda(), "hv_map": hv2.cuda(), "nuclei_type_map": tp2.cuda()}, coords2)
    a, b, c = otf.accumulate(np2, hv2, tp2, coords2, S, halo, alpha, W, H, a, b, c)
    assert np.abs(st.tp_map.cpu().numpy() - c).max() <= 2e-6
    assert st.allocated_tiles() <= st.tiles_x * st.tiles_y
    with pytest.raises(ValueError):
        st.accumulate_batch_torch({"np": np_l.cuda(), "hv": hv.cuda(), "tp": tp_l[:, :1].cuda()}, coords)


@pytest.mark.gpu
def test_overlapping_tiles_of_one_batch_paste_in_order():
    """The reference pastes tile after tile (tilefuse.py:262-278): where interiors of one batch overlap, the later tile
    wins.  The device path detects the overlap and pastes in order instead of racing."""
    from wsinsight_b200.modellib.tilefuse import TileRemapStitcher

    B, K, h, S, halo, W, H = 6, 3, 24, 48, 4, 200, 160
    g = torch.Generator().manual_seed(7)
    np_l, hv, tp_l = torch.randn((B, 2, h, h), generator=g), torch.randn((B, 2, h, h), generator=g), torch.randn((B, K, h, h), generator=g)
    xs = torch.tensor([0, 20, 40, 100, 110, 30])
    ys = torch.tensor([0, 10, 30, 60, 70, 20])
    coords = torch.stack([xs - halo, ys - halo, torch.full_like(xs, S + 2 * halo), torch.full_like(xs, S + 2 * halo)], 1)
    st = TileRemapStitcher(K, W, H, S, halo, slide_mpp=1.0, model_mpp=1.0, canvas_tile=64)
    st.accumulate_batch_torch({"np": np_l.cuda(), "hv": hv.cuda(), "tp": tp_l.cuda()}, coords)
    a, b, c = otf.accumulate(np_l, hv, tp_l, coords, S, halo, 1.0, W, H)
    assert np.abs(st.np_map.cpu().numpy() - a).max() <= 2e-6
    assert np.abs(st.hv_map.cpu().numpy() - b).max() <= 2e-5
    assert np.abs(st.tp_map.cpu().numpy() - c).max() <= 2e-6
