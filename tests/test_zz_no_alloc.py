"""`wsi_patch_grid_static`: the allocation-free variant of `wsi_patch_grid` (scratch in a file-scope device array of the
library; no counterpart in the reference).

Collected LAST on purpose (file name): a failure here must not hide other tests.
Integer work: the bar is bit-exact equality with the existing entry point, on two calls in a row that share the array.
"""

import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _outer_mask(seed, ht, wt):
    """A 0/1 'outermost components' raster with blobs, holes filled or not -- wsi_patch_grid only reads it."""
    rng = np.random.default_rng(seed)
    m = np.zeros((ht, wt), np.uint8)
    yy, xx = np.mgrid[0:ht, 0:wt]
    for _ in range(12):
        cy, cx = rng.integers(0, ht), rng.integers(0, wt)
        ry, rx = rng.integers(ht // 12 + 2, ht // 4 + 3), rng.integers(wt // 12 + 2, wt // 4 + 3)
        m |= (((yy - cy) / ry) ** 2 + ((xx - cx) / rx) ** 2 <= 1).astype(np.uint8)
    m[rng.random((ht, wt)) < 0.02] = 0          # pinholes: ragged rows of kept candidates
    return m


def _run(torch, _cabi, static, outer, slide_w, slide_h, half, step):
    ni, nj = len(range(half, slide_w, step)), len(range(half, slide_h, step))
    cap = max(1, ni * nj)
    coords = torch.full((cap, 2), -7, dtype=torch.int32, device="cuda")
    count = torch.full((1,), -1, dtype=torch.int32, device="cuda")
    td = (C.c_int * 2)()
    head = (outer.data_ptr(), int(outer.shape[0]), int(outer.shape[1]), slide_w, slide_h, half, step,
            coords.data_ptr(), cap, count.data_ptr(), td)
    if static:
        _cabi.call("wsi_patch_grid_static", *head, _cabi.stream_ptr())
    else:
        ws_bytes = max(int(_cabi.load().wsi_patch_grid_workspace_bytes(slide_w, slide_h, half, step)), 256)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
        _cabi.call("wsi_patch_grid", *head, ws.data_ptr(), ws_bytes, _cabi.stream_ptr())
    torch.cuda.synchronize()
    n = int(count.item())
    return n, coords[:min(n, cap)].cpu().numpy(), (int(td[0]), int(td[1]))


def test_patch_grid_static_matches_patch_grid_on_two_calls_in_a_row():
    from wsinsight_b200 import _cabi

    torch = _cabi.require_cuda()
    # (seed, thumbnail Ht x Wt, slide W x H, half, step): a large grid first, then a small one that leaves most of the
    # shared array holding the first call's offsets, then the large one again
    cases = [
        (1, 819, 1024, 100_000, 80_000, 87, 175),      # the C2 geometry: 571 x 457 candidates, 1020 blocks
        (2, 200, 260, 9_000, 7_000, 64, 96),           # 94 x 73 candidates, 27 blocks
        (3, 819, 1024, 100_000, 80_000, 50, 131),      # overlapping patches: 763 x 611 candidates
    ]
    for seed, ht, wt, sw, sh, half, step in cases:
        outer = torch.from_numpy(_outer_mask(seed, ht, wt)).cuda()
        n_ref, c_ref, td_ref = _run(torch, _cabi, False, outer, sw, sh, half, step)
        n_a, c_a, td_a = _run(torch, _cabi, True, outer, sw, sh, half, step)
        n_b, c_b, td_b = _run(torch, _cabi, True, outer, sw, sh, half, step)     # second call in a row, same inputs
        assert n_ref > 0, "the case would be vacuous"
        assert (n_a, td_a) == (n_ref, td_ref) and (n_b, td_b) == (n_ref, td_ref)
        assert np.array_equal(c_a, c_ref) and np.array_equal(c_b, c_ref)


def test_patch_grid_static_rejects_grids_beyond_its_scratch():
    from wsinsight_b200 import _cabi

    torch = _cabi.require_cuda()
    outer = torch.ones((64, 64), dtype=torch.uint8, device="cuda")
    coords = torch.empty((16, 2), dtype=torch.int32, device="cuda")
    count = torch.zeros(1, dtype=torch.int32, device="cuda")
    # 16384 x 16384 = 2^28 candidates > WSI_PATCH_GRID_STATIC_MAX_CANDIDATES (2^24)
    with pytest.raises(Exception, match="exceed the static scratch"):
        _cabi.call("wsi_patch_grid_static", outer.data_ptr(), 64, 64, 1 << 20, 1 << 20, 0, 64, coords.data_ptr(), 16,
                   count.data_ptr(), None, _cabi.stream_ptr())
    # exactly at the limit is accepted: 4096 x 4096 candidates (capacity 16: only the count is complete)
    _cabi.call("wsi_patch_grid_static", outer.data_ptr(), 64, 64, 4096 * 64, 4096 * 64, 0, 64, coords.data_ptr(), 16,
               count.data_ptr(), None, _cabi.stream_ptr())
    torch.cuda.synchronize()
    assert int(count.item()) > 0
