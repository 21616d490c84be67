"""Per-class density heat-map grids of classified patches / cells (BASELINE config 5).

The reference ships no heat-map code; the grid is defined from its one upstream primitive, the object
centre ``np.rint(minx + width / 2)`` (``wsinsight/insightlib/insight_helpers.py:26-27``): object *i* adds 1 to
``grid[argmax_k prob[i, k]][center_y // bin_px][center_x // bin_px]``.  Each rank bins its own shard of the objects
on its GPU (``csrc/density.cu``); the grids are summed with ONE NCCL all-reduce -- the only collective of the path
besides the gather of the probabilities (north_star).
"""

from __future__ import annotations

import ctypes as C

from .. import _cabi


def grid_shape(slide_width: int, slide_height: int, bin_px: int):
    """(Gy, Gx) bins covering the slide."""
    if bin_px < 1:
        raise ValueError("bin_px must be >= 1")
    return (-(-int(slide_height) // int(bin_px)), -(-int(slide_width) // int(bin_px)))


def density_grid_device(xywh, probs, slide_width: int, slide_height: int, bin_px: int, out=None):
    """Bin this rank's objects.  ``xywh``: CUDA int32 (n, 4); ``probs``: CUDA float32 (n, K) (or (n,) for K == 1).

    Returns ``(grid, dropped)``: int32 CUDA tensor (K, Gy, Gx) and the number of objects whose centre lies outside
    the slide.  ``out`` accumulates into an existing grid."""
    torch = _cabi.require_cuda()
    if probs.dim() == 1:
        probs = probs[:, None]
    n, k = int(probs.shape[0]), int(probs.shape[1])
    if xywh.shape != (n, 4) or xywh.dtype != torch.int32 or probs.dtype != torch.float32:
        raise ValueError("xywh must be int32 (n, 4) and probs float32 (n, K)")
    xywh, probs = xywh.contiguous(), probs.contiguous()
    gy, gx = grid_shape(slide_width, slide_height, bin_px)
    grid = torch.zeros((k, gy, gx), dtype=torch.int32, device=probs.device) if out is None else out
    if tuple(grid.shape) != (k, gy, gx) or grid.dtype != torch.int32:
        raise ValueError("out has the wrong shape or dtype")
    dropped = torch.zeros(1, dtype=torch.int32, device=probs.device)
    with torch.cuda.device(probs.device):
        _cabi.call("wsi_density_accumulate", xywh.data_ptr(), probs.data_ptr(), n, k, int(bin_px), gx, gy,
                   grid.data_ptr(), dropped.data_ptr(), _cabi.stream_ptr())
    return grid, dropped


def allreduce_grids(grid, group=None):
    """Sum the per-rank grids in place over the process group (NCCL on GPUs, gloo in the CPU tests)."""
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(grid, op=dist.ReduceOp.SUM, group=group)
    return grid
