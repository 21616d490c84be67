"""Density grids: kernel vs the numpy statement, and the all-reduce over ranks (gloo on CPU)."""

import os

import numpy as np
import pytest
import torch

from oracle import density as oden


def _objects(n, w, h, k, seed):
    rng = np.random.default_rng(seed)
    xywh = np.stack([rng.integers(-40, w, n), rng.integers(-40, h, n), rng.integers(1, 400, n), rng.integers(1, 400, n)],
                    1).astype(np.int32)
    xywh[: n // 4, 2:] = 175                      # odd sizes: centres on .5, rounded half to even
    xywh[n // 4: n // 2, 2:] = 350
    probs = rng.random((n, k)).astype(np.float32)
    probs[::7] = 0.5                              # exact ties: first maximum wins
    return xywh, probs


def test_oracle_centres_round_half_to_even():
    cx, cy = oden.centers(np.array([[0, 1, 175, 175], [1, 2, 175, 175], [10, 10, 350, 100]]))
    assert cx.tolist() == [88, 88, 185] and cy.tolist() == [88, 90, 60]
    g, dropped = oden.density_grid(np.array([[0, 0, 10, 10], [95, 95, 20, 20]]), np.array([[.2, .8], [.9, .1]]), 100, 100, 50)
    assert g.shape == (2, 2, 2) and g[1, 0, 0] == 1 and dropped == 1


@pytest.mark.gpu
@pytest.mark.parametrize("n,k,bin_px", [(50000, 2, 100), (1000, 1, 64), (33, 5, 7), (200000, 3, 1000)])
def test_density_kernel_matches_oracle(n, k, bin_px):
    from wsinsight_b200.insightlib import density_grid_device

    w, h = 20000, 15000
    xywh, probs = _objects(n, w, h, k, n + k)
    grid, dropped = density_grid_device(torch.from_numpy(xywh).cuda(), torch.from_numpy(probs).cuda(), w, h, bin_px)
    want, want_dropped = oden.density_grid(xywh, probs, w, h, bin_px)
    assert np.array_equal(grid.cpu().numpy(), want)
    assert int(dropped.item()) == want_dropped
    assert int(grid.sum().item()) + want_dropped == n


def _worker(rank, world, port, tmp):
    import torch.distributed as dist

    from wsinsight_b200.insightlib import allreduce_grids

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    xywh, probs = _objects(4000, 5000, 4000, 2, 9)
    lo, hi = len(xywh) * rank // world, len(xywh) * (rank + 1) // world
    part, _ = oden.density_grid(xywh[lo:hi], probs[lo:hi], 5000, 4000, 250)     # stand-in for the kernel on CPU
    total = allreduce_grids(torch.from_numpy(part))
    full, _ = oden.density_grid(xywh, probs, 5000, 4000, 250)
    ok = np.array_equal(total.numpy(), full)
    open(os.path.join(tmp, f"ok{rank}"), "w").write("1" if ok else "0")
    dist.destroy_process_group()


def test_allreduce_grids_world2_gloo(tmp_path):
    import torch.multiprocessing as mp

    port = 29650 + os.getpid() % 200
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok0").read_text() == "1" and (tmp_path / "ok1").read_text() == "1"
