"""GPU parity: fused crop/resize/ToTensor/Normalize kernel vs PIL + torchvision (the reference's calls)."""

import numpy as np
import pytest
import torch

from oracle import preprocess

pytestmark = pytest.mark.gpu

MEAN, STD = [0.7238, 0.5716, 0.6779], [0.112, 0.1459, 0.1089]


def _slide(h, w, seed):
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 256, size=(h, w, 4), dtype=np.uint8)
    # smooth-ish content + noise so both flat and high-frequency regions occur
    a[: h // 2] = (a[: h // 2] // 8) + 100
    return a


def _coords(h, w, ps, n, seed):
    rng = np.random.default_rng(seed)
    xs = rng.integers(-ps // 2, w - ps // 2, n)
    ys = rng.integers(-ps // 2, h - ps // 2, n)
    c = np.stack([xs, ys], 1).astype(np.int32)
    c[0] = (0, 0)
    c[1] = (w - ps, h - ps)
    c[2] = (w - ps // 3, h - ps // 3)       # overhangs bottom/right: zero padded
    c[3] = (-ps // 3, -ps // 4)             # overhangs top/left
    return c


@pytest.mark.parametrize("ps,S", [(350, 224), (175, 224), (700, 224), (100, 299), (200, 299), (224, 224), (37, 64)])
def test_resize_u8_bit_exact_vs_pil(ps, S):
    from wsinsight_b200.modellib.data import PreprocessPlan
    from wsinsight_b200.modellib.transforms import FusedTransform
    from wsinsight_b200.wsi import DeviceSlide

    h, w = 900, 1100
    host = _slide(h, w, ps + S)
    slide = DeviceSlide.from_numpy(host, 0.25)
    plan = PreprocessPlan(slide, ps, FusedTransform(S, MEAN, STD))
    coords = _coords(h, w, ps, 12, ps)
    got = plan.run_u8(torch.from_numpy(coords).cuda()).cpu().numpy()
    assert got[..., 3].max() == 0
    for i, (x, y) in enumerate(coords.tolist()):
        crop = preprocess.read_region_rgba(host, x, y, ps, ps)[..., :3]
        want = preprocess.resize_lib_u8(np.ascontiguousarray(crop), S) if ps != S else crop
        bad = np.argwhere(got[i, ..., :3] != want)
        assert len(bad) == 0, f"patch {i} at {(x, y)}: {len(bad)} mismatches, first {bad[:4].tolist()} plan {plan.info()}"


@pytest.mark.parametrize("ps,S,pad", [(350, 224, 3), (175, 224, 3), (100, 299, 0)])
def test_bf16_batch_vs_torchvision(ps, S, pad):
    from wsinsight_b200.modellib.data import PreprocessPlan
    from wsinsight_b200.modellib.transforms import FusedTransform
    from wsinsight_b200.wsi import DeviceSlide

    h, w = 800, 900
    host = _slide(h, w, 7)
    slide = DeviceSlide.from_numpy(host, 0.25)
    plan = PreprocessPlan(slide, ps, FusedTransform(S, MEAN, STD), S + 2 * pad, S + 2 * pad, pad, pad)
    coords = _coords(h, w, ps, 8, 3)
    out = torch.zeros((8, S + 2 * pad, S + 2 * pad, 4), dtype=torch.bfloat16, device="cuda")
    plan.run(torch.from_numpy(coords).cuda(), out)
    got = out.float().cpu().numpy()
    if pad:
        assert got[:, :pad].max() == 0 and got[:, :, :pad].max() == 0 and got[:, -pad:].max() == 0
    assert np.abs(got[..., 3]).max() == 0
    want = preprocess.preprocess_batch_lib(host, np.concatenate([coords, np.full_like(coords, ps)], 1), S, MEAN, STD)
    want_bf16 = want.to(torch.bfloat16).float().permute(0, 2, 3, 1).numpy()     # RNE, same as the kernel
    inner = got[:, pad:pad + S, pad:pad + S, :3]
    assert np.array_equal(inner, want_bf16), f"max abs diff {np.abs(inner - want_bf16).max()}"


def test_synthetic_slide_host_device_identical():
    from wsinsight_b200.wsi import DeviceSlide, synth_window_host

    s = DeviceSlide.synthetic(3000, 2000, 0.5, seed=5)
    dev = s.rgba[:, :3000].cpu().numpy()
    for (x, y, w, h) in [(0, 0, 64, 64), (2900, 1900, 100, 100), (1234, 777, 333, 50)]:
        assert np.array_equal(synth_window_host(x, y, w, h, 3000, 2000, 5), dev[y:y + h, x:x + w])
    assert (dev[..., 3] == 255).all()
    th = s.thumbnail_device((256, 256), method="box").cpu().numpy()
    assert th.shape == (171, 256, 3)
    # exact integer box mean (the stand-in thumbnail; the Pillow-exact one is covered by tests/test_thumbnail.py)
    j, i = 100, 50
    x0, x1, y0, y1 = j * 3000 // 256, (j + 1) * 3000 // 256, i * 2000 // 171, (i + 1) * 2000 // 171
    box = dev[y0:y1, x0:x1, :3].reshape(-1, 3).astype(np.int64)
    assert th[i, j].tolist() == ((box.sum(0) + len(box) // 2) // len(box)).tolist()
