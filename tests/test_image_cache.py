"""N4: the optional /images patch cache -- extraction from the slide and classification from the cache."""

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _slide():
    from wsinsight_b200.wsi import DeviceSlide

    return DeviceSlide.synthetic(3001, 2003, mpp=0.5, seed=11)


def test_extract_patches_matches_read_region():
    """reference patchlib/io.py:19-48: read_region(...).convert("RGB") per patch, zero padded outside the slide."""
    from wsinsight_b200.patchlib.io import extract_patches_from_slide

    s = _slide()
    coords = np.array([[0, 0], [100, 200], [2900, 1900], [3000, 2002], [1234, 567], [2826, 1828]], dtype=np.int32)
    ps = 175
    got = extract_patches_from_slide(s, coords, ps, as_numpy=True)
    assert got.shape == (len(coords), ps, ps, 3) and got.dtype == np.uint8
    for i, (x, y) in enumerate(coords):
        want = np.asarray(s.read_region((int(x), int(y)), 0, (ps, ps)).convert("RGB"))
        assert np.array_equal(got[i], want), f"patch {i} at {(x, y)}"
    assert got[3, 1:, :].max() == 0 and got[3, :, 1:].max() == 0       # only the corner pixel is inside the slide
    with pytest.raises(ValueError):
        extract_patches_from_slide(s, np.zeros((3, 4)), ps)
    assert extract_patches_from_slide(s, np.zeros((0, 2)), ps).shape == (0, ps, ps, 3)


@pytest.mark.parametrize("arch", ["resnet34"])
def test_classify_from_cache_identical_to_slide(arch):
    """Same pixels through data.py:243-268 (cache) and :270-281 (slide) give the same probabilities, bit for bit."""
    import torch

    from oracle import forward as oforward
    from wsinsight_b200.modellib.engine import B200Network
    from wsinsight_b200.modellib.run_inference import classify_patches_device, classify_patches_from_images
    from wsinsight_b200.modellib.transforms import make_compose_from_transform_config
    from wsinsight_b200.patchlib.io import extract_patches_from_slide

    s = _slide()
    ps = 175
    rng = np.random.default_rng(0)
    coords = np.stack([rng.integers(0, 3001 - ps, 21), rng.integers(0, 2003 - ps, 21)], 1).astype(np.int32)
    coords[0] = (2950, 1950)                        # partly outside: zero padded in both paths
    xywh = np.concatenate([coords, np.full_like(coords, ps)], 1)
    cfg = oforward.BREAST_RESNET34_CONFIG
    tf = make_compose_from_transform_config(cfg["transform"])
    net = B200Network.from_torchscript(torch.jit.freeze(torch.jit.script(oforward.make_resnet34(2, 0))), 2,
                                       input_size=224, max_batch=8)
    want = classify_patches_device(s, xywh, net, tf, batch_size=8).cpu()
    images = extract_patches_from_slide(s, coords, ps)
    got = classify_patches_from_images(images, net, tf, batch_size=8).cpu()
    assert torch.equal(got, want)
    got_np = classify_patches_from_images(images.cpu().numpy(), net, tf, batch_size=5).cpu()      # numpy input, ragged batches
    assert torch.equal(got_np, want)
    with pytest.raises(ValueError):
        classify_patches_from_images(np.zeros((2, 10, 12, 3), np.uint8), net, tf, 4)


@pytest.mark.parametrize("ps", [175, 100, 37, 350])
def test_fast_pack_equals_reference_pack(ps):
    """wsi_pack_patches_rgba_fast (shared-memory staged, any first patch) == wsi_pack_patches_rgba, bit for bit."""
    import torch

    from wsinsight_b200 import _cabi

    n = 9
    g = torch.Generator().manual_seed(ps)
    rgb = torch.randint(0, 256, (n * ps * ps * 3 + 64,), dtype=torch.uint8, generator=g).cuda()
    pitch = (ps + 3) // 4 * 4 * 4
    want = torch.zeros((n * ps, pitch), dtype=torch.uint8, device="cuda")
    got = torch.zeros_like(want)
    _cabi.call("wsi_pack_patches_rgba", rgb.data_ptr(), n, ps, pitch, want.data_ptr(), _cabi.stream_ptr())
    _cabi.call("wsi_pack_patches_rgba_fast", rgb.data_ptr(), 0, 4, ps, pitch, got.data_ptr(), _cabi.stream_ptr())
    _cabi.call("wsi_pack_patches_rgba_fast", rgb.data_ptr(), 4, n - 4, ps, pitch, got.data_ptr(), _cabi.stream_ptr())
    assert torch.equal(got, want)


def test_hostpipe_rgb_buffers_equal_resident_path():
    """Host-buffer entry with dense RGB patches (host_channels = 3): same probabilities, bit for bit, as the slide-resident
    path and as the RGBA host buffers."""
    import torch

    from oracle import forward as oforward
    from wsinsight_b200.modellib.engine import B200Network
    from wsinsight_b200.modellib.run_inference import HostPatchClassifier, classify_patches_device
    from wsinsight_b200.modellib.transforms import make_compose_from_transform_config
    from wsinsight_b200.patchlib.io import extract_patches_from_slide

    s = _slide()
    ps, n, B = 175, 300, 256        # 256 >= 4 * 64: the batch goes over in four pieces
    rng = np.random.default_rng(1)
    coords = np.stack([rng.integers(0, 3001 - ps, n), rng.integers(0, 2003 - ps, n)], 1).astype(np.int32)
    coords[0] = (2950, 1950)
    xywh = np.concatenate([coords, np.full_like(coords, ps)], 1)
    tf = make_compose_from_transform_config(oforward.BREAST_RESNET34_CONFIG["transform"])
    net = B200Network.from_torchscript(torch.jit.freeze(torch.jit.script(oforward.make_resnet34_calibrated(2, 0))), 2,
                                       input_size=224, max_batch=B)
    want = classify_patches_device(s, xywh, net, tf, batch_size=B).cpu()
    rgb = extract_patches_from_slide(s, coords, ps).cpu()
    for ch in (3, 4):
        hp = HostPatchClassifier(net, ps, tf, B, depth=2, host_channels=ch)
        out = torch.zeros((n, 2), dtype=torch.float32).pin_memory()
        for lo in range(0, n, B):
            hi = min(n, lo + B)
            px, _ = hp.alloc_host_batch(hi - lo)
            if ch == 3:
                px.copy_(rgb[lo:hi])
            else:
                px[:, :, :ps, :3].copy_(rgb[lo:hi])
                px[:, :, :ps, 3] = 255
            hp.submit(px, out[lo:hi])
            hp.sync()
        hp.close()
        assert torch.equal(out, want), f"host_channels={ch}: max diff {(out - want).abs().max()}"
