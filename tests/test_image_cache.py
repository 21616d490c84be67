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
