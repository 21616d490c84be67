"""N1: the LANCZOS thumbnail chain.  CPU: the oracle restatement against live Pillow and the C-ABI
geometry against the oracle.  GPU: the kernels (csrc/thumb.cu) against Pillow, bit for bit."""

import ctypes as C

import numpy as np
import pytest
from PIL import Image

from oracle import thumbnail as othumb

SHAPES = [  # (H, W, requested size)
    (300, 400, (64, 64)),          # factor 3, partial right-edge blocks
    (777, 1001, (100, 100)),       # factors (5, 4): unequal, partial blocks on both edges
    (512, 512, (128, 128)),        # 2x2
    (1024, 768, (128, 128)),       # 4x4
    (997, 2003, (200, 50)),        # (10, 9)
    (1200, 900, (1000, 1000)),     # no reduction, Lanczos from full resolution
    (3001, 2002, (37, 64)),        # factor 27
    (200, 300, (512, 512)),        # already small enough: returned unchanged
]


def _pil_thumbnail(a, size):
    im = Image.fromarray(a, "RGB")
    im.thumbnail(size, Image.LANCZOS)
    return np.asarray(im)


def _image(h, w, seed):
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w]
    base = (127 + 100 * np.sin(xx / 37.0) * np.cos(yy / 23.0))[..., None] + rng.normal(0, 40, (h, w, 3))
    a = np.clip(base, 0, 255).astype(np.uint8)
    a[: h // 7, : w // 5] = 255          # saturated region: exercises the clip of the Lanczos overshoot
    a[h // 2: h // 2 + 9] = 0
    return a


@pytest.mark.parametrize("h,w,size", SHAPES)
def test_oracle_matches_pillow(h, w, size):
    a = _image(h, w, h + w)
    want = _pil_thumbnail(a, size)
    got = othumb.thumbnail_u8(a, size)
    assert got.shape == want.shape and np.array_equal(got, want)


@pytest.mark.parametrize("f", [(2, 2), (3, 3), (4, 4), (5, 5), (6, 6), (8, 8), (24, 24), (24, 25), (2, 3)])
def test_oracle_reduce_matches_pillow(f):
    a = _image(53, 71, sum(f))
    assert np.array_equal(othumb.reduce_u8(a, *f), np.asarray(Image.fromarray(a, "RGB").reduce(f)))


def test_cabi_geometry_matches_oracle():
    from wsinsight_b200 import _cabi

    lib = _cabi.load()
    ow, oh, fx, fy = C.c_int(), C.c_int(), C.c_int(), C.c_int()
    cases = [(100000, 80000, 2048, 2048), (8192, 8192, 2048, 2048), (4096, 3072, 512, 512), (1001, 777, 100, 100),
             (123, 4567, 2048, 64), (4567, 123, 64, 2048), (50, 60, 100, 100), (99999, 3, 2048, 2048)]
    for (w, h, mx, my) in cases:
        assert lib.wsi_thumbnail_geometry(w, h, mx, my, C.byref(ow), C.byref(oh), C.byref(fx), C.byref(fy)) == 0
        size = othumb.thumbnail_size(w, h, (mx, my))
        assert (ow.value, oh.value) == size, (w, h, mx, my)
        assert (fx.value, fy.value) == othumb.reduce_factors(w, h, *size)
        assert lib.wsi_thumbnail_workspace_bytes(w, h, mx, my) > 0


@pytest.mark.gpu
@pytest.mark.parametrize("h,w,size", SHAPES + [(2048, 1537, (256, 256)), (1111, 4099, (512, 512))])
def test_device_thumbnail_bit_exact_vs_pillow(h, w, size):
    from wsinsight_b200.wsi import DeviceSlide

    a = _image(h, w, 3 * h + w)
    s = DeviceSlide.from_numpy(a, mpp=0.25)
    got = s.thumbnail_device(size).cpu().numpy()
    want = _pil_thumbnail(a, size)
    assert got.shape == want.shape
    assert np.array_equal(got, want), f"{int((got != want).sum())} of {got.size} values differ"


@pytest.mark.gpu
def test_device_thumbnail_synthetic_slide_and_get_thumbnail():
    """The slide protocol member the reference calls (pipeline.py:102) on a generated slide, vs Pillow on a host copy."""
    from wsinsight_b200.wsi import DeviceSlide

    s = DeviceSlide.synthetic(6000, 4500, mpp=0.5, seed=4)
    host = s.rgba[:, :s.width, :3].cpu().numpy()
    th = s.get_thumbnail((512, 512))
    assert th.mode == "RGB" and th.size == (512, 384)
    assert np.array_equal(np.asarray(th), _pil_thumbnail(np.ascontiguousarray(host), (512, 512)))


@pytest.mark.gpu
def test_device_thumbnail_unit_axis_factor_is_refused():
    from wsinsight_b200._cabi import WsiKernelError
    from wsinsight_b200.wsi import DeviceSlide

    s = DeviceSlide.from_numpy(_image(100, 3000, 1), mpp=0.25)      # factors (3, 1)... aspect rule keeps both axes
    ow, oh = othumb.thumbnail_size(3000, 100, (400, 400))
    fx, fy = othumb.reduce_factors(3000, 100, ow, oh)
    if (fx == 1) != (fy == 1):
        with pytest.raises(WsiKernelError):
            s.thumbnail_device((400, 400))
