import json
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden_purple():
    return json.loads((ROOT / "tests" / "golden" / "purple_coords.json").read_text())


def blob_mask(h, w, seed, density=0.5, smooth=9):
    """Random blob-structured boolean mask (smoothed noise threshold) with holes and islands."""
    import cv2

    rng = np.random.default_rng(seed)
    f = rng.random((h, w)).astype(np.float32)
    f = cv2.GaussianBlur(f, (0, 0), smooth)
    thr = np.quantile(f, 1.0 - density)
    return f > thr


def blob_thumbnail(h, w, seed, density=0.5, smooth=9):
    """RGB thumbnail: H&E-like noisy tissue on near-white background + salt noise near threshold."""
    rng = np.random.default_rng(seed + 1000)
    m = blob_mask(h, w, seed, density, smooth)
    img = np.empty((h, w, 3), dtype=np.uint8)
    bg = rng.integers(236, 250, size=(h, w, 3))
    fg = np.stack([rng.integers(150, 230, (h, w)), rng.integers(60, 150, (h, w)), rng.integers(140, 210, (h, w))], -1)
    img[:] = np.where(m[..., None], fg, bg)
    # weakly saturated speckle so the median / threshold stage is exercised around S == 7
    speck = rng.random((h, w)) < 0.08
    img[speck] = rng.integers(200, 256, size=(int(speck.sum()), 3))
    return img


def synth_nuclei(h, w, seed, K=5, n=None):
    """HoVer-Net-like maps of a field of elliptical nuclei (some touching): nuclei probability (h, w), horizontal /
    vertical distance maps (h, w, 2) in [-1, 1] per nucleus, type probabilities (h, w, K).  float32."""
    import cv2

    rng = np.random.default_rng(seed)
    n = n if n is not None else max(3, h * w // 450)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    owner = np.full((h, w), -1, np.int32)
    hv = np.zeros((h, w, 2), np.float32)
    cls = rng.integers(0, K, n)
    for i in range(n):
        cx, cy = rng.uniform(0, w), rng.uniform(0, h)
        rx, ry = rng.uniform(3.5, 9.0), rng.uniform(3.5, 9.0)
        th = rng.uniform(0, np.pi)
        dx, dy = xx - cx, yy - cy
        u = (dx * np.cos(th) + dy * np.sin(th)) / rx
        v = (-dx * np.sin(th) + dy * np.cos(th)) / ry
        inside = (u * u + v * v <= 1.0) & (owner < 0)
        owner[inside] = i
        ext_x = max(rx, ry)
        hv[inside, 0] = np.clip(dx[inside] / ext_x, -1, 1)
        hv[inside, 1] = np.clip(dy[inside] / ext_x, -1, 1)
    fg = owner >= 0
    npm = np.where(fg, 0.92, 0.06).astype(np.float32) + rng.normal(0, 0.03, (h, w)).astype(np.float32)
    npm = cv2.GaussianBlur(npm, (3, 3), 0.6)
    hv += rng.normal(0, 0.02, hv.shape).astype(np.float32)
    tp = rng.uniform(0.02, 0.12, (h, w, K)).astype(np.float32)
    tp[fg, cls[owner[fg]]] += 0.8
    tp /= tp.sum(-1, keepdims=True)
    return np.clip(npm, 0, 1).astype(np.float32), hv.astype(np.float32), tp.astype(np.float32)
