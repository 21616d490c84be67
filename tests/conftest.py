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
