"""GeoJSON box overlays from the per-patch CSVs.

Mirrors the behaviour of the reference's tile exporter (``wsinsight/write_geojson.py:39-143`` box path,
``:199-254`` CSV builder, ``write_geojsons``) for the outputs of the patch-classification path: one QuPath
``Feature`` per patch with the shrunken (overlap-aware) box ring, the ``prob_*`` measurements and, optionally,
the arg-max classification with a distinct colour.  The polygon/WKT branch of the reference belongs to the
cell-detection path (out of scope, needs shapely/geopandas) and raises ``NotImplementedError`` here.

Pure host code: the ring arithmetic is a handful of vectorised integer operations on (N, 4) coordinates and
the time goes into JSON serialisation, so there is no kernel behind this module.
"""

from __future__ import annotations

import json
import uuid
from colorsys import hsv_to_rgb
from pathlib import Path
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

try:  # the reference serialises with orjson; fall back to the stdlib encoder (same JSON document)
    import orjson as _orjson

    def _dumps(obj: dict) -> bytes:
        return _orjson.dumps(obj)
except Exception:  # pragma: no cover - depends on the image
    def _dumps(obj: dict) -> bytes:
        return json.dumps(obj, separators=(",", ":")).encode()


def _make_distinct_colors(n: int, s: float = 0.70, v: float = 0.90, shuffle: bool = True,
                          seed: Optional[int] = None) -> List[dict]:
    """``n`` evenly spaced hues (reference ``write_geojson.py:39-65``): with ``shuffle`` the hues are taken
    alternately from both ends so that neighbouring classes differ; ``seed`` is accepted and unused."""
    if n <= 0:
        raise ValueError("n must be > 0")
    idx = list(range(n))
    if shuffle and n > 2:
        lo, hi, order = 0, n - 1, []
        while lo <= hi:
            order.append(lo)
            if lo != hi:
                order.append(hi)
            lo, hi = lo + 1, hi - 1
        idx = order
    colors = []
    for i in idx:
        h = i / n
        rgb = tuple(int(round(c * 255)) for c in hsv_to_rgb(h, s, v))
        colors.append({"hex": "#%02X%02X%02X" % rgb, "rgb": rgb, "hsv": (h, s, v)})
    return colors


def box_rings(xywh: np.ndarray, overlap: float) -> np.ndarray:
    """(N, 4) ``minx, miny, width, height`` -> closed rings (N, 5, 2) int64 (reference ``:85-105``).

    The patch is shrunk to its non-overlapping core: side ``rint(w * (1 - overlap))`` centred with a
    ``rint`` margin; ring order is (maxx,miny) (maxx,maxy) (minx,maxy) (minx,miny) (maxx,miny)."""
    xywh = np.asarray(xywh, dtype=np.int64).reshape(-1, 4)
    minx, miny, w, h = xywh.T
    pw = np.rint(w * (1.0 - overlap)).astype(np.int64)
    ph = np.rint(h * (1.0 - overlap)).astype(np.int64)
    x0 = minx + np.rint((w - pw) * 0.5).astype(np.int64)
    y0 = miny + np.rint((h - ph) * 0.5).astype(np.int64)
    x1, y1 = x0 + pw, y0 + ph
    xs = np.stack([x1, x1, x0, x0, x1], axis=1)
    ys = np.stack([y0, y1, y1, y0, y0], axis=1)
    return np.stack([xs, ys], axis=2)


def _class_names(prob_cols: Sequence[str], prefix: str) -> List[str]:
    return [c if c.startswith(f"{prefix}_") else f"{prefix}_{c}" for c in prob_cols]


def features_from_arrays(xywh: np.ndarray, probs: np.ndarray, prob_cols: Sequence[str], overlap: float, *,
                         prefix: str = "prob", object_type: str = "tile", set_classification: bool = False,
                         color_list: Optional[List[dict]] = None) -> dict:
    """FeatureCollection straight from the arrays the device path produces ((N,4) coords, (N,K) probs)."""
    xywh = np.asarray(xywh).reshape(-1, 4)
    probs = np.asarray(probs, dtype=np.float32).reshape(len(xywh), len(prob_cols))
    keep = ~np.isnan(probs).any(axis=1)            # the reference drops rows with a missing probability
    xywh, probs = xywh[keep], probs[keep]
    rings = box_rings(xywh, overlap)
    arg = probs.argmax(axis=1) if len(probs) else np.zeros(0, dtype=np.int64)
    if color_list is None:
        color_list = _make_distinct_colors(len(prob_cols))
    names = _class_names(prob_cols, prefix)
    feats = []
    for i in range(len(xywh)):
        props = {"isLocked": True,
                 "measurements": {c: float(probs[i, j]) for j, c in enumerate(prob_cols)},
                 "objectType": object_type}
        if set_classification:
            k = int(arg[i])
            props["classification"] = {"name": names[k], "color": list(color_list[k]["rgb"])}
        feats.append({"type": "Feature", "id": str(uuid.uuid4()),
                      "geometry": {"type": "Polygon", "coordinates": [rings[i].tolist()]},
                      "properties": props})
    return {"type": "FeatureCollection", "features": feats}


def _dataframe_to_geojson_box_fast(df, prob_cols: List[str], overlap: float, *, prefix: str = "prob",
                                   object_type: str = "tile", set_classification: bool = False,
                                   color_list: Optional[List[dict]] = None) -> dict:
    """Same signature as the reference (``:70-143``); ``df`` is a pandas DataFrame with the CSV columns."""
    xywh = df[["minx", "miny", "width", "height"]].to_numpy(dtype=np.int64)
    probs = df[list(prob_cols)].to_numpy(dtype=np.float32)
    return features_from_arrays(xywh, probs, prob_cols, overlap, prefix=prefix, object_type=object_type,
                                set_classification=set_classification, color_list=color_list)


def _build_geojson_dict_from_csv(csv, *, overlap: float, results_dir, output_dir, prefix: str = "prob",
                                 object_type: str = "tile", set_classification: bool = False,
                                 annotation_shape: str = "box", usecols=None, dtype=None) -> Tuple[Path, dict]:
    """CSV -> (destination path, FeatureCollection) (reference ``:199-254``)."""
    import pandas as pd

    df = pd.read_csv(csv, usecols=usecols, dtype=dtype)
    prob_cols = [c for c in df.columns if c.startswith(f"{prefix}_")]
    if not prob_cols:
        raise KeyError(f"No {prefix}_* columns in {csv}")
    if annotation_shape != "box":
        if "polygon_wkt" not in df.columns:
            raise KeyError("polygon_wkt column is required for annotation_shape='polygon'")
        raise NotImplementedError("polygon annotations belong to the cell-detection path (out of scope)")
    gj = _dataframe_to_geojson_box_fast(df, prob_cols, overlap, prefix=prefix, object_type=object_type,
                                        set_classification=set_classification,
                                        color_list=_make_distinct_colors(len(prob_cols)))
    return Path(results_dir) / Path(output_dir) / f"{Path(csv).stem}.geojson", gj


def write_geojsons(csvs: Sequence, overlap: float, results_dir, output_dir="model-outputs-geojson", prefix: str = "prob",
                   num_workers: int = 1, object_type: str = "tile", set_classification: bool = False,
                   annotation_shape: str = "box", show_progress: bool = False, print_timings: bool = False) -> List[Path]:
    """One ``<stem>.geojson`` per CSV under ``results_dir/output_dir`` (serial: the device path produces one CSV
    per slide and the exporter is not on the timed path)."""
    written = []
    for csv in csvs:
        out, gj = _build_geojson_dict_from_csv(csv, overlap=overlap, results_dir=results_dir, output_dir=output_dir,
                                               prefix=prefix, object_type=object_type,
                                               set_classification=set_classification, annotation_shape=annotation_shape)
        out.parent.mkdir(parents=True, exist_ok=True)
        out.write_bytes(_dumps(gj))
        written.append(out)
    return written
