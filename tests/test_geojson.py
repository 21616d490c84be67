"""GeoJSON exporter: mirrors the reference's live tests (tests/test_all.py:60-168)."""

import json

import numpy as np
import pandas as pd
import pytest

from oracle import geojson as ogeo
from wsinsight_b200 import write_geojson as wg


def _write_minimal_csv(path):
    path.parent.mkdir(parents=True, exist_ok=True)
    pd.DataFrame({"minx": [0, 350], "miny": [0, 350], "width": [350, 350], "height": [350, 350],
                  "prob_Other": [0.9, 0.2], "prob_Tumor": [0.1, 0.8]}).to_csv(path, index=False)


def test_make_distinct_colors_unique_hex():
    colors = wg._make_distinct_colors(5, seed=123, shuffle=False)
    assert len(colors) == 5 and len({c["hex"] for c in colors}) == 5
    for c in colors:
        assert c["hex"].startswith("#") and len(c["hex"]) == 7
        assert all(0 <= ch <= 255 for ch in c["rgb"])
    with pytest.raises(ValueError):
        wg._make_distinct_colors(0)
    # shuffle takes hues alternately from both ends (reference :47-55)
    assert [round(c["hsv"][0] * 5) for c in wg._make_distinct_colors(5)] == [0, 4, 1, 3, 2]


def test_dataframe_to_geojson_box_fast_sets_properties():
    df = pd.DataFrame({"minx": [0, 20], "miny": [0, 20], "width": [10, 10], "height": [10, 10],
                       "prob_background": [0.25, 0.75], "prob_tumor": [0.75, 0.25]})
    res = wg._dataframe_to_geojson_box_fast(df, prob_cols=["prob_background", "prob_tumor"], overlap=0.0,
                                            set_classification=True)
    assert res["type"] == "FeatureCollection" and len(res["features"]) == len(df)
    names = []
    for f in res["features"]:
        p = f["properties"]
        assert set(p["measurements"]) == {"prob_background", "prob_tumor"}
        assert p["classification"]["name"].startswith("prob_") and p["isLocked"] is True and p["objectType"] == "tile"
        assert f["geometry"]["type"] == "Polygon"
        names.append(p["classification"]["name"])
    assert names == ["prob_tumor", "prob_background"]
    assert res["features"][0]["geometry"]["coordinates"] == [[[10, 0], [10, 10], [0, 10], [0, 0], [10, 0]]]


@pytest.mark.parametrize("overlap", [0.0, 0.1, 0.25, 0.5, 0.333])
def test_box_rings_match_oracle(overlap):
    rng = np.random.default_rng(int(overlap * 1000))
    xywh = np.concatenate([rng.integers(0, 100000, (200, 2)), rng.integers(1, 900, (200, 1)).repeat(2, 1)], 1)
    xywh[:50, 3] = rng.integers(1, 900, 50)            # non-square too
    got = wg.box_rings(xywh, overlap)
    want = np.array([ogeo.box_ring(*map(int, r), overlap) for r in xywh])
    assert got.dtype == np.int64 and np.array_equal(got, want)


def test_nan_rows_dropped_and_empty():
    xywh = np.array([[0, 0, 10, 10], [10, 0, 10, 10]])
    probs = np.array([[0.5, 0.5], [np.nan, 0.2]], dtype=np.float32)
    gj = wg.features_from_arrays(xywh, probs, ["prob_a", "prob_b"], 0.0)
    assert len(gj["features"]) == 1
    assert wg.features_from_arrays(np.zeros((0, 4)), np.zeros((0, 2)), ["prob_a", "prob_b"], 0.0)["features"] == []


def test_build_geojson_dict_from_csv_box(tmp_path):
    results = tmp_path / "results"
    csv = results / "model-outputs-csv" / "sample.csv"
    _write_minimal_csv(csv)
    from pathlib import Path
    out, gj = wg._build_geojson_dict_from_csv(csv, overlap=0.0, results_dir=results, output_dir=Path("geojson"),
                                              prefix="prob", object_type="tile", set_classification=True,
                                              annotation_shape="box")
    assert out == results / "geojson" / "sample.geojson"
    assert gj["type"] == "FeatureCollection" and len(gj["features"]) == 2
    with pytest.raises(KeyError):
        wg._build_geojson_dict_from_csv(csv, overlap=0.0, results_dir=results, output_dir="g", prefix="score")
    with pytest.raises(KeyError):
        wg._build_geojson_dict_from_csv(csv, overlap=0.0, results_dir=results, output_dir="g",
                                        annotation_shape="polygon")


def test_write_geojsons_creates_output(tmp_path):
    results = tmp_path / "results"
    for s in ("slide_a", "slide_b"):
        _write_minimal_csv(results / "model-outputs-csv" / f"{s}.csv")
    wg.write_geojsons(csvs=sorted((results / "model-outputs-csv").glob("*.csv")), overlap=0.0, results_dir=results,
                      output_dir="geojson-out", set_classification=True)
    for s in ("slide_a", "slide_b"):
        data = json.loads((results / "geojson-out" / f"{s}.geojson").read_text())
        assert data["features"] and data["features"][1]["properties"]["classification"]["name"] == "prob_Tumor"
