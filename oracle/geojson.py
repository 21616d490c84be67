"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's GeoJSON box ring arithmetic.

Follows wsinsight/write_geojson.py:85-105 (shrunken, overlap-aware box; closed ring of five vertices) one
row at a time in plain Python.  Pin: the reference's own live tests (tests/test_all.py:71-168) check structure
only (feature count, property keys, geometry type); the numeric ring values are pinned by this restatement.
"""

from __future__ import annotations


def _rint(x: float) -> int:
    # numpy.rint: round half to even
    return int(round(x))


def box_ring(minx: int, miny: int, w: int, h: int, overlap: float):
    pw, ph = _rint(w * (1.0 - overlap)), _rint(h * (1.0 - overlap))
    x0, y0 = minx + _rint((w - pw) * 0.5), miny + _rint((h - ph) * 0.5)
    x1, y1 = x0 + pw, y0 + ph
    return [[x1, y0], [x1, y1], [x0, y1], [x0, y0], [x1, y0]]
