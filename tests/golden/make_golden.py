"""Generate tests/golden/purple_coords.json from the reference's golden CSVs.

Run in the build container only (reads /root/reference, which does not exist on the GPU box):

    python tests/golden/make_golden.py

The reference's disabled integration suite (tests/test_all.py:283-357) compares `wsinsight run`
on a uniform purple 4096x4096 slide at 0.25 um/px against tests/reference/<model>/purple.csv.
Only the coordinate columns are usable offline (the probabilities need the real zoo weights), so
this script keeps, per model: patch size and the ordered (minx, miny) list.
"""

import csv
import json
from pathlib import Path

REF = Path("/root/reference/tests/reference")
OUT = Path(__file__).resolve().parent / "purple_coords.json"

golden = {}
for d in sorted(REF.iterdir()):
    f = d / "purple.csv"
    if not f.exists():
        continue
    with open(f) as fh:
        rows = list(csv.DictReader(fh))
    widths = {int(r["width"]) for r in rows} | {int(r["height"]) for r in rows}
    assert len(widths) == 1, (d.name, widths)
    golden[d.name] = {
        "slide_width": 4096, "slide_height": 4096, "slide_mpp": 0.25, "rgb": [160, 32, 240],
        "patch_size": widths.pop(),
        "columns": list(rows[0].keys()),
        "coords": [[int(r["minx"]), int(r["miny"])] for r in rows],
    }
OUT.write_text(json.dumps(golden, indent=0, separators=(",", ":")))
print({k: (v["patch_size"], len(v["coords"])) for k, v in golden.items()})
