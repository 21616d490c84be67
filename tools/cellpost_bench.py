#!/usr/bin/env python
"""Time the device finalize (csrc/cellpost.cu) on one full-size job: 2048-px tile + 64 px padding of a nuclei field."""
import json
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def main():
    import torch

    import __graft_entry__ as ge

    ge.build()
    from conftest import synth_nuclei
    from wsinsight_b200.modellib.tilefuse import TileRemapStitcher

    K = 6
    base = synth_nuclei(272, 272, 0, K=K)
    reps = 8
    npm, hv, tp = (np.tile(a, (reps, reps) + (1,) * (a.ndim - 2)) for a in base)
    h, w = npm.shape
    st = TileRemapStitcher(K, w, h, 256, 0, 1.0, 1.0, min_object_size=10, canvas_tile=4096)
    t = st._tile(0, 0)
    t[0].copy_(torch.from_numpy(npm)); t[1].copy_(torch.from_numpy(hv)); t[2].copy_(torch.from_numpy(tp))
    out = {}
    for tile, pad in ((2048, 64),):
        ms = []
        for _ in range(3):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            boxes, probs, pxy, poff = st.finalize_arrays(tile_size=tile, padding_size=pad)
            torch.cuda.synchronize()
            ms.append((time.perf_counter() - t0) * 1e3)
        jobs = len(st.finalize_jobs(h, w, tile, pad))
        out[f"tile{tile}_pad{pad}"] = {"canvas": [h, w], "jobs": jobs, "instances": int(len(boxes)), "polygon_points": int(len(pxy)),
                                       "ms_runs": ms, "ms_per_job": min(ms) / jobs,
                                       "mpx_per_s": h * w / (min(ms) * 1e-3) / 1e6}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
