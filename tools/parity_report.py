#!/usr/bin/env python
"""Run the end-to-end forward parity harness (tests/parity_harness.py) and write JSON records.

usage: python tools/parity_report.py OUTDIR [n_patches]      (on a GPU box; test/bench infrastructure)"""
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def main():
    import __graft_entry__ as ge

    ge.build()
    from parity_harness import run_parity

    out = Path(sys.argv[1])
    out.mkdir(parents=True, exist_ok=True)
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 10240
    for arch, K, nn, mode, batch in (("resnet34", 2, n, "cpu", 1024), ("resnet34", 1, min(n, 4096), "cpu", 1024),
                                     ("inceptionv4", 2, n, "cudnn_fp32", 512)):
        rec = run_parity(arch, K, n_patches=nn, batch=batch, oracle_mode=mode, log=lambda *a: print(*a, flush=True))
        (out / f"r02_parity_{arch}_k{K}.json").write_text(json.dumps(rec, indent=1))
        print(arch, K, json.dumps({k: rec[k] for k in ("spread_frac_p_in_0.05_0.95", "b200_resident",
                                                       "cudnn_bf16_channels_last",
                                                       "fp32_with_bf16_rounded_conv_weights_only", "seconds")}),
              flush=True)


if __name__ == "__main__":
    main()
