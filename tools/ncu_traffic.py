"""Summarise an `ncu --set full` capture of the conv kernels into profiles/<ROUND>_ncu_conv_traffic.json (ROUND from the
environment variable WSI_ROUND, default r02).

usage: python tools/ncu_traffic.py <model> <report.ncu-rep> [<csv out>]
Reads dram__bytes_read.sum + dram__bytes_write.sum and gpu__time_duration.sum per launch (ncu --page raw --csv)
and stores the per-launch average that bench.py reports as roofline.traffic.
"""
import csv
import io
import json
import os
import subprocess
import sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0,
        "usecond": 1e-6, "msecond": 1e-3, "nsecond": 1e-9, "second": 1.0}


def main():
    model, rep = sys.argv[1], sys.argv[2]
    metrics = ("gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,"
               "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active,"
               "l1tex__m_xbar2l1tex_read_bytes.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed")
    if rep.endswith(".csv"):      # exported on the GPU box: `--page raw --csv` (wide) or a `--metrics ... --csv` log (long)
        out = open(rep).read()
        out = out[out.index('"ID"'):]
    else:
        out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--metrics", metrics], capture_output=True,
                             text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    if "Metric Name" in rows[0]:  # long format -> wide
        c = {h: i for i, h in enumerate(rows[0])}
        wide, unit_of = {}, {}
        for r in rows[1:]:
            if len(r) < len(rows[0]):
                continue
            w = wide.setdefault(r[c["ID"]], {"Kernel Name": r[c["Kernel Name"]]})
            w[r[c["Metric Name"]]] = r[c["Metric Value"]].replace(",", "")
            unit_of[r[c["Metric Name"]]] = r[c["Metric Unit"]]
        names = ["Kernel Name"] + sorted(unit_of)
        rows = [names, [""] + [unit_of[n] for n in names[1:]]] + [[w.get(n, "0") for n in names] for w in wide.values()]
    if len(sys.argv) > 3:
        keep = [i for i, h in enumerate(rows[0]) if h in ("ID", "Kernel Name", "Grid Size", "Block Size") or h in metrics.split(",")]
        with open(sys.argv[3], "w", newline="") as f:
            csv.writer(f).writerows([[r[i] for i in keep] for r in rows])
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}

    def val(r, name):
        return float(r[col[name]].replace(",", "")) * UNIT.get(units[col[name]], 1.0)

    per = {}
    tot_b = tot_t = 0.0
    for r in data:
        k = r[col["Kernel Name"]].split("(")[0].replace("void ", "")
        b = val(r, "dram__bytes_read.sum") + val(r, "dram__bytes_write.sum")
        t = val(r, "gpu__time_duration.sum")
        d = per.setdefault(k, {"launches": 0, "dram_bytes": 0.0, "time_s": 0.0})
        d["launches"] += 1
        d["dram_bytes"] += b
        d["time_s"] += t
        tot_b += b
        tot_t += t
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "profiles", os.environ.get("WSI_ROUND", "r02") + "_ncu_conv_traffic.json")
    try:
        j = json.load(open(path))
    except Exception:
        j = {}
    j[model] = {"dram_bytes_per_launch_avg": tot_b / len(data), "launches": len(data),
                "ncu_time_per_launch_avg_us": tot_t / len(data) * 1e6, "per_kernel": per,
                "source": os.path.basename(rep), "note": "ncu --set full --clock-control none, one forward at batch 1024"}
    json.dump(j, open(path, "w"), indent=1)
    print(json.dumps(j[model])[:600])


if __name__ == "__main__":
    main()
