#!/usr/bin/env python
"""profiles/r2/traffic.json: DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) per dependency_kernel launch, what
bench.py reports as roofline.traffic.
  N = 1: from the `ncu --set full` report of every dependency launch of one full-size step (tools/measure_r2.sh step 3)
  N > 1: from the metrics-only captures of the middle band of the sharded frame alone on one GPU (step 4)

usage: traffic_r2.py <tag> [WxHxP boxes]   (reads gpurun_out/dep_fullhd_<tag>.ncu-rep and gpurun_out/traffic_band_n*_<tag>.csv)"""
import csv
import glob
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
TIME = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "usecond": 1e-3, "msecond": 1.0, "nsecond": 1e-6, "second": 1e3}


def num(x):
    try:
        return float(x.replace(",", ""))
    except ValueError:
        return float("nan")


def from_raw_page(txt, kernel):
    rows = list(csv.reader(io.StringIO(txt)))
    h, units = rows[0], rows[1]
    ir, iw, it, ik = h.index("dram__bytes_read.sum"), h.index("dram__bytes_write.sum"), h.index("gpu__time_duration.sum"), h.index("Kernel Name")
    out = []
    for r in rows[2:]:
        if kernel in r[ik]:
            out.append({"dram_bytes": num(r[ir]) * UNIT[units[ir]] + num(r[iw]) * UNIT[units[iw]], "ms": num(r[it]) * TIME[units[it]]})
    return out


def from_metric_csv(path, kernel):
    """`ncu --csv --log-file` of a --metrics run: one row per (launch, metric)."""
    lines = [l for l in open(path, errors="replace") if l.startswith('"')]
    rows = list(csv.DictReader(io.StringIO("".join(lines))))
    per = {}
    for r in rows:
        if kernel not in r["Kernel Name"]:
            continue
        d = per.setdefault(r["ID"], {"dram_bytes": 0.0, "ms": None})
        v, u = num(r["Metric Value"]), r["Metric Unit"]
        if r["Metric Name"].startswith("dram__bytes"):
            d["dram_bytes"] += v * UNIT.get(u, 1)
        elif r["Metric Name"] == "gpu__time_duration.sum":
            d["ms"] = v * TIME.get(u, 1e-6)
    return [per[k] for k in sorted(per, key=int)]


def entry(per, source, note):
    ok = [p["dram_bytes"] for p in per if p["dram_bytes"] == p["dram_bytes"] and p["dram_bytes"] > 0]
    return {"dram_bytes_per_launch": sum(ok) / len(ok), "launches_with_dram_counters": len(ok), "launches": per, "source": source, "note": note}


def main():
    tag = sys.argv[1]
    shape = sys.argv[2] if len(sys.argv) > 2 else "1920x1080x8"
    boxes = sys.argv[3] if len(sys.argv) > 3 else "55,35,17,7"
    out = os.path.join(ROOT, "profiles", "r2", "traffic.json")
    data = json.load(open(out)) if os.path.exists(out) else {}
    rep = os.path.join(ROOT, "gpurun_out", f"dep_fullhd_{tag}.ncu-rep")
    if os.path.exists(rep):
        txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        per = from_raw_page(txt, "dependency_kernel")
        data[f"{shape}:{boxes}:n1"] = {"dependency_kernel": entry(per, os.path.basename(rep), "ncu --set full --clock-control none, every dependency launch of one step")}
    for path in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", f"traffic_band_n*_{tag}.csv"))):
        n = int(re.search(r"_n(\d+)_", os.path.basename(path)).group(1))
        per = from_metric_csv(path, "dependency_kernel")
        if per:
            data[f"{shape}:{boxes}:n{n}"] = {"dependency_kernel": entry(per, os.path.basename(path),
                                             f"the middle band of {n} alone on one GPU (bench.py --band), ncu --metrics dram__bytes_*.sum --clock-control none")}
    json.dump(data, open(out, "w"), indent=1)
    for k, v in data.items():
        print(k, {kk: round(vv["dram_bytes_per_launch"] / 1e9, 3) for kk, vv in v.items()}, "GB per launch")


if __name__ == "__main__":
    main()
