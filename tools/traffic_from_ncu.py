#!/usr/bin/env python
"""profiles/r1/traffic.json from an `ncu --set full` report: DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) per
launch of a kernel, averaged over the captured launches.

usage: traffic_from_ncu.py <report.ncu-rep> <kernel name> <workload key, e.g. 1280x720x8:55,35,17,7>"""
import csv
import io
import json
import os
import subprocess
import sys

UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def main():
    rep, kernel, key = sys.argv[1:4]
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    h, units = rows[0], rows[1]
    ir, iw, it, ik = h.index("dram__bytes_read.sum"), h.index("dram__bytes_write.sum"), h.index("gpu__time_duration.sum"), h.index("Kernel Name")
    per = []
    for r in rows[2:]:
        if kernel not in r[ik]:
            continue
        b = float(r[ir].replace(",", "")) * UNIT[units[ir]] + float(r[iw].replace(",", "")) * UNIT[units[iw]]
        per.append({"dram_bytes": b, "duration": r[it] + " " + units[it]})
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "r1", "traffic.json")
    data = json.load(open(out)) if os.path.exists(out) else {}
    ok = [p["dram_bytes"] for p in per if p["dram_bytes"] == p["dram_bytes"]]  # ncu leaves -nan where a replay pass was cut short
    data.setdefault(key, {})[kernel] = {"dram_bytes_per_launch": sum(ok) / len(ok), "launches_with_dram_counters": len(ok), "launches": per,
                                        "source": os.path.basename(rep), "note": "ncu --set full --clock-control none, replayed launches"}
    json.dump(data, open(out, "w"), indent=1)
    print(json.dumps(data[key][kernel])[:400])


if __name__ == "__main__":
    main()
