#!/usr/bin/env python
"""Attribute an ncu SASS-level source page to CUDA source lines.

usage: ncu_by_line.py <report.ncu-rep> <cubin> <kernel-substring> [top_n] [mangled-substring]
(the last argument selects one template instantiation in the cubin when the kernel name alone is ambiguous)
Joins `ncu --page source --csv` (per-SASS-instruction executed counts and stall samples) with the
`//## File ..., line N` annotations of `nvdisasm -g` (built with -lineinfo) by instruction order."""
import csv
import io
import re
import subprocess
import sys
from collections import defaultdict


def main():
    rep, cubin, kern = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    dkey = sys.argv[5] if len(sys.argv) > 5 else kern
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "-k", f"regex:{kern}"], capture_output=True,
                         text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hi = next(i for i, r in enumerate(rows) if "Source" in r and "Address" in r)
    hdr = rows[hi]
    ci, cs, csrc = hdr.index("Instructions Executed"), hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Source")
    sass = []
    for r in rows[hi + 1:]:  # first matching launch only (a report may hold several kernels / launches)
        if len(r) <= max(ci, cs) or not r[ci].isdigit():
            if sass:
                break
            continue
        sass.append((r[csrc].strip(), int(r[ci]), int(r[cs])))
    dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
    lines, cur, inside = [], None, False
    for ln in dis:
        if ln.startswith(".text."):
            inside = dkey in ln
            continue
        if not inside:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(.*?);", ln)
        if m:
            lines.append((cur, m.group(1).strip()))
    if len(lines) != len(sass):
        print(f"warning: {len(lines)} disassembled vs {len(sass)} profiled instructions", file=sys.stderr)
    agg = defaultdict(lambda: [0, 0, 0])
    for (loc, _), (_, ne, ns) in zip(lines, sass):
        a = agg[loc]
        a[0] += ne
        a[1] += ns
        a[2] += 1
    ti, ts = sum(a[0] for a in agg.values()), sum(a[1] for a in agg.values())
    src_cache = {}
    print(f"total warp instructions {ti}, stall samples {ts}")
    print(f"{'inst%':>6} {'stall%':>6} {'sass':>5}  location")
    for loc, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
        text = ""
        if loc:
            f = src_cache.setdefault(loc[0], None)
            if f is None:
                import glob
                cand = glob.glob(f"**/{loc[0]}", recursive=True)
                src_cache[loc[0]] = f = open(cand[0]).read().splitlines() if cand else []
            text = f[loc[1] - 1].strip()[:90] if 0 < loc[1] <= len(f) else ""
        print(f"{a[0] / ti * 100:6.2f} {a[1] / max(ts, 1) * 100:6.2f} {a[2]:5d}  {loc[0] if loc else '?'}:{loc[1] if loc else 0}  {text}")


if __name__ == "__main__":
    main()
