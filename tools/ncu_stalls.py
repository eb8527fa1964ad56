#!/usr/bin/env python
"""Summarise an ncu report's source page: top SASS instructions by stall samples, with the dominant stall reason.
Usage: tools/ncu_stalls.py <report.ncu-rep> [topN]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
kernel = None
hdr = None
data = []
for r in rows:
    if len(r) >= 2 and r[0] == "Kernel Name":
        if kernel is not None:
            break   # first kernel instance only
        kernel = r[1]
        continue
    if r and r[0] == "Address":
        hdr = r
        continue
    if hdr and len(r) == len(hdr):
        data.append(r)
print("kernel:", kernel)
ci = {h: i for i, h in enumerate(hdr)}
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
total = sum(int(r[ci["# Samples"]] or 0) for r in data)
agg = {h: sum(int(r[ci[h]] or 0) for r in data) for h in stall_cols}
print("total samples:", total)
print("stall mix:", ", ".join(f"{h[6:]}={100 * v / max(total, 1):.1f}%" for h, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
data.sort(key=lambda r: -int(r[ci["# Samples"]] or 0))
for r in data[:top]:
    s = int(r[ci["# Samples"]] or 0)
    dom = max(stall_cols, key=lambda h: int(r[ci[h]] or 0))
    print(f"{100 * s / max(total, 1):5.1f}%  {dom[6:]:<12} {r[ci['Source']].strip()[:100]}")
