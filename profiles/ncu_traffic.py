#!/usr/bin/env python
"""DRAM bytes per launch of every hot-path kernel from `ncu -i X.ncu-rep --page raw --csv` (one whole step captured
with --set full) -> profiles/ncu_traffic.json, which bench.py reports as roofline.traffic.
Usage: python profiles/ncu_traffic.py raw.csv ["source note"] > profiles/ncu_traffic.json"""
import csv
import json
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
names, units = rows[hdr], rows[hdr + 1]
col = {n: i for i, n in enumerate(names)}
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
acc = {}
seen_dup = False
for r in rows[hdr + 2:]:
    if len(r) < len(names):
        continue
    full = r[col["Kernel Name"]]
    m = re.search(r"(\w+_kernel)(<(\d+)[^>]*>)?", full)
    if not m:
        continue
    key = m.group(1)
    if key == "dup_scan_kernel":
        seen_dup = True
    if key == "preprocess_fwd_kernel":
        seen_dup = False  # a new step
    if key == "onesweep_kernel":  # the passes before the duplication kernel order the Gaussians by depth, the ones
        key += "_instances" if seen_dup else "_depth_order"  # after it sort the instance words on the tile bits
        if not seen_dup and float(r[col["gpu__time_duration.sum"]].replace(",", "")) < 5.0 and \
                units[col["gpu__time_duration.sum"]] in ("us", "usecond"):
            key += "_skipped"  # the pass above the key width returns at once
    total = 0.0
    for c in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        total += float(r[col[c]].replace(",", "")) * scale[units[col[c]]]
    a = acc.setdefault(key, [0, 0.0])
    a[0] += 1
    a[1] += total
out = {k: {"launches": n, "dram_bytes_per_launch": b / n} for k, (n, b) in acc.items()}
if len(sys.argv) > 2:
    out["_source"] = sys.argv[2]
print(json.dumps(out, indent=1))
