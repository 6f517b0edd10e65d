#!/usr/bin/env python
"""From an ncu launch list (csv, gpu__time_duration of every launch of `bench.py --steps 2 --warmup 3`): how many
hot-path kernels precede the last complete forward+backward step, i.e. the `-s` for a `-k regex:... -c 14` capture of
exactly that step.  Usage: skip_for_last_step.py launches.csv 'regex'  -> prints "SKIP COUNT"."""
import csv
import re
import sys

rx = re.compile(sys.argv[2])
lines = [l for l in open(sys.argv[1]) if l.startswith('"')]
r = csv.reader(lines)
hdr = next(r)
ki = hdr.index("Kernel Name")
names = [row[ki] for row in r]
hot = [n for n in names if rx.search(n)]
last_bwd = max(i for i, n in enumerate(hot) if "preprocess_bwd_kernel" in n)
first = max(i for i in range(last_bwd) if "preprocess_fwd_kernel" in hot[i])
print(first, last_bwd - first + 1)
