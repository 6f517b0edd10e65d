#!/usr/bin/env python
"""Aggregates `ncu --page source --csv` (SASS view): executed warp-instructions and stall samples per opcode, plus the
hottest individual instructions."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
si, ii, smp = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
ops, samp = collections.Counter(), collections.Counter()
tot = 0
body = []
for r in rows[2:]:
    if r and r[0] == "Kernel Name": break  # (some exports list the kernel a second time: first section only)
    if len(r) <= smp: continue
    try: n, s = int(r[ii]), int(r[smp])
    except ValueError: continue
    op = r[si].split()[0] if not r[si].strip().startswith('@') else r[si].split()[1]
    op = op.split('.')[0]
    ops[op] += n; samp[op] += s; tot += n
    body.append((s, n, r[si].strip()))
print("total warp instr", tot, "samples", sum(samp.values()))
for op, n in ops.most_common(25):
    print(f"  {op:10s} {n:12d} {100*n/tot:5.1f}%  samples {samp[op]}")
top = int(sys.argv[2]) if len(sys.argv) > 2 else 15
print("hottest by samples:")
for s, n, t in sorted(body, reverse=True)[:top]:
    print(f"  {s:6d} {n:10d}  {t[:100]}")
