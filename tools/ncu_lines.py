#!/usr/bin/env python
"""Per-source-line instruction counts of one captured launch:
   ncu -i X.ncu-rep --page source --csv --print-source sass,cuda > src.csv; python tools/ncu_lines.py src.csv UNITS
UNITS = work items of the launch (series), to print instructions per item."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
units = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
agg, stall = collections.Counter(), collections.Counter()
text = {}
cur = hdr = None
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur, hdr = r[1].split("/")[-1], None
        continue
    if r[0] == "Line No":
        hdr = r
        i_n, i_s = hdr.index("Instructions Executed"), hdr.index("Warp Stall Sampling (All Samples)")
        continue
    if hdr is None or r[2] != "-":
        continue
    try:
        ln, n, s = int(r[0]), int(r[i_n] or 0), int(r[i_s] or 0)
    except ValueError:
        continue
    agg[(cur, ln)] += n
    stall[(cur, ln)] += s
    text[(cur, ln)] = r[1].strip()[:70]
tot, ts = sum(agg.values()), sum(stall.values())
print(f"total warp instructions {tot}, per item {tot / units:.1f}; stall samples {ts}")
for key, n in sorted(agg.items(), key=lambda x: -x[1])[:int(sys.argv[3]) if len(sys.argv) > 3 else 45]:
    print(f"{key[0]}:{key[1]:<5d} {n / units:8.1f} instr/item {100 * stall[key] / max(ts, 1):5.1f}% stalls  {text[key]}")
