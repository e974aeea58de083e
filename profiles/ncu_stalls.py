#!/usr/bin/env python3
"""Aggregate warp-stall samples of one kernel from `ncu --page source --csv --kernel-name K`."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
starts = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
inst = int(sys.argv[2]) if len(sys.argv) > 2 else 0
hdr = rows[starts[inst]]
blk = rows[starts[inst] + 1: (starts[inst + 1] - 1 if len(starts) > inst + 1 else len(rows))]
cols = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
tot = {hdr[i]: 0 for i in cols}
for r in blk:
    for i in cols:
        if i < len(r) and r[i].isdigit():
            tot[hdr[i]] += int(r[i])
s = sum(tot.values())
for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
    if v:
        print(f"{k:28s} {v:8d} {100*v/s:5.1f}%")
