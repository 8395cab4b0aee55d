#!/usr/bin/env python
"""Condensed SASS listing from `ncu --page source --csv --print-source sass`:
per instruction the executed count (relative to the first instruction) and stall samples."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
start = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[start]
ix = {h: i for i, h in enumerate(hdr)}
data = []
for r in rows[start + 1:]:
    if len(r) < len(hdr):
        break          # next kernel instance
    data.append(r)
base = float(data[0][ix["Instructions Executed"]]) or 1.0
tot_s = sum(int(r[ix["# Samples"]] or 0) for r in data)
tot_i = sum(float(r[ix["Instructions Executed"]] or 0) for r in data)
print("instructions: %d, executed/warp-launch-unit: %.1f, samples %d" % (len(data), tot_i / base, tot_s))
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.0
for n, r in enumerate(data):
    s = int(r[ix["# Samples"]] or 0)
    e = float(r[ix["Instructions Executed"]] or 0) / base
    if 100.0 * s / max(1, tot_s) >= thr:
        print("%4d %6.2f%% x%7.2f  %s" % (n, 100.0 * s / max(1, tot_s), e, r[ix["Source"]].strip()))
