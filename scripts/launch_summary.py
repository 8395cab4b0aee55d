#!/usr/bin/env python
"""Per-kernel totals of an ncu launch-list CSV (gpu__time_duration.sum)."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hdr = None; data = []
for r in rows:
    if r and r[0] == 'ID': hdr = r; continue
    if hdr and len(r) == len(hdr): data.append(dict(zip(hdr, r)))
steps = float(sys.argv[2]) if len(sys.argv) > 2 else 5.0
agg = collections.OrderedDict()
for d in data:
    k = d['Kernel Name'].split('(')[0]
    a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += float(d['Metric Value'].replace(',', ''))
tot = sum(a[1] for a in agg.values())
print("%d launches, %.3f ms per step (sum of kernel times / %g steps)" % (len(data), tot / steps / 1e6, steps))
for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
    print("%-48s n/step=%5.1f  ms/step=%7.3f  share=%5.1f%%  avg_us=%9.1f" % (k[:48], a[0] / steps, a[1] / steps / 1e6, 100 * a[1] / tot, a[1] / a[0] / 1e3))
