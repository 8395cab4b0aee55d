#!/usr/bin/env python
"""Executed warp-instructions per CUDA source line (ncu --page source --print-source cuda,sass)."""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
div = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
rows = list(csv.reader(out.splitlines()))
cur_file = None; hdr = None; per = {}
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur_file = r[1].split("/")[-1]; continue
    if r[0] == "Line No": hdr = r; continue
    if hdr and len(r) == len(hdr) and r[0].isdigit():
        ix = hdr.index("Instructions Executed")
        try: e = float(r[ix])
        except ValueError: continue
        per[(cur_file, int(r[0]), r[1].strip()[:90])] = per.get((cur_file, int(r[0]), r[1].strip()[:90]), 0) + e
tot = sum(per.values())
print("total %.1f per unit" % (tot / div))
for (f, ln, src), e in sorted(per.items(), key=lambda x: -x[1])[:45]:
    print("%6.1f %5.1f%%  %s:%d  %s" % (e / div, 100 * e / tot, f, ln, src))
