#!/usr/bin/env python
"""Executed warp-instructions per opcode (and per pipe guess) from a SASS source page CSV."""
import csv, sys, collections, re
rows = list(csv.reader(open(sys.argv[1])))
start = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[start]; ix = {h: i for i, h in enumerate(hdr)}
ops = collections.Counter(); total = 0
for r in rows[start + 1:]:
    if len(r) < len(hdr): break
    src = r[ix["Source"]].strip()
    m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_.]+)", src)
    op = m.group(2).split(".")[0] if m else src
    e = float(r[ix["Instructions Executed"]] or 0)
    ops[op] += e; total += e
div = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
for op, e in ops.most_common(25):
    print("%-10s %12.1f  %5.1f%%" % (op, e / div, 100 * e / total))
print("total %.1f" % (total / div))
