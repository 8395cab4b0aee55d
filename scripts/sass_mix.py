#!/usr/bin/env python
"""Executed instruction mix of one kernel instance of an .ncu-rep, by SASS opcode.

  python scripts/sass_mix.py REPORT.ncu-rep KERNEL_SUBSTRING [UNITS]

Reads `ncu --page source --csv --print-source sass` (the report must have been captured with
--import-source on / -lineinfo), sums "Thread Instructions Executed" and "Instructions Executed" per
opcode (modifiers stripped: LDS.128 -> LDS) and prints them per UNITS (e.g. reads voted) when given."""
import collections
import csv
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
units = float(sys.argv[3]) if len(sys.argv) > 3 else None
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--kernel-name",
                      "regex:" + kern], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
start = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[start]
ix = {h: i for i, h in enumerate(hdr)}
warp, thread, static = collections.Counter(), collections.Counter(), collections.Counter()
for r in rows[start + 1:]:
    if len(r) < len(hdr):
        break
    src = r[ix["Source"]].strip()
    toks = src.split()
    if toks and toks[0].startswith("@"):
        toks = toks[1:]
    if not toks:
        continue
    op = toks[0].split(".")[0]
    warp[op] += float(r[ix["Instructions Executed"]] or 0)
    thread[op] += float(r[ix["Thread Instructions Executed"]] or 0)
    static[op] += 1
tw, tt = sum(warp.values()), sum(thread.values())
print("kernel %s: %d SASS instructions, %.4g warp-instructions executed, %.4g thread-instructions, %.1f active threads per warp-instruction"
      % (kern, sum(static.values()), tw, tt, tt / max(tw, 1)))
if units:
    print("per unit (%g units): %.1f warp-instructions x32 = %.1f thread slots, %.1f thread-instructions" % (units, tw / units, 32 * tw / units, tt / units))
print("%-10s %8s %14s %8s %s" % ("opcode", "static", "warp-instr", "share", "per unit (warp-instr x 32)" if units else ""))
for op, v in warp.most_common():
    print("%-10s %8d %14.4g %7.2f%% %s" % (op, static[op], v, 100 * v / tw, "%.2f" % (32 * v / units) if units else ""))
