#!/usr/bin/env python
"""Markdown table of every kernel instance of an .ncu-rep (one step): time, DRAM bytes, DRAM GB/s against the
measured peak, issue-slot use.  python scripts/ncu_step_table.py REPORT.ncu-rep [peak GB/s]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
peak = float(sys.argv[2]) if len(sys.argv) > 2 else 6552.6
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1, "ms": 1e-3, "us": 1e-6, "ns": 1e-9, "s": 1, "msecond": 1e-3, "usecond": 1e-6, "nsecond": 1e-9, "second": 1}


def val(d, name):
    return float(d[name].replace(",", "")) * scale.get(units[hdr.index(name)], 1)


print("| kernel | time us | DRAM read MB | DRAM write MB | DRAM GB/s | of %.0f GB/s | issue slots busy %% |" % peak)
print("|---|---|---|---|---|---|---|")
tot_t = tot_r = tot_w = 0.0
for r in rows[2:]:
    d = dict(zip(hdr, r))
    t, rd, wr = val(d, "gpu__time_duration.sum"), val(d, "dram__bytes_read.sum"), val(d, "dram__bytes_write.sum")
    tot_t += t; tot_r += rd; tot_w += wr
    name = d["Kernel Name"].split("(")[0].replace("void ", "").replace("ccgpu::", "")
    print("| `%s` | %.1f | %.0f | %.0f | %.0f | %.0f %% | %.0f |" % (name, t * 1e6, rd / 1e6, wr / 1e6, (rd + wr) / t / 1e9,
          100 * (rd + wr) / t / 1e9 / peak, float(d["smsp__issue_active.avg.pct_of_peak_sustained_active"])))
print("\nSum of the kernel times: %.3f ms; DRAM read %.0f MB, written %.0f MB." % (tot_t * 1e3, tot_r / 1e6, tot_w / 1e6))
