#!/usr/bin/env python
"""Write a small JSON summary of one kernel instance of an .ncu-rep (for profiles/)."""
import csv, json, subprocess, sys
rep, out = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, r = rows[0], rows[1], rows[2]
d = dict(zip(hdr, r)); u = dict(zip(hdr, units))
def val(name, scale={"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1, "ms": 1e-3, "us": 1e-6, "ns": 1e-9, "s": 1}):
    return float(d[name].replace(",", "")) * scale.get(u[name], 1)
keep = ["gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "smsp__inst_executed.sum"]
s = {"kernel": d["Kernel Name"].split("(")[0], "grid": d["Grid Size"], "block": d["Block Size"],
     "duration_s": val("gpu__time_duration.sum"),
     "dram_bytes_read": val("dram__bytes_read.sum"), "dram_bytes_write": val("dram__bytes_write.sum"),
     "metrics": {k: float(d[k].replace(",", "")) for k in keep if k in d},
     "source": "ncu --set full --clock-control none (profiled run: cold caches, serialised; use shares and byte counts, not the time)"}
s["dram_bytes_per_launch"] = s["dram_bytes_read"] + s["dram_bytes_write"]
json.dump(s, open(out, "w"), indent=1)
print(json.dumps(s)[:400])
