#!/usr/bin/env python
"""Key metrics of every kernel instance in an .ncu-rep (reads `ncu --page raw --csv`)."""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'smsp__warps_eligible.avg.per_cycle_active', 'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'launch__occupancy_limit_registers',
        'sm__maximum_warps_per_active_cycle_pct', 'launch__waves_per_multiprocessor']
for r in rows[2:]:
    d = dict(zip(hdr, r))
    print(d.get('Kernel Name', '')[:60], d.get('Grid Size'), d.get('Block Size'))
    for w in want:
        if w in d:
            print("   %-72s %s %s" % (w, d[w], units[hdr.index(w)]))
    st = [(float(d[h]), h) for h in hdr if 'warp_issue_stalled' in h and h.endswith('_per_warp_active.pct') and d[h]]
    for v, h in sorted(st, reverse=True)[:6]:
        print("   stall %-66s %.1f" % (h.replace('smsp__warp_issue_stalled_', '').replace('_per_warp_active.pct', ''), v))
