#!/usr/bin/env python
"""ncu_summary.py <report.ncu-rep> — the handful of raw metrics the round notes quote, as markdown."""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
want = ['Kernel Name', 'gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__occupancy_limit_registers', 'launch__grid_size',
        'launch__block_size', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct',
        'lts__t_bytes.sum', 'smsp__inst_executed.sum', 'sass__inst_executed_local_loads', 'sass__inst_executed_local_stores',
        'sass__inst_executed_global_loads']
stalls = [h for h in hdr if h.startswith('smsp__average_warps_issue_stalled_') and h.endswith('_per_issue_active.ratio')]
print("| metric | unit | " + " | ".join("launch %d" % (i + 1) for i in range(len(data))) + " |")
print("|---|---|" + "---|" * len(data))
for w in want + stalls:
    if w in hdr:
        i = hdr.index(w)
        vals = [r[i][:60] for r in data]
        if w in stalls and all(float(v or 0) < 0.05 for v in vals):
            continue
        print("| %s | %s | %s |" % (w.replace('smsp__average_warps_issue_stalled_', 'stall_').replace('_per_issue_active.ratio', ''), units[i], " | ".join(vals)))
