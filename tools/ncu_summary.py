#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU): one row per captured launch with the counters the roofline
discussion uses.  usage: tools/ncu_summary.py gpurun_out/x.ncu-rep"""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
want = [("Kernel Name", "kernel"), ("gpu__time_duration.sum", "us"), ("smsp__thread_inst_executed_per_inst_executed.ratio", "lanes/inst"),
        ("smsp__inst_executed.sum", "warp_inst"), ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"), ("launch__registers_per_thread", "regs"),
        ("dram__bytes_read.sum", "dramR"), ("dram__bytes_write.sum", "dramW"), ("lts__t_sector_hit_rate.pct", "L2hit%"),
        ("l1tex__t_sector_hit_rate.pct", "L1hit%"), ("lts__t_sectors.sum", "L2sectors"), ("l1tex__t_sectors.sum", "L1sectors"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2%"), ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "L1%"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM%")]
cols = [(hdr.index(k), n) for k, n in want if k in hdr]
print(" | ".join(n + ("(" + units[i] + ")" if units[i] else "") for i, n in cols))
for r in data:
    print(" | ".join(r[i][:28] for i, n in cols))
