#!/usr/bin/env python
"""profiles/traffic.json from an .ncu-rep (read here, no GPU): per-launch counters of the first k_trace<false> launch that
traced a full 1080p queue -- what bench.py's roofline object quotes beside its own CUDA-event numbers.
usage: tools/ncu_traffic.py gpurun_out/x.ncu-rep [out.json]"""
import csv, io, json, subprocess, sys
rep = sys.argv[1]; out = sys.argv[2] if len(sys.argv) > 2 else "profiles/traffic.json"
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, data = rows[0], rows[2:]
col = lambda r, k: float(r[hdr.index(k)].replace(",", "")) if k in hdr and r[hdr.index(k)] not in ("", "n/a") else None
res = {}
for r in data:
    name = r[hdr.index("Kernel Name")]
    key = "k_trace<false>" if "k_trace<0>" in name or "k_trace<false>" in name else "k_trace_pair" if "k_trace_pair" in name else "k_tail" if "k_tail" in name else None
    if key is None or key in res:
        continue
    res[key] = {
        "duration_us": col(r, "gpu__time_duration.sum"),
        "dram_bytes_per_launch": (col(r, "dram__bytes_read.sum") or 0) + (col(r, "dram__bytes_write.sum") or 0),
        "lts_bytes_per_launch": (col(r, "lts__t_sectors.sum") or 0) * 32,
        "l1_bytes_per_launch": (col(r, "l1tex__t_sectors.sum") or 0) * 32,
        "issue_active_pct": col(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
        "lanes_per_inst": col(r, "smsp__thread_inst_executed_per_inst_executed.ratio"),
        "l1_hit_pct": col(r, "l1tex__t_sector_hit_rate.pct"), "l2_hit_pct": col(r, "lts__t_sector_hit_rate.pct"),
        "dram_pct_of_peak": col(r, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
        "l2_pct_of_peak": col(r, "lts__throughput.avg.pct_of_peak_sustained_elapsed"),
        "registers": col(r, "launch__registers_per_thread"), "occupancy_pct": col(r, "sm__warps_active.avg.pct_of_peak_sustained_active"),
        "source": rep.split("/")[-1] + " (ncu --set full, cold cache, units as ncu prints them: byte counters may be in Mbyte -- see the unit row)"}
units = dict(zip(hdr, rows[1]))
for k in res.values():
    for name, metric in (("dram_bytes_per_launch", "dram__bytes_read.sum"),):
        u = units.get(metric, "")
        mult = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1}.get(u, 1)
        k[name] *= mult
json.dump(res, open(out, "w"), indent=1)
print(json.dumps(res, indent=1))
