#!/usr/bin/env python
"""Per-source-line hot spots from an .ncu-rep captured with --import-source on (needs -lineinfo).
usage: tools/ncu_lines.py rep [launch_index] [top_n]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; which = int(sys.argv[2]) if len(sys.argv) > 2 else 0; top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
launch = -1; files_seen = set(); cur_file = None; hdr = None; agg = {}; seen_funcs = []
i = 0
for r in rows:
    if not r: continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        if launch < 0 or cur_file in files_seen: launch += 1; files_seen = set()
        files_seen.add(cur_file)
        continue
    if r[0] == "Function Name":
        if len(seen_funcs) <= launch: seen_funcs.append(r[1][10:34])
        continue
    if r[0] == "Kernel Name":
        launch += 1; seen_funcs.append(r[1][:40]); continue
    if r[0] == "Line No": hdr = r; continue
    if launch != which and launch != -1: pass
    if hdr and r[0].isdigit():
        if launch in (-1, which) or True:
            key = (launch, cur_file, int(r[0]))
            def col(name):
                try: return float(r[hdr.index(name)])
                except Exception: return 0.0
            a = agg.setdefault(key, [r[1][:90], 0, 0, 0])
            a[1] += col("# Samples"); a[2] += col("Instructions Executed"); a[3] += col("Thread Instructions Executed")
launches = sorted(set(k[0] for k in agg))
print("launch ids in report:", launches, seen_funcs[:8])
sel = {k: v for k, v in agg.items() if k[0] == (launches[which] if which < len(launches) else launches[0])}
tot_i = sum(v[2] for v in sel.values()); tot_s = sum(v[1] for v in sel.values())
print("total warp inst %.3g, samples %d" % (tot_i, tot_s))
for k, v in sorted(sel.items(), key=lambda kv: -kv[1][2])[:top]:
    print("%5.1f%% inst %5.1f%% smp  lanes %4.1f  %s:%d  %s" % (100 * v[2] / tot_i, 100 * v[1] / max(1, tot_s), v[3] / max(1, v[2]), k[1], k[2], v[0].strip()))
