#!/usr/bin/env python
"""profiles/tools/summarise_ncu.py <capture.ncu-rep> <out-prefix>: the raw page of an `ncu --set full` capture as CSV
(<out-prefix>_raw.csv, as in round 1) plus the 25 source lines with the most stall samples (<out-prefix>_hot_lines.txt)."""
import csv
import io
import subprocess
import sys

rep, out = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
open(out + "_raw.csv", "w").write(raw)
rows = list(csv.reader(io.StringIO(raw)))
if len(rows) >= 3:
    head, vals = rows[0], rows[2]
    keep = ("Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
            "sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
            "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit_shared_mem",
            "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "lts__t_sector_hit_rate.pct")
    for k in keep:
        for i, h in enumerate(head):
            if h == k:
                print("%-80s %s %s" % (k, vals[i], rows[1][i]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
lines = list(csv.reader(io.StringIO(src)))
hdr = None
for i, r in enumerate(lines):
    if "Source" in r and any("Sampling" in c for c in r):
        hdr = i
        break
if hdr is not None:
    h = lines[hdr]
    si = h.index("Source")
    samp = [i for i, c in enumerate(h) if c.startswith("# Samples") or c == "Warp Stall Sampling (All Samples)" or c == "Warp Stall Sampling (All Cycles)"]
    col = samp[0] if samp else None
    stall_cols = [i for i, c in enumerate(h) if c.startswith("stall_")]
    body = []
    for r in lines[hdr + 1:]:
        try:
            n = float(r[col])
        except Exception:
            continue
        top = sorted(((float(r[i] or 0), h[i]) for i in stall_cols), reverse=True)[:3]
        body.append((n, r[si].strip()[:150], ", ".join("%s=%g" % (nm, v) for v, nm in top if v > 0)))
    total = sum(b[0] for b in body) or 1.0
    body.sort(reverse=True)
    with open(out + "_hot_lines.txt", "w") as f:
        f.write("# %s: source lines by warp-stall samples (share of all samples | line | top stall reasons)\n" % rep)
        for n, s, st in body[:25]:
            f.write("%5.1f%% | %s | %s\n" % (100.0 * n / total, s, st))
    print(open(out + "_hot_lines.txt").read())
else:
    print("no source page with sampling columns; header rows:", [r[:6] for r in lines[:3]])
