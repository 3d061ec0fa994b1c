#!/usr/bin/env python
"""Summarise an .ncu-rep here (no GPU needed): headline raw metrics + stall samples per source line.
usage: python tools/ncu_summary.py gpurun_out/x.ncu-rep [top_n]"""
import csv, io, subprocess, sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "lts__t_sector_hit_rate.pct", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "launch__grid_size", "launch__block_size", "smsp__warps_eligible.avg.per_cycle_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "sm__cycles_elapsed.max", "launch__shared_mem_per_block_dynamic", "sm__inst_executed_pipe_fp64.sum",
        "smsp__inst_executed_pipe_fp64.sum", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.sum", "smsp__thread_inst_executed_per_inst_executed.ratio"]
for r in rows[2:]:
    d = dict(zip(rows[0], r))
    print("kernel:", d.get("Kernel Name", "?")[:100])
    for h, u in zip(rows[0], rows[1]):
        if h in want:
            print("  %-68s %-14s %s" % (h, u, d[h]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
cur, hdr, agg = None, None, []
for r in csv.reader(io.StringIO(src)):
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1]
    elif r[0] == "Line No":
        hdr = r
    elif hdr and r[0].isdigit() and len(r) > 4 and r[2] == "-":
        d = dict(zip(hdr[4:], r[4:]))
        agg.append((cur.split("/")[-1], int(r[0]), r[1], int(d["# Samples"] or 0), int(d["Instructions Executed"] or 0)))
ts, ti = sum(a[3] for a in agg) or 1, sum(a[4] for a in agg) or 1
print("total samples %d, warp instructions %d" % (ts, ti))
for a in sorted(agg, key=lambda a: -a[3])[:top]:
    print("%-20s %4d  samples %5.1f%%  inst %5.1f%%  %s" % (a[0][:20], a[1], 100 * a[3] / ts, 100 * a[4] / ti, a[2][:100]))
# instruction / sample share by source region (line ranges given as name:lo-hi,... in argv[3])
if len(sys.argv) > 3:
    print("regions:")
    for spec in sys.argv[3].split(","):
        name, rng = spec.split(":")
        lo, hi = map(int, rng.split("-"))
        sel = [a for a in agg if a[0].startswith("place_kernel") and lo <= a[1] <= hi]
        print("  %-12s lines %4d-%4d  samples %5.1f%%  inst %5.1f%%" % (name, lo, hi, 100 * sum(a[3] for a in sel) / ts, 100 * sum(a[4] for a in sel) / ti))
    other = [a for a in agg if not a[0].startswith("place_kernel")]
    print("  %-12s                  samples %5.1f%%  inst %5.1f%%" % ("other files", 100 * sum(a[3] for a in other) / ts, 100 * sum(a[4] for a in other) / ti))
