#!/usr/bin/env python
"""development aid: digest `ncu --page raw/source --csv` exports of one kernel.
usage: ncu_summary.py raw.csv src.csv [top_n]"""
import collections
import csv
import sys

raw, src = sys.argv[1], sys.argv[2]
top_n = int(sys.argv[3]) if len(sys.argv) > 3 else 25
rows = list(csv.reader(open(raw)))
hdr = rows[0]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__issue_active.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "smsp__inst_executed.sum", "l1tex__t_sectors.sum",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers"]
for i, h in enumerate(hdr):
    if h in want:
        print(f"{h:70s} {rows[2][i]} {rows[1][i]}")
for i, h in enumerate(hdr):
    if "issue_stalled" in h and h.endswith("per_issue_active.ratio") and float(rows[2][i] or 0) > 0.1:
        print(f"{h.split('issue_stalled_')[1].split('_per_issue')[0]:30s} {float(rows[2][i]):.2f}")

rows = list(csv.reader(open(src)))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[hi]
ix = {h: i for i, h in enumerate(hdr)}
out = []
for r in rows[hi + 1:]:
    if r and r[0] == "Kernel Name":
        break
    out.append(r)
base = int(out[0][ix["Address"]], 16)
tot_s = sum(int(r[ix["# Samples"]] or 0) for r in out)
tot_i = sum(int(r[ix["Instructions Executed"]] or 0) for r in out)
print("static instr", len(out), "samples", tot_s, "warp instr executed", tot_i)
w = collections.OrderedDict()
for r in out:
    a = int(r[ix["Address"]], 16) - base
    x = w.setdefault(a // 0x400, [0, 0])
    x[0] += int(r[ix["Instructions Executed"]] or 0)
    x[1] += int(r[ix["# Samples"]] or 0)
print("per 0x400 window: offset, instr executed (M), samples")
for k, (n, sm) in w.items():
    if n > tot_i / 400 or sm > tot_s / 200:
        print(f"  {k * 0x400:#7x} {n / 1e6:8.2f} {sm:6d}")
print("top stall sites")
for r in sorted(out, key=lambda r: -int(r[ix["# Samples"]] or 0))[:top_n]:
    a = int(r[ix["Address"]], 16) - base
    print(f"  {a:#7x} {r[ix['# Samples']]:>6s} exec {r[ix['Instructions Executed']]:>8s} lsb {r[ix['stall_long_sb']]:>5s} ssb {r[ix['stall_short_sb']]:>5s} "
          f"wait {r[ix['stall_wait']]:>4s} br {r[ix['stall_branch_resolving']]:>4s} | {r[ix['Source']][:64]}")
