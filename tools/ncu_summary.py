#!/usr/bin/env python3
"""Dev tool: text summary of one kernel of an .ncu-rep (raw page + per-opcode / per-line stall
sampling from the source page).  usage: tools/ncu_summary.py report.ncu-rep [top_lines]"""
import collections
import csv
import io
import re
import subprocess
import sys

rep = sys.argv[1]
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 16


def ncu(*args):
    return subprocess.run(["ncu", "-i", rep, *args], capture_output=True, text=True).stdout


raw = list(csv.reader(io.StringIO(ncu("--page", "raw", "--csv"))))
hdr, units, vals = raw[0], raw[1], raw[-1]
want = """gpu__time_duration.sum smsp__inst_executed.sum sm__issue_active.avg.pct_of_peak_sustained_elapsed
sm__warps_active.avg.pct_of_peak_sustained_active launch__registers_per_thread dram__bytes_read.sum
dram__bytes_write.sum l1tex__t_sector_hit_rate.pct lts__t_sector_hit_rate.pct
l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed
sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active
smsp__warps_eligible.avg.per_cycle_active sm__cycles_elapsed.avg""".split()
print("=== ", vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "")
for i, h in enumerate(hdr):
    if h in want:
        print(f"  {h:85s} {vals[i]} {units[i]}")
st = [(float(vals[i]), h) for i, h in enumerate(hdr)
      if "issue_stalled" in h and h.endswith("per_issue_active.ratio")]
print("  stalls per issue: " + ", ".join(
    f"{h.split('issue_stalled_')[1].replace('_per_issue_active.ratio', '')}={v:.2f}"
    for v, h in sorted(st, reverse=True)[:10]))

rows = list(csv.reader(io.StringIO(ncu("--page", "source", "--csv", "--print-source", "sass"))))
for i, r in enumerate(rows):
    if "Source" in r and "Instructions Executed" in r:
        h2, start = r, i + 1
        break
iS, iE = h2.index("Source"), h2.index("Instructions Executed")
iA = h2.index("Warp Stall Sampling (All Samples)")
cols = [i for i, c in enumerate(h2) if c.startswith("stall_") and "Not Issued" not in c]


def I(x):
    try:
        return int(x)
    except ValueError:
        return 0


data = rows[start:]
tot, totE = sum(I(r[iA]) for r in data), sum(I(r[iE]) for r in data)
maxE = max(I(r[iE]) for r in data)
blocks = collections.Counter(I(r[iE]) for r in data if I(r[iE]) > 0).most_common(1)[0][0]
pro = [r for r in data if 0 < I(r[iE]) <= blocks * 1.05]
print(f"  once-per-warp instructions: {sum(I(r[iE]) for r in pro) / totE * 100:.1f}% of executed, "
      f"{sum(I(r[iA]) for r in pro) / tot * 100:.1f}% of stall samples")
ops = collections.defaultdict(lambda: [0, 0])
for r in data:
    m = re.match(r"\s*(@!?U?P\w+\s+)?([A-Z0-9_]+)", r[iS])
    if m:
        ops[m.group(2)][0] += I(r[iA])
        ops[m.group(2)][1] += I(r[iE])
print("  opcode: samples% / executed%")
print("   " + "  ".join(f"{op} {a / tot * 100:.1f}/{e / totE * 100:.1f}"
                         for op, (a, e) in sorted(ops.items(), key=lambda kv: -kv[1][1])[:18]))
print("  top lines by stall samples:")
for k in sorted(range(len(data)), key=lambda k: -I(data[k][iA]))[:top_n]:
    r = data[k]
    s3 = sorted(((I(r[i]), h2[i][6:]) for i in cols), reverse=True)[:2]
    print(f"   {I(r[iA]) / tot * 100:4.1f}%  x{I(r[iE]) / 1e6:7.1f}M  {r[iS].strip()[:58]:58s} "
          + " ".join(f"{n}:{c}" for c, n in s3))
