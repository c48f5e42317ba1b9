"""Summary of one-kernel ncu reports: python tools/prof_summary.py report.ncu-rep  (totals, stall reasons, opcode mix, hottest source lines)"""
import collections
import csv
import subprocess
import sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, vals = rows[0], rows[2]
want = ["gpu__time_duration.sum", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread"]
for h, v in zip(hdr, vals):
    if h in want:
        print(f"{h:70s} {v}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
hi = next(i for i, r in enumerate(rows) if "Source" in r)
hdr, data = rows[hi], rows[hi + 1:]
cols = [i for i, c in enumerate(hdr) if c.startswith("stall_") and "Not Issued" not in c]
tot = {hdr[i]: sum(float(r[i] or 0) for r in data) for i in cols}
s = sum(tot.values())
print("stalls:", ", ".join(f"{k[6:]} {v / s * 100:.1f}%" for k, v in sorted(tot.items(), key=lambda kv: -kv[1])[:8]))
ie, si = hdr.index("Instructions Executed"), hdr.index("Source")
ops = collections.Counter()
for r in data:
    t = r[si].strip().split()
    if t:
        ops[(t[1] if t[0].startswith("@") else t[0]).split(".")[0]] += float(r[ie] or 0)
t = sum(ops.values())
print("opcodes:", ", ".join(f"{k} {v / t * 100:.1f}%" for k, v in ops.most_common(16)))
