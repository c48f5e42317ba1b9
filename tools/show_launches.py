"""Print the last launches of an ncu --csv launch list (gpu__time_duration.sum): python tools/show_launches.py file.csv [n]"""
import csv
import sys

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
n = int(sys.argv[2]) if len(sys.argv) > 2 else 6
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
for r in rows[-n:]:
    v = float(r[vi].replace(",", ""))
    if r[ui] == "ns":
        v /= 1000
    print(f"  {r[ki][:70]:70s} {v:9.1f} us")
