"""Turns an .ncu-rep into the tracked summary kept under profiles/: one metric,unit,value table per captured
kernel (ncu --page raw), and optionally the launch list of a bench run into per-kernel totals.

  python tools/profile_summary.py rep gpurun_out/prof_x.ncu-rep profiles/r1b_x.summary.csv
  python tools/profile_summary.py launches gpurun_out/r1b_launches.csv profiles/r1b_launches.summary.txt [first_n]
"""
import collections, csv, subprocess, sys


def rep(src, dst):
    out = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(out.splitlines()))
    h, unit = r[0], r[1]
    with open(dst, "w", newline="") as f:
        w = csv.writer(f)
        for row in r[2:]:
            w.writerow(["metric", "unit", "value"])
            for k, u, v in zip(h, unit, row):
                if k in ("ID", "Process ID", "Process Name", "Host Name", "Context", "Stream", "Device", "CC"):
                    continue
                w.writerow([k, u, v])
            w.writerow([])


def launches(src, dst, first_n=None):
    rows = list(csv.reader(open(src)))
    hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
    h = rows[hi]
    kn, mv = h.index("Kernel Name"), h.index("Metric Value")
    body = [r for r in rows[hi + 1:] if len(r) > mv]
    if first_n:
        body = body[: int(first_n)]
    agg = collections.OrderedDict()
    for r in body:
        agg.setdefault(r[kn].split("(")[0], []).append(float(r[mv].replace(",", "")))
    tot = sum(sum(v) for v in agg.values())
    with open(dst, "w") as f:
        f.write(f"# {src}: {len(body)} launches, gpu__time_duration.sum (ns); cold-cache, serialised by ncu\n")
        for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
            f.write(f"{k:62s} n={len(v):4d} avg={sum(v) / len(v) / 1e3:10.2f} us share={100 * sum(v) / tot:5.1f}%\n")


if __name__ == "__main__":
    {"rep": rep, "launches": launches}[sys.argv[1]](*sys.argv[2:])
