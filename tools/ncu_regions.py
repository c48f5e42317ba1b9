"""Splits a kernel's SASS (from an .ncu-rep source page) into runs of instructions with the same execution count and
prints, per run, instructions per unit and the share of stall samples with the top stall reasons.
  python tools/ncu_regions.py rep.ncu-rep units [min_share_pct]"""
import collections, csv, subprocess, sys
rep, units = sys.argv[1], float(sys.argv[2])
min_share = float(sys.argv[3]) if len(sys.argv) > 3 else 0.7
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]
ia, ie, isamp = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
body = [r for r in rows[2:] if len(r) > isamp and r[ie].isdigit()]
tot = sum(int(r[ie]) for r in body); tots = sum(int(r[isamp]) for r in body)
blocks, cur = [], None
for i, r in enumerate(body):
    e, s = int(r[ie]), int(r[isamp])
    op = r[ia].strip().split()
    op = op[1] if op[0].startswith("@") else op[0]
    if cur and abs(e - cur["e"]) <= 0.03 * max(cur["e"], 1):
        pass
    else:
        if cur: blocks.append(cur)
        cur = {"i": i, "e": e, "n": 0, "sum": 0, "s": 0, "ops": collections.Counter(), "st": collections.Counter()}
    cur["n"] += 1; cur["sum"] += e; cur["s"] += s; cur["ops"][op.split(".")[0]] += 1
    for c in stall_cols:
        if r[c].isdigit(): cur["st"][hdr[c][6:]] += int(r[c])
blocks.append(cur)
print(f"total {tot / units:.1f} instr/unit, {tots} samples")
for b in blocks:
    if 100 * b["s"] / tots >= min_share:
        st = " ".join(f"{k}:{100 * v / max(1, sum(b['st'].values())):.0f}%" for k, v in b["st"].most_common(3))
        print(f"{b['i']:6d} n={b['n']:4d} exec/unit={b['e'] / units:7.3f} instr/unit={b['sum'] / units:6.1f} samples={100 * b['s'] / tots:5.1f}%  "
              f"{' '.join(f'{k}{v}' for k, v in b['ops'].most_common(5))} | {st}")
