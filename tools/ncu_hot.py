"""Per-opcode instruction counts, stall reasons and the most-sampled SASS lines of one kernel in an .ncu-rep.

  python tools/ncu_hot.py gpurun_out/prof_x.ncu-rep [units]     units: divide the counts by this (e.g. warp-words)
"""
import collections, csv, re, subprocess, sys


def main():
    rep = sys.argv[1]
    units = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = rows[1]
    ia, ie, isamp = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
    body = [r for r in rows[2:] if len(r) > isamp and r[ie].isdigit()]
    ops = collections.Counter()
    for r in body:
        s = re.sub(r"^@!?U?P\d\s+", "", r[ia].strip())
        ops[s.split()[0].split(".")[0]] += int(r[ie])
    tot = sum(ops.values())
    print(f"instructions {tot} = {tot / units:.2f} per unit")
    print(" ".join(f"{k} {v / units:.2f}" for k, v in ops.most_common(24)))
    cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    st = collections.Counter()
    for r in body:
        for i in cols:
            if r[i].isdigit():
                st[hdr[i]] += int(r[i])
    n = sum(st.values())
    print(" ".join(f"{k[6:]} {100 * v / n:.1f}%" for k, v in st.most_common(10)))
    samples = sum(int(r[isamp]) for r in body)
    top = sorted(((int(r[isamp]), i, r) for i, r in enumerate(body)), reverse=True)[:25]
    for s, i, r in top:
        print(f"{100 * s / samples:5.2f}% {i:6d} {r[ie]:>9s} {r[ia].strip()[:80]}")


main()
