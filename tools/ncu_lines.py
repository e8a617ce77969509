#!/usr/bin/env python
"""Per-source-line sample/instruction shares of one kernel from an .ncu-rep (cuda,sass view).
usage: ncu_lines.py <rep> <kernel> [n] [instance]"""
import csv, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 30
inst = int(sys.argv[4]) if len(sys.argv) > 4 else 0
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kern, "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout.splitlines()
rows = list(csv.reader(raw))
heads = [i for i, r in enumerate(rows) if r and r[0] == "Line No"]
# one "Line No" header per (kernel instance, file); group by instance = restart at "Function Name"? keep simple: pick by order
seen = {}
groups = []
for hi, h in enumerate(heads):
    end = heads[hi + 1] - 2 if hi + 1 < len(heads) else len(rows)
    groups.append((rows[h - 2][1] if h >= 2 else "", rows[h], rows[h + 1:end]))
files = [g[0] for g in groups]
per_inst = len(set(files)) or 1
sel = groups[inst * per_inst:(inst + 1) * per_inst]
tot_ex = tot_s = 0.0
lines = []
def f(x):
    try:
        return float(x)
    except ValueError:
        return 0.0
for fname, hdr, body in sel:
    iex, ismp = hdr.index("Instructions Executed"), hdr.index("# Samples")
    for r in body:
        if len(r) == len(hdr) and r[0] != "":
            lines.append((f(r[ismp]), f(r[iex]), fname.split("/")[-1], r[0], r[1]))
            tot_ex += f(r[iex]); tot_s += f(r[ismp])
print("instance %d: total warp-inst %.0f samples %.0f" % (inst, tot_ex, tot_s))
for s, e, fn, ln, src in sorted(lines, key=lambda x: -x[0])[:n]:
    print("%6.2f%% smp %6.2f%% exe  %s:%-4s %s" % (100 * s / max(tot_s, 1), 100 * e / max(tot_ex, 1), fn[:12], ln, src.strip()[:100]))
if len(sys.argv) > 5:  # line ranges "a-b,c-d" -> share of executed instructions
    for rng in sys.argv[5].split(","):
        lo, hi = (int(v) for v in rng.split("-"))
        e = sum(x[1] for x in lines if x[2].startswith("extract") and lo <= int(x[3]) <= hi)
        sm = sum(x[0] for x in lines if x[2].startswith("extract") and lo <= int(x[3]) <= hi)
        print("lines %d-%d: %.1f%% exe %.1f%% smp  (%.0f warp-inst)" % (lo, hi, 100 * e / tot_ex, 100 * sm / tot_s, e))
