#!/usr/bin/env python
"""Top SASS instructions / source lines of one kernel from an .ncu-rep (source page).
usage: ncu_hot.py <rep> <kernel-name> [n]"""
import csv, subprocess, sys, collections
rep, kern = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kern, "--print-source", "sass"],
                     capture_output=True, text=True).stdout.splitlines()
# first kernel instance only
start = [i for i, l in enumerate(raw) if l.startswith('"Kernel Name"')]
blk = raw[start[0] + 1:(start[1] if len(start) > 1 else len(raw))]
rows = list(csv.reader(blk))
hdr = rows[0]
ia, isrc, iex, ismp = hdr.index("Address"), hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
tot_ex = sum(float(r[iex] or 0) for r in rows[1:])
tot_s = sum(float(r[ismp] or 0) for r in rows[1:])
print("total warp-inst %.0f  samples %.0f  sass lines %d" % (tot_ex, tot_s, len(rows) - 1))
ops = collections.Counter()
for r in rows[1:]:
    op = r[isrc].split()[0] if not r[isrc].startswith("@") else r[isrc].split()[1]
    ops[op.split(".")[0]] += float(r[iex] or 0)
print("by opcode:", ", ".join("%s %.1f%%" % (k, 100 * v / tot_ex) for k, v in ops.most_common(14)))
print("-- top by samples")
for r in sorted(rows[1:], key=lambda r: -float(r[ismp] or 0))[:n]:
    print("%6.2f%% smp %6.2f%% exe  %s  %s" % (100 * float(r[ismp] or 0) / max(tot_s, 1), 100 * float(r[iex] or 0) / tot_ex, r[ia][-5:], r[isrc][:90]))
