#!/usr/bin/env python
"""SASS listings of liborb_gpu.so for profiles/: one gzip'd listing per kernel (encodings stripped) and a
SUMMARY.txt with instruction counts and opcode histograms (the mnemonics that matter here: VABSDIFF4,
VIMNMX3.S16x2, IDP (dp2a/dp4a), POPC, LOP3, LDGSTS for cp.async).  usage: dump_sass.py [out_dir]"""
import collections, gzip, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "sass")
os.makedirs(out, exist_ok=True)
lib = os.path.join(ROOT, "c_orb_slam_b200", "liborb_gpu.so")
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
names = subprocess.run(["cu++filt"], input="\n".join(re.findall(r"Function : (\S+)", txt)), capture_output=True, text=True).stdout.split("\n")
blocks = re.split(r"\n\s*Function : ", txt)[1:]
summary = []
for blk, dem in zip(blocks, names):
    mangled, body = blk.split("\n", 1)
    short = re.sub(r"\(.*", "", dem).replace("orbgpu::", "").replace("void ", "")
    short = re.sub(r"[<>, ]+", "_", short).strip("_")
    targs = re.search(r"kernelI((?:L[ib]\d+E)+)E", mangled)  # template arguments, e.g. ILi2ELb1EE -> _2_1
    if targs:
        short += "_" + "_".join(re.findall(r"L[ib](\d+)E", targs.group(1)))
    lines = []
    ops = collections.Counter()
    for ln in body.splitlines():
        m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", ln)
        if m:
            ins = m.group(2).strip()
            lines.append("%s  %s" % (m.group(1), ins))
            op = ins.split()[1] if ins.startswith("@") else ins.split()[0]
            ops[op.split(".")[0] if not op.startswith(("VIMNMX", "IDP", "VABSDIFF", "LDGSTS", "POPC")) else op] += 1
    with gzip.open(os.path.join(out, short + ".sass.gz"), "wt") as f:
        f.write("// %s\n" % dem + "\n".join(lines) + "\n")
    summary.append("%-40s %6d instr | %s" % (short, len(lines), ", ".join("%s %d" % kv for kv in ops.most_common(14))))
open(os.path.join(out, "SUMMARY.txt"), "w").write(
    "# cuobjdump -sass c_orb_slam_b200/liborb_gpu.so (sm_100a); full listings: <kernel>.sass.gz\n" + "\n".join(summary) + "\n")
print("\n".join(summary))
