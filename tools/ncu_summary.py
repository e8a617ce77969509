#!/usr/bin/env python
"""Summarises ncu outputs brought back in gpurun_out/ into small text files under profiles/.
usage: ncu_summary.py launches <launches.csv> <out.txt> | full <report.ncu-rep> <out.txt>"""
import collections
import csv
import subprocess
import sys

mode, src, dst = sys.argv[1:4]
out = []
if mode == "launches":
    rows = [r for r in csv.reader(open(src)) if len(r) > 10]
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        v = float(r[vi].replace(",", ""))
        v = v / 1e3 if r[ui] in ("ns", "nsecond") else v * 1e3 if r[ui] in ("ms", "msecond") else v
        k = r[ki].split("(")[0]
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(v[1] for v in agg.values())
    out.append("# per-kernel device time from `ncu --metrics gpu__time_duration.sum --clock-control none` (cold-cache,")
    out.append("# serialised launches: compare SHARES, not absolutes).  source: %s" % src)
    out.append("%-64s %6s %12s %7s" % ("kernel", "n", "total_us", "share"))
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        out.append("%-64s %6d %12.1f %7.3f" % (k[:64], v[0], v[1], v[1] / tot))
else:
    raw = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr = rows[0]
    want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
            "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
            "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static",
            "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "inst_executed",
            "TPC.TriageCompute.sm__inst_executed_pipe_alu_realtime.avg.pct_of_peak_sustained_elapsed",
            "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
            "sm__inst_executed.sum",
            "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_lsu.sum", "sm__inst_executed_pipe_xu.sum"]
    idx = [(w, hdr.index(w)) for w in want if w in hdr]
    out.append("# `ncu --set full --clock-control none --import-source on`, raw page; source: %s" % src)
    for r in rows[2:]:
        out.append("-" * 100)
        for w, i in idx:
            out.append("%-62s %s %s" % (w, r[i][:70], rows[1][i] if w != "Kernel Name" else ""))
open(dst, "w").write("\n".join(out) + "\n")
print("\n".join(out[:60]))
