#!/usr/bin/env python
"""DRAM bytes and warp instructions per frame and stage from a full ncu capture of one extraction step
(tools/gpu_ncu.sh) -> profiles/traffic_<tag>.json, which bench.py reads for roofline.traffic and roofline.issue.
usage: ncu_traffic.py <report.ncu-rep> <frames per launch> <out.json>"""
import csv
import json
import subprocess
import sys

rep, frames, dst = sys.argv[1], int(sys.argv[2]), sys.argv[3]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
col = {n: hdr.index(n) for n in ("Kernel Name", "dram__bytes_read.sum", "dram__bytes_write.sum", "inst_executed",
                                 "gpu__time_duration.sum")}
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
stage_of = {"resize_level": "pyramid", "repack_rows": "pyramid", "fast_cells": "fast", "quadtree": "quadtree",
            "blur": "blur", "orient_desc": "orient_desc"}
byt, inst, us, seen = {}, {}, {}, {}
for r in rows[2:]:
    name = r[col["Kernel Name"]]
    name = name[5:] if name.startswith("void ") else name  # templated kernels are listed with their return type
    st = next((v for k, v in stage_of.items() if name.startswith(k)), None)
    if st is None:
        continue
    b = sum(float(r[col[c]].replace(",", "")) * scale[units[col[c]]] for c in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
    byt[st] = byt.get(st, 0.0) + b
    inst[st] = inst.get(st, 0.0) + float(r[col["inst_executed"]].replace(",", ""))
    us[st] = us.get(st, 0.0) + float(r[col["gpu__time_duration.sum"]].replace(",", ""))
    seen.setdefault(st, set()).add(name.split("(")[0])
out = {"source": "ncu --set full --clock-control none (%s), one step of %d frames; dram__bytes_read.sum + dram__bytes_write.sum "
                 "and inst_executed per stage, divided by the frames of the launch (pyramid = the seven resize launches)" % (rep, frames),
       "kernels": {k: sorted(v) for k, v in seen.items()},
       "bytes_per_frame": {k: round(v / frames) for k, v in byt.items()},
       "warp_inst_per_frame": {k: round(v / frames) for k, v in inst.items()},
       "us_per_step_under_ncu": {k: round(v, 1) for k, v in us.items()}}
json.dump(out, open(dst, "w"), indent=1)
print(json.dumps(out, indent=1))
