#!/usr/bin/env python3
"""profiles/stencil_traffic.json from an `ncu --set full` raw page: DRAM bytes and warp instructions per launch of the two
forms of k_stencil (bench.py reads it for `roofline.traffic` and the issue-slot figures).

    ncu -i REPORT.ncu-rep --page raw --csv > raw.csv
    python tools/make_stencil_traffic.py raw.csv BATCH ROWS COLS "source text" > profiles/stencil_traffic.json
"""
import csv
import json
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, body = rows[0], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
units = rows[1]
sys.path.insert(0, str(__import__("pathlib").Path(__file__).resolve().parents[1]))
from acs_visual_odometry_b200.build import source_hash  # noqa: E402

# the hash is taken from the working tree: run this tool on the code the capture ran
out = {"batch": int(sys.argv[2]), "rows": int(sys.argv[3]), "cols": int(sys.argv[4]), "source": sys.argv[5],
       "stencil_code_sha1": source_hash("stencil.cu")}


def scale(col):
    u = units[ix[col]].lower()
    return {"byte": 1.0, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1.0)


for r in body:
    name = r[ix["Kernel Name"]]
    m = re.search(r"k_stencil<(\d+), (\d+), (\d+), (\d+)", name.replace("(int)", "").replace("(bool)", ""))
    if not m:
        continue
    kind = "fused" if int(m.group(4)) > 0 else "unfused"
    rd = float(r[ix["dram__bytes_read.sum"]]) * scale("dram__bytes_read.sum")
    wr = float(r[ix["dram__bytes_write.sum"]]) * scale("dram__bytes_write.sum")
    out[kind] = {"kernel": name.split("(St")[0], "dram_bytes_per_launch": rd + wr, "dram_read": rd, "dram_write": wr,
                 "warp_instructions_per_launch": float(r[ix["smsp__inst_executed.sum"]]),
                 "duration_ms_under_ncu": float(r[ix["gpu__time_duration.sum"]]) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(units[ix["gpu__time_duration.sum"]].lower(), 1e-3),
                 "issue_active_pct": float(r[ix["smsp__issue_active.avg.pct_of_peak_sustained_active"]]),
                 "warps_active_pct": float(r[ix["sm__warps_active.avg.pct_of_peak_sustained_active"]])}
print(json.dumps(out, indent=1))
