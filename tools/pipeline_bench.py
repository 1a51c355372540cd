#!/usr/bin/env python3
"""BASELINE.json config 5 in miniature: the pipelined VisualOdometry::run driver (host/vo_pipeline.cpp) on a
KITTI-shaped synthetic PNG sequence with a synthetic ground-truth pose file - decode, GPU front end, host RANSAC +
pose, pose CSV.  Reports front-end and end-to-end frames/s separately.

    python tools/pipeline_bench.py [frames=1000] [threads=all] [batch=64] [pgm|png]
"""
import json
import os
import sys
import tempfile
import time
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from acs_visual_odometry_b200 import host  # noqa: E402
from acs_visual_odometry_b200.synth import make_sequence  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
threads = int(sys.argv[2]) if len(sys.argv) > 2 and int(sys.argv[2]) > 0 else (os.cpu_count() or 8)
batch = int(sys.argv[3]) if len(sys.argv) > 3 else 64
fmt = sys.argv[4] if len(sys.argv) > 4 else "png"

with tempfile.TemporaryDirectory(dir=os.environ.get("TMPDIR", "/tmp")) as tmp:
    d = Path(tmp) / "seq"
    d.mkdir()
    t0 = time.time()
    frames = make_sequence(n, 376, 1241, seed=0)
    write = host.write_png if fmt == "png" else host.write_pgm
    with ThreadPoolExecutor(threads) as ex:  # the frame files always carry the .png name the driver asks for
        list(ex.map(lambda i: write(d / f"{i:06d}.png", frames[i]), range(n)))
    gt = Path(tmp) / "poses.txt"
    gt.write_text("".join(f"1 0 0 {0.02 * i:.6f} 0 1 0 0 0 0 1 {0.8 * i:.6f}\n" for i in range(n)))
    t_gen = time.time() - t0
    host.pipeline_run(str(d) + "/", min(n, 2 * batch + 1), gt, Path(tmp) / "warm.csv", num_threads=threads, batch=batch, seed=1)
    st = host.pipeline_run(str(d) + "/", n, gt, Path(tmp) / "out.csv", num_threads=threads, batch=batch, seed=1)
    lines = (Path(tmp) / "out.csv").read_text().count("\n")
out = {
    "workload": f"{n} synthetic KITTI-shaped 1241x376 {fmt} frames, preset vo_gpu_ref, REF32 matching, host RANSAC + pose",
    "host_threads": threads, "batch": batch, "csv_lines": lines, "generate_s": round(t_gen, 2),
    "e2e_frames_per_s": st["frames"] / st["wall_s"],
    "frontend_frames_per_s": st["frames"] / st["frontend_s"] if st["frontend_s"] else None,
    "ransac_pose_pairs_per_s": (st["pairs_ok"] + st["skipped_frames"]) / st["host_s"] if st["host_s"] else None,
    "decode_frames_per_s": st["frames"] / st["decode_s"] if st["decode_s"] else None,
    **st,
}
print(json.dumps(out))
