#!/usr/bin/env python3
"""Per-kernel time of the resident hot path (CUDA events around each launch, vo_profile_*), for quick
before/after comparisons of one kernel.

    python tools/kernel_times.py [frames] [batch] [match_bits] [rows] [cols] [preset] [max_kp] [rects]
"""
import json
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from acs_visual_odometry_b200 import Params, VoFront  # noqa: E402
from acs_visual_odometry_b200.synth import make_sequence  # noqa: E402

arg = sys.argv[1:]
frames_n = int(arg[0]) if len(arg) > 0 else 256
batch = int(arg[1]) if len(arg) > 1 else 64
bits = int(arg[2]) if len(arg) > 2 else 32
rows = int(arg[3]) if len(arg) > 3 else 376
cols = int(arg[4]) if len(arg) > 4 else 1241
preset = arg[5] if len(arg) > 5 else "vo_gpu_ref"
max_kp = int(arg[6]) if len(arg) > 6 else 0
rects = int(arg[7]) if len(arg) > 7 else 0
kw = {"n_rects": rects} if rects else {}
frames = make_sequence(frames_n, rows, cols, seed=0, **kw)
p = getattr(Params, preset)()
p.match_bits = bits
if max_kp:
    p.max_keypoints = max_kp
with VoFront(rows, cols, p.max_keypoints, 2 * batch) as vo:
    vo.sequence_upload(frames)
    for _ in range(3):
        vo.sequence_run_resident(p, batch)
    vo.sync()
    vo.timer_start()
    for _ in range(5):
        vo.sequence_run_resident(p, batch)
    total = vo.timer_stop() / 5
    vo.profile_enable(True)
    vo.profile_read(reset=True)
    for _ in range(5):
        vo.sequence_run_resident(p, batch)
    vo.sync()
    prof = vo.profile_read(reset=True)
    vo.profile_enable(False)
    nk, _, nm, _ = vo.sequence_fetch(want_kp=False, want_pairs=False)
out = {"frames": frames_n, "batch": batch, "rows": rows, "cols": cols, "match_bits": bits,
       "ms_per_pass": round(total, 4), "frames_per_s": round(frames_n / total * 1e3, 1),
       "keypoints_per_frame": float(np.mean(nk)), "matches_per_pair": float(np.mean(nm)),
       "us_per_launch": {k: round(v[0] / max(v[1], 1) * 1e3, 2) for k, v in prof.items()},
       "ms_per_pass_by_kernel": {k: round(v[0] / 5, 4) for k, v in prof.items()}}
print(json.dumps(out))
