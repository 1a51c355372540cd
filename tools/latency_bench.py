#!/usr/bin/env python3
"""Frame-at-a-time rate of the drop-in entry points, i.e. what an unchanged VisualOdometry::run loop sees
(VisualOdometry.cpp:64-100: extract frame i, match against frame i-1, synchronously, one frame per call):
vo_extract (H2D + six kernels) + vo_frame_fetch (keypoints + 512-byte descriptors, as the reference's API returns
them) + vo_match (device-resident descriptors).  Prints one JSON line."""
import json
import sys
import time
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from acs_visual_odometry_b200 import MATCH_REF32, Params, VoFront  # noqa: E402
from acs_visual_odometry_b200.synth import make_sequence  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 200
frames = make_sequence(n, 376, 1241, seed=0)
p = Params.vo_gpu_ref()
res = {}
with VoFront(376, 1241, p.max_keypoints, 4) as vo:
    for mode in ("adapter: extract+fetch(desc512)+match_host(desc512)", "extract+fetch(desc512)+match", "extract+fetch(packed)+match", "extract+match"):
        for rep in range(2):
            t0 = time.perf_counter()
            nm = 0
            prev = None
            for i in range(n):
                vo.extract(frames[i], p, slot=i & 1)
                if mode.startswith("adapter"):  # what include/vofront_adapter.hpp does with the reference's vector<vector<uint8_t>> API
                    _, d512, _ = vo.fetch(i & 1, want_desc512=True, want_packed=False)
                    if prev is not None:
                        nm += len(vo.match_host(prev, d512, p.ratio, MATCH_REF32))
                    prev = d512
                    continue
                if mode != "extract+match":
                    vo.fetch(i & 1, want_desc512=mode.startswith("extract+fetch(desc512)"), want_packed=True)
                if i:
                    nm += len(vo.match((i - 1) & 1, i & 1, p.ratio, MATCH_REF32))
            dt = time.perf_counter() - t0
        res[mode] = {"frames_per_s": round(n / dt, 1), "ms_per_frame": round(dt / n * 1e3, 3), "matches_per_pair": round(nm / (n - 1), 1)}
print(json.dumps({"workload": f"{n} KITTI-shaped frames, one synchronous call per frame (preset vo_gpu_ref, REF32)", **res}))
