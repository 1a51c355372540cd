#!/usr/bin/env python3
"""Per-kernel time of the resident hot path (CUDA events around each launch, vo_profile_*) on synthetic frames of
any BASELINE.json config; prints one JSON line.

    python tools/kernel_times.py [--frames N] [--batch B] [--bits 32|512] [--rows R] [--cols C] [--preset P]
                                 [--max-kp K] [--rects N] [--threshold T] [--passes P] [--tag NAME]

configs[1] (KITTI)   : defaults
configs[2] (1080p)   : --rows 1080 --cols 1920 --max-kp 8000 --rects 6000 --threshold 6000 --bits 512 --frames 64 --batch 16
configs[3] (4K sweep): --rows 2160 --cols 3840 --rects 20000 --frames 64 --batch 1..64   (read us_per_launch.stencil)
"""
import argparse
import json
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from acs_visual_odometry_b200 import Params, VoFront  # noqa: E402
from acs_visual_odometry_b200.synth import make_sequence  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--frames", type=int, default=256)
ap.add_argument("--batch", type=int, default=100)
ap.add_argument("--bits", type=int, default=32)
ap.add_argument("--rows", type=int, default=376)
ap.add_argument("--cols", type=int, default=1241)
ap.add_argument("--preset", default="vo_gpu_ref")
ap.add_argument("--max-kp", type=int, default=0)
ap.add_argument("--rects", type=int, default=0)
ap.add_argument("--threshold", type=float, default=0.0)
ap.add_argument("--passes", type=int, default=5)
ap.add_argument("--tag", default="")
a = ap.parse_args()

kw = {"n_rects": a.rects} if a.rects else {}
frames = make_sequence(a.frames, a.rows, a.cols, seed=0, **kw)
p = getattr(Params, a.preset)()
p.match_bits = a.bits
if a.max_kp:
    p.max_keypoints = a.max_kp
if a.threshold:
    p.threshold = a.threshold
P = a.passes
with VoFront(a.rows, a.cols, p.max_keypoints, 2 * a.batch) as vo:
    vo.sequence_upload(frames)
    for _ in range(3):
        vo.sequence_run_resident(p, a.batch)
    vo.sync()
    vo.timer_start()
    for _ in range(P):
        vo.sequence_run_resident(p, a.batch)
    total = vo.timer_stop() / P
    vo.profile_enable(True)
    vo.profile_read(reset=True)
    for _ in range(P):
        vo.sequence_run_resident(p, a.batch)
    vo.sync()
    prof = vo.profile_read(reset=True)
    vo.profile_enable(False)
    nk, _, nm, _ = vo.sequence_fetch(want_kp=False, want_pairs=False)
    # the stencil kernel alone, both forms (vo_bench_stencil): response-writing (6 B/px) and fused NMS (2 B/px)
    nb = min(a.batch, a.frames)
    st_unfused = vo.bench_stencil(p, nb, fused=False, reps=5)
    try:
        st_fused = vo.bench_stencil(p, nb, fused=True, reps=5)
    except Exception:
        st_fused = None
px = a.rows * a.cols
st_ms, st_n = prof["stencil"]
compares = float(np.sum(nk[:-1].astype(np.float64) * nk[1:]))
out = {"tag": a.tag, "frames": a.frames, "batch": a.batch, "rows": a.rows, "cols": a.cols, "match_bits": a.bits,
       "threshold": p.threshold, "max_keypoints": p.max_keypoints,
       "ms_per_pass": round(total, 4), "frames_per_s": round(a.frames / total * 1e3, 1),
       "keypoints_per_frame": float(np.mean(nk)), "matches_per_pair": float(np.mean(nm)) if len(nm) else 0.0,
       "us_per_launch": {k: round(v[0] / max(v[1], 1) * 1e3, 2) for k, v in prof.items()},
       "ms_per_pass_by_kernel": {k: round(v[0] / P, 4) for k, v in prof.items()},
       "stencil_alone_us": {"write_R_6B_per_px": round(st_unfused * 1e3, 2), "fused_nms_2B_per_px": round(st_fused * 1e3, 2) if st_fused else None,
                            "write_R_GBps": round(6.0 * px * nb / (st_unfused / 1e3) / 1e9, 1),
                            "fused_GBps": round(2.0 * px * nb / (st_fused / 1e3) / 1e9, 1) if st_fused else None},
       "stencil_GBps_algorithmic_6B_per_px": round(6.0 * px * a.frames * P / (st_ms / 1e3) / 1e9, 1) if st_ms else None,
       "compares_per_s_matcher_only": round(compares * P / (prof["match"][0] / 1e3), 1) if prof["match"][0] else None,
       "popc32_per_s_matcher_only": round(compares * P * (16 if a.bits == 512 else 1) / (prof["match"][0] / 1e3), 1) if prof["match"][0] else None}
print(json.dumps(out))
