#!/usr/bin/env python3
"""Small fixed workload for ncu: a few batched passes of the hot path on KITTI-shaped synthetic frames.

    python tools/profile_run.py [frames] [batch] [match_bits]
"""
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from acs_visual_odometry_b200 import Params, VoFront  # noqa: E402
from acs_visual_odometry_b200.synth import make_sequence  # noqa: E402

frames_n = int(sys.argv[1]) if len(sys.argv) > 1 else 500
batch = int(sys.argv[2]) if len(sys.argv) > 2 else 250
bits = int(sys.argv[3]) if len(sys.argv) > 3 else 32
frames = make_sequence(frames_n, 376, 1241, seed=0)
p = Params.vo_gpu_ref()
p.match_bits = bits
with VoFront(376, 1241, p.max_keypoints, 2 * batch) as vo:
    vo.sequence_upload(frames)
    for _ in range(3):
        vo.sequence_run_resident(p, batch)
    vo.sync()
    nk, _, nm, _ = vo.sequence_fetch(want_kp=False, want_pairs=False)
    # the response-writing form of the stencil kernel is not part of the step: launch it alone so that a capture of this
    # program holds both forms (vo_bench_stencil mode 0, then mode 1 = the fused form alone)
    print("stencil alone ms: unfused", vo.bench_stencil(p, min(batch, frames_n), fused=False, reps=2), "fused", vo.bench_stencil(p, min(batch, frames_n), fused=True, reps=2))
    print("keypoints/frame", float(np.mean(nk)), "matches/pair", float(np.mean(nm)), "launches", vo.launch_count())
