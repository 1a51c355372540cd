#!/usr/bin/env python3
"""Small end-to-end workload for compute-sanitizer (memcheck / racecheck / initcheck): every kernel family, aligned and
unaligned rows, edge strips, tails, both matchers.

    compute-sanitizer --tool memcheck python tools/sanitize_run.py
"""
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from acs_visual_odometry_b200 import MATCH_FULL512, MATCH_REF32, Params, VoFront  # noqa: E402
from acs_visual_odometry_b200.synth import make_sequence  # noqa: E402

for rows, cols in ((130, 251), (96, 128), (200, 333)):
    frames = make_sequence(9, rows, cols, seed=rows, n_rects=120)
    p = Params(20000.0, 3, 300)
    with VoFront(rows, cols, 300, 8) as vo:
        for bits in (MATCH_REF32, MATCH_FULL512):
            p.match_bits = bits
            nk, kp, nm, pr = vo.sequence_run(frames, p, 4)          # unaligned staging rows
            vo.extract_batch(frames[:4], p, 0)                        # aligned slot rows
            m = vo.match(0, 1, p.ratio, bits)
            print(rows, cols, bits, int(nk.sum()), int(nm.sum()), len(m))
        vo.extract_bgr(np.repeat(frames[0][:, :, None], 3, axis=2), p, 0)
        b = vo.blur(frames[0]); R = vo.response(b, 20000.0); k = vo.nms(R, 3, 300); d = vo.describe(b, k)
print("sanitize workload done")
