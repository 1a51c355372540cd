#!/usr/bin/env python3
"""Where the end-to-end step's time above the bare H2D copy goes: vo_sequence_run with and without the result copies."""
import sys, time
from pathlib import Path
import numpy as np
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from acs_visual_odometry_b200 import Params, VoFront
from acs_visual_odometry_b200.api import PinnedBuffer
from acs_visual_odometry_b200.synth import make_sequence
n = 1000
pinned = PinnedBuffer((n, 376, 1241), np.uint8)
frames = make_sequence(n, 376, 1241, seed=0, out=pinned.array)
p = Params.vo_gpu_ref()
mk = p.max_keypoints
bufs = (PinnedBuffer((n,), np.int32), PinnedBuffer((n, mk, 2), np.float32), PinnedBuffer((n - 1,), np.int32), PinnedBuffer((n - 1, mk, 2), np.int32))
full = tuple(b.array for b in bufs)
with VoFront(376, 1241, 2000, 500) as vo:
    for batch in (50, 100, 200, 250):
        for kw in (dict(out=full), dict(out=(full[0], None, full[2], None))):
            for _ in range(2):
                vo.sequence_run(frames, p, batch, **kw)
            t0 = time.perf_counter()
            for _ in range(5):
                vo.sequence_run(frames, p, batch, **kw)
            print(batch, "all results" if kw["out"][1] is not None else "counts only", round((time.perf_counter() - t0) / 5 * 1e3, 3), "ms")
