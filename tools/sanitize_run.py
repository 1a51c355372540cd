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

# fused NMS with both window sizes it takes (3 and 5), on scene frames and on a dense frame (uniform noise: a local
# maximum every few pixels, so a warp stages more than 32 candidates between flushes), then the same through the
# persistent-task form of the stencil kernel (VO_STENCIL_CAP is read by vo_create)
import os  # noqa: E402

rows, cols = 150, 420
frames = make_sequence(6, rows, cols, seed=11, n_rects=150)
frames[2] = np.random.default_rng(5).integers(0, 256, (rows, cols), dtype=np.uint8)
for cap_env in (None, "2"):
    if cap_env is None:
        os.environ.pop("VO_STENCIL_CAP", None)
    else:
        os.environ["VO_STENCIL_CAP"] = cap_env
    with VoFront(rows, cols, 2000, 8, cand_cap=rows * cols) as vo:
        for window in (3, 5):
            p = Params(20000.0, window, 2000)
            nk, kp, nm, pr = vo.sequence_run(frames, p, 3)
            print("fused window", window, "stencil cap", cap_env, nk.tolist(), nm.tolist())
os.environ.pop("VO_STENCIL_CAP", None)

# the context-free matcher (static device scratch): two calls in a row, second one smaller
import ctypes as C  # noqa: E402

import torch  # noqa: E402

from acs_visual_odometry_b200 import load_library  # noqa: E402

lib = load_library()
rng = np.random.default_rng(9)
q = torch.empty((700, 16), dtype=torch.int32, device="cuda")
t = torch.empty((700, 16), dtype=torch.int32, device="cuda")
pairs = torch.empty((700, 2), dtype=torch.int32, device="cuda")
count = torch.empty((1,), dtype=torch.int32, device="cuda")
for n1, n2, bits in ((700, 650, MATCH_FULL512), (130, 9, MATCH_REF32), (257, 700, MATCH_FULL512)):
    d = rng.integers(0, 2**31, (700, 16)).astype(np.int32)
    q.copy_(torch.from_numpy(d)); t.copy_(torch.from_numpy(np.roll(d, 3, axis=0)))
    rc = lib.vo_match_static(C.c_void_p(q.data_ptr()), n1, C.c_void_p(t.data_ptr()), n2, 0.75, bits,
                             C.c_void_p(pairs.data_ptr()), C.c_void_p(count.data_ptr()), None)
    torch.cuda.synchronize()
    print("vo_match_static", n1, n2, bits, rc, int(count.cpu()[0]))
print("sanitize workload done")
