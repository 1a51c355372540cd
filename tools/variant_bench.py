#!/usr/bin/env python3
"""Time kernel-tuning variants of libvofront.so against each other on the bench shape, and prove that they compute the
same thing.  One subprocess per library (the binding caches its dlopen):

    python tools/variant_bench.py [--frames 1000] [--batch 1000] LIB [LIB ...]      # driver
    VO_LIB=LIB python tools/variant_bench.py --child ...                              # one library

Per library, one JSON line: per-kernel ms per pass (CUDA events around every launch, vo_profile_*), the stencil kernel
alone in both forms (vo_bench_stencil: 0 = response written, 1 = fused NMS), the un-instrumented step time, and a SHA-1 of
every output of the pass (keypoint counts, keypoints, match counts, pairs) plus of the descriptors, blurred image and
response map of one frame - variants must agree on all of them.
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))


def child(a):
    import numpy as np

    from acs_visual_odometry_b200 import Params, VoFront
    from acs_visual_odometry_b200 import api
    from acs_visual_odometry_b200.synth import make_sequence

    frames = make_sequence(a.frames, a.rows, a.cols, seed=0)
    p = Params.vo_gpu_ref() if a.preset == "vo_gpu_ref" else Params.cpu_ref()
    p.match_bits = a.bits
    if a.max_kp:
        p.max_keypoints = a.max_kp
    if a.threshold:
        p.threshold = a.threshold
    out = {"lib": os.environ.get("VO_LIB", "in-tree"), "frames": a.frames, "batch": a.batch}
    with VoFront(a.rows, a.cols, p.max_keypoints, 2 * a.batch) as vo:
        vo.sequence_upload(frames)
        for _ in range(2):
            vo.sequence_run_resident(p, a.batch)
        vo.sync()
        vo.timer_start()
        for _ in range(a.passes):
            vo.sequence_run_resident(p, a.batch)
        out["step_ms"] = round(vo.timer_stop() / a.passes, 4)
        nk, kp, nm, pr = vo.sequence_fetch()
        h = hashlib.sha1()
        for i in range(len(nk)):
            h.update(np.ascontiguousarray(kp[i, :nk[i]]).tobytes())
        for i in range(len(nm)):
            h.update(np.ascontiguousarray(pr[i, :nm[i]]).tobytes())
        h.update(nk.tobytes()); h.update(nm.tobytes())
        out["sha_sequence"] = h.hexdigest()[:16]
        out["keypoints_per_frame"] = float(np.mean(nk)); out["matches_per_pair"] = float(np.mean(nm))
        vo.profile_enable(True)
        vo.profile_read(reset=True)
        for _ in range(a.passes):
            vo.sequence_run_resident(p, a.batch)
        vo.sync()
        prof = vo.profile_read(reset=True)
        vo.profile_enable(False)
        out["kernel_ms"] = {k: round(v[0] / a.passes, 4) for k, v in prof.items()}
        sb = min(a.batch, a.frames)
        out["stencil_fused_ms"] = round(vo.bench_stencil(p, sb, True, reps=5), 4)
        out["stencil_unfused_ms"] = round(vo.bench_stencil(p, sb, False, reps=5), 4)
        px = a.rows * a.cols * sb
        out["stencil_unfused_GBs"] = round(6.0 * px / out["stencil_unfused_ms"] / 1e6, 1)
        # single-frame outputs of every stage
        h = hashlib.sha1()
        g = frames[a.frames // 2]
        n = vo.extract(g, p, slot=0)
        kp1, desc, packed = vo.fetch(0)
        h.update(kp1.tobytes()); h.update(packed.tobytes())
        out["sha_desc"] = h.hexdigest()[:16]
        b = vo.blur(g)
        R = vo.response(b, p.threshold)
        out["sha_blur_R"] = hashlib.sha1(b.tobytes() + R.tobytes()).hexdigest()[:16]
        p7 = Params(p.threshold, 7, p.max_keypoints)
        vo.extract(g, p7, slot=1)
        out["sha_unfused_kp"] = hashlib.sha1(vo.fetch(1)[0].tobytes()).hexdigest()[:16]
        out["sha_R_dbg"] = hashlib.sha1(vo.debug(1, api.DBG_RESPONSE, a.rows, a.cols).tobytes()).hexdigest()[:16]
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("libs", nargs="*")
    ap.add_argument("--child", action="store_true")
    ap.add_argument("--frames", type=int, default=1000)
    ap.add_argument("--batch", type=int, default=1000)
    ap.add_argument("--rows", type=int, default=376)
    ap.add_argument("--cols", type=int, default=1241)
    ap.add_argument("--bits", type=int, default=32)
    ap.add_argument("--passes", type=int, default=3)
    ap.add_argument("--preset", default="vo_gpu_ref")
    ap.add_argument("--max-kp", type=int, default=0)
    ap.add_argument("--threshold", type=float, default=0.0)
    a = ap.parse_args()
    if a.child:
        return child(a)
    fwd = [x for x in sys.argv[1:] if x not in a.libs]
    for lib in a.libs or [""]:
        env = dict(os.environ)
        if lib:
            env["VO_LIB"] = str(Path(lib).resolve())
        r = subprocess.run([sys.executable, __file__, "--child", *fwd], env=env, capture_output=True, text=True)
        sys.stdout.write(r.stdout)
        if r.returncode:
            print(json.dumps({"lib": lib, "error": r.stderr[-600:]}), flush=True)


if __name__ == "__main__":
    main()
