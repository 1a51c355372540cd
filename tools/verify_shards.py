#!/usr/bin/env python3
"""Hardware check of frame-pair sharding ("1/2/4/8 devices give byte-identical per-pair outputs to
1 device"): the SAME sequence is extracted and matched once by a single GPU and once sharded over all ranks
(acs_visual_odometry_b200/shard.py, no collective on the data path), and every frame's keypoint list and every pair's
match list are compared through SHA-1 digests.  One blank frame is planted exactly on a shard boundary (both
neighbouring shards extract it; its two pairs have no matches).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
        tools/verify_shards.py [--frames 257] [--batch 64]

Rank 0 prints one JSON line with "shards_equal_single".  torch.distributed carries only the digests (reporting).
"""
import argparse
import hashlib
import json
import os
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))


def digests(vo, frames, params, batch):
    """Per-frame and per-pair SHA-1 of what the front end returns for `frames` (consecutive pairs matched)."""
    nk, kp, nm, pr = vo.sequence_run(frames, params, batch)
    fd = [hashlib.sha1(np.ascontiguousarray(kp[f, :nk[f]]).tobytes() + bytes([0]) + int(nk[f]).to_bytes(4, "little")).hexdigest()
          for f in range(len(frames))]
    pd = [hashlib.sha1(np.ascontiguousarray(pr[p, :nm[p]]).tobytes() + bytes([1]) + int(nm[p]).to_bytes(4, "little")).hexdigest()
          for p in range(len(frames) - 1)]
    return fd, pd, int(nk.sum()), int(nm.sum())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=257)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--rows", type=int, default=376)
    ap.add_argument("--cols", type=int, default=1241)
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", rank))
    import torch
    import torch.distributed as dist
    from acs_visual_odometry_b200 import Params, VoFront
    from acs_visual_odometry_b200.shard import shard_frames, shard_pairs
    from acs_visual_odometry_b200.synth import make_sequence
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    frames = make_sequence(a.frames, a.rows, a.cols, seed=7)
    planted = None
    if world > 1:
        planted = shard_pairs(a.frames, 0, world)[1]       # last frame of shard 0 = first frame of shard 1
        frames[planted] = 128
    p = Params.vo_gpu_ref()
    flo, fhi = shard_frames(a.frames, rank, world)
    plo, phi = shard_pairs(a.frames, rank, world)
    with VoFront(a.rows, a.cols, p.max_keypoints, 2 * a.batch, device=local) as vo:
        fd, pd, _, _ = digests(vo, frames[flo:fhi], p, a.batch) if fhi > flo else ([], [], 0, 0)
        mine = {"frames": {flo + k: d for k, d in enumerate(fd)}, "pairs": {plo + k: d for k, d in enumerate(pd)}}
        assert len(pd) == phi - plo
        if world > 1:
            gathered = [None] * world
            dist.all_gather_object(gathered, mine)
        else:
            gathered = [mine]
        if rank == 0:
            whole_f, whole_p, n_kp, n_m = digests(vo, frames, p, a.batch)
            bad_frames, bad_pairs, seen_pairs, boundary = [], [], set(), 0
            frame_hits = {}
            for g in gathered:
                for f, d in g["frames"].items():
                    frame_hits[f] = frame_hits.get(f, 0) + 1
                    if whole_f[f] != d:
                        bad_frames.append(f)
                for q, d in g["pairs"].items():
                    seen_pairs.add(q)
                    if whole_p[q] != d:
                        bad_pairs.append(q)
            boundary = sum(1 for c in frame_hits.values() if c == 2)
            ok = (not bad_frames and not bad_pairs and seen_pairs == set(range(a.frames - 1))
                  and set(frame_hits) == set(range(a.frames)) and boundary == world - 1)
            print(json.dumps({"check": "sharded front end == single GPU, SHA-1 per frame (keypoints) and per pair (match list)",
                              "shards_equal_single": bool(ok), "n_gpus": world, "frames": a.frames, "pairs": a.frames - 1,
                              "shape": [a.rows, a.cols], "batch": a.batch, "keypoints_total": n_kp, "matches_total": n_m,
                              "boundary_frames_extracted_twice": boundary, "planted_blank_frame": planted,
                              "mismatching_frames": bad_frames[:10], "mismatching_pairs": bad_pairs[:10]}), flush=True)
            rc = 0 if ok else 1
        else:
            rc = 0
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    sys.exit(rc)


if __name__ == "__main__":
    main()
