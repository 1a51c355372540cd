#!/usr/bin/env python3
"""BASELINE.json configs[4]: a KITTI-shaped sequence sharded by frame pairs over the GPUs of one box, each rank
running the pipelined VisualOdometry::run driver (host/vo_pipeline.cpp: decode -> GPU front end -> host RANSAC + pose)
on its shard; rank 0 stitches the shard trajectories with the one sequential step (shard.compose_shard_poses) and
writes the pose CSV.  No data-path collective: torch.distributed (gloo) carries only the barrier and the gather of
the per-shard pose lists and timings.

    python tools/pipeline_shard_bench.py [--frames 10000] [--fmt pgm|png] [--batch 64]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/pipeline_shard_bench.py ...
"""
import argparse
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
from acs_visual_odometry_b200.shard import compose_shard_poses, shard_frames  # noqa: E402
from acs_visual_odometry_b200.synth import make_sequence  # noqa: E402


def run_shard_pipeline(frames, fmt, batch, threads, out_path, rank, world, local, dist):
    """One rank's share of the job; `dist` is an initialised torch.distributed (any backend) or None when world == 1.
    Returns the summary dict on rank 0, None elsewhere."""
    def barrier():
        if world > 1:
            dist.barrier()

    threads = threads if threads > 0 else max(1, (os.cpu_count() or 8) // world)
    lo, hi = shard_frames(frames, rank, world)  # frames [lo, hi): this rank's pairs plus the shared right boundary frame
    n = hi - lo
    with tempfile.TemporaryDirectory(dir=os.environ.get("TMPDIR", "/tmp")) as tmp:
        d = Path(tmp) / "seq"
        d.mkdir()
        t0 = time.time()
        write = host.write_png if fmt == "png" else host.write_pgm
        step = 250  # generate and write in pieces: a 10 k-frame sequence is 4.7 GB
        for s0 in range(0, n, step):
            fr = make_sequence(min(step, n - s0), 376, 1241, seed=0, start=lo + s0)
            with ThreadPoolExecutor(threads) as ex:  # the files always carry the .png name the driver asks for
                list(ex.map(lambda i: write(d / f"{s0 + i:06d}.png", fr[i]), range(len(fr))))
        gt = Path(tmp) / "poses.txt"
        gt.write_text("".join(f"1 0 0 {0.02 * i:.6f} 0 1 0 0 0 0 1 {0.8 * i:.6f}\n" for i in range(lo, hi)))
        t_gen = time.time() - t0
        csv = Path(tmp) / "out.csv"
        host.pipeline_run(str(d) + "/", min(n, 2 * batch + 1), gt, Path(tmp) / "warm.csv", num_threads=threads, device=local,
                          batch=batch, seed=1)
        barrier()
        t0 = time.perf_counter()
        st = host.pipeline_run(str(d) + "/", n, gt, csv, num_threads=threads, device=local, batch=batch, seed=1 + lo)
        t_local = time.perf_counter() - t0
        barrier()
        t_all = time.perf_counter() - t0
        poses = []
        for line in csv.read_text().splitlines():
            v = [float(x) for x in line.replace(",", " ").split()]
            if len(v) >= 12:
                poses.append(np.vstack([np.array(v[:12]).reshape(3, 4), [0, 0, 0, 1.0]]))
    mine = {**st, "rank": rank, "frames": n, "wall_s": t_local, "generate_s": t_gen, "poses": poses}
    gathered = [mine]
    if world > 1:
        gathered = [None] * world
        dist.all_gather_object(gathered, mine)
    if rank != 0:
        return None
    gathered.sort(key=lambda g: g["rank"])
    t0 = time.perf_counter()
    traj = compose_shard_poses([g["poses"] for g in gathered])
    t_compose = time.perf_counter() - t0
    if out_path:
        with open(out_path, "w") as f:
            for T in traj:
                f.write(host.format_pose_csv(T).rstrip("\n") + "\n")
    wall = max(g["wall_s"] for g in gathered)
    return {
        "workload": f"{frames} synthetic KITTI-shaped 1241x376 {fmt} frames sharded by frame pairs over {world} GPU(s), "
                    f"preset vo_gpu_ref, REF32 matching, host RANSAC + pose per shard, prefix product on rank 0",
        "n_gpus": world, "host_threads_per_rank": threads, "batch": batch, "trajectory_poses": len(traj),
        "e2e_frames_per_s": frames / (wall + t_compose), "wall_s_max_over_ranks": wall, "barrier_to_barrier_s": t_all,
        "compose_s": t_compose,
        "frontend_frames_per_s_sum_over_ranks": sum(g["frames"] / g["frontend_s"] for g in gathered if g["frontend_s"]),
        "ransac_pose_pairs_per_s_sum_over_ranks": sum((g["pairs_ok"] + g["skipped_frames"]) / g["host_s"] for g in gathered if g["host_s"]),
        "decode_frames_per_s_sum_over_ranks": sum(g["frames"] / g["decode_s"] for g in gathered if g["decode_s"]),
        "pairs_ok": sum(g["pairs_ok"] for g in gathered), "skipped_frames": sum(g["skipped_frames"] for g in gathered),
        "mean_inliers": float(np.mean([g["mean_inliers"] for g in gathered])),
        "per_rank_wall_s": [round(g["wall_s"], 3) for g in gathered],
        # A shard whose LAST pair failed hands the next shard a boundary frame that the sequential run would not have
        # used as its query frame (it keeps matching against the last valid frame, VisualOdometry.cpp:164-166, and that
        # state does not cross shards): the stitched trajectory then differs from the 1-GPU one from that boundary on.
        # A skipped frame repeats the previous pose, so it shows in the shard's own pose list.
        "boundary_pairs_failed": sum(1 for g in gathered[:-1] if len(g["poses"]) >= 2 and np.allclose(g["poses"][-1], g["poses"][-2])),
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=10000)
    ap.add_argument("--fmt", default="pgm", choices=["pgm", "png"])
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--out", default="")
    ap.add_argument("--threads", type=int, default=0, help="host threads per rank (0 = all cores / ranks); RANSAC's chunking, and so "
                    "its result, depends on it (ransac.cpp:152-157), so trajectories are comparable only at equal thread counts")
    a = ap.parse_args()
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    import torch.distributed as dist
    if world > 1:
        dist.init_process_group("gloo")
    out = run_shard_pipeline(a.frames, a.fmt, a.batch, a.threads, a.out, rank, world, local, dist if world > 1 else None)
    if out is not None:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
