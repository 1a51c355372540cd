"""world_size-2 gloo test of the multi-GPU host logic: frame-pair shards computed by independent processes
(no data-path collective) reproduce the single-process match lists.  The per-shard worker here is the CPU
oracle (tests may use it); on the GPU box the same sharding drives one vo_ctx per rank (bench.py)."""
import os
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]


def _worker(rank, world, port, n_frames, q):
    sys.path.insert(0, str(ROOT))
    import torch.distributed as dist
    from acs_visual_odometry_b200.shard import shard_frames, shard_pairs
    from acs_visual_odometry_b200.synth import make_sequence
    from oracle.pyoracle import Oracle
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    o = Oracle()
    frames = make_sequence(n_frames, 150, 200, seed=3, n_rects=120)
    flo, fhi = shard_frames(n_frames, rank, world)
    plo, phi = shard_pairs(n_frames, rank, world)
    feats = [o.extract(frames[f], 25000.0, 5, 1500) for f in range(flo, fhi)]
    mine = {p: o.match(feats[p - flo][1], feats[p + 1 - flo][1], 32).tolist() for p in range(plo, phi)}
    dist.barrier()
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)   # reporting only; the data path itself exchanges nothing
    if rank == 0:
        merged = {}
        for g in gathered:
            merged.update(g)
        q.put(merged)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_sharding_reproduces_single_process():
    import torch.multiprocessing as mp
    from acs_visual_odometry_b200.synth import make_sequence
    from oracle.pyoracle import Oracle
    n_frames, world, port = 5, 2, 29500 + os.getpid() % 2000
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_frames, q)) for r in range(world)]
    for p in procs:
        p.start()
    merged = q.get(timeout=240)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    o = Oracle()
    frames = make_sequence(n_frames, 150, 200, seed=3, n_rects=120)
    feats = [o.extract(f, 25000.0, 5, 1500) for f in frames]
    assert sorted(merged) == list(range(n_frames - 1))
    for p in range(n_frames - 1):
        assert merged[p] == o.match(feats[p][1], feats[p + 1][1], 32).tolist()
    assert sum(len(v) for v in merged.values()) > 0


def test_shard_pose_composition_equals_the_unsharded_trajectory():
    """compose_shard_poses (the one sequential step of the sharded VisualOdometry::run: T_curr = T_curr * T_rel,
    VisualOdometry.cpp:184) stitches per-shard CSV pose lists into the trajectory a single driver writes; shards of the
    synthetic sequence equal slices of the whole."""
    from acs_visual_odometry_b200.shard import compose_shard_poses, shard_frames
    from acs_visual_odometry_b200.synth import make_sequence
    whole_frames = make_sequence(6, 60, 80, seed=3, n_rects=20)
    assert np.array_equal(whole_frames[2:5], make_sequence(3, 60, 80, seed=3, n_rects=20, start=2))
    rng = np.random.default_rng(0)
    n = 23
    rels = []
    for _ in range(n - 1):
        th = rng.normal(0, 0.05)
        T = np.eye(4)
        T[:3, :3] = [[np.cos(th), 0, np.sin(th)], [0, 1, 0], [-np.sin(th), 0, np.cos(th)]]
        T[:3, 3] = rng.normal(0, 1, 3)
        rels.append(T)
    F = np.diag([1, 1, -1, 1.0])

    def driver(lo, hi):  # poses a driver writes for frames lo..hi: identity, then flipZ * T_curr
        T, out = np.eye(4), [np.eye(4)]
        for i in range(lo + 1, hi + 1):
            T = T @ rels[i - 1]
            out.append(F @ T)
        return out

    whole = driver(0, n - 1)
    for world in (1, 2, 3, 5, 8):
        shards = [driver(*(lambda lo, hi: (lo, hi - 1))(*shard_frames(n, r, world))) for r in range(world)]
        got = compose_shard_poses(shards)
        assert len(got) == n
        assert max(np.abs(g - w).max() for g, w in zip(got, whole)) < 1e-12


def test_failed_boundary_pair_is_the_stated_limit_of_sharding():
    """The reference advances its query frame only on success (VisualOdometry.cpp:164-166) and that state does not cross
    shard boundaries.  With the sequential frame-loop oracle on scripted stages: when the pair that ENDS a shard
    succeeds, the stitched shards equal the sequential run; when it fails, the next shard starts from a frame the
    sequential run never uses as its query and the trajectories part at the boundary (shard.py states the restriction,
    tools/pipeline_shard_bench.py counts such boundaries)."""
    from acs_visual_odometry_b200.shard import compose_shard_poses, shard_frames
    from oracle.vo_loop_oracle import run_loop
    n = 13
    gt = np.zeros((n, 3, 4))
    gt[:, :, :3] = np.eye(3)
    gt[:, 2, 3] = np.arange(n) * 2.0

    def run(lo, hi, bad):  # frames lo..hi inclusive as one driver would process them; `bad` frames have no features
        frames = [np.full((2, 2), i, np.uint8) for i in range(lo, hi + 1)]

        def extract(img):
            i = int(img[0, 0])
            return np.c_[np.arange(32, dtype=np.float32) + i, np.arange(32, dtype=np.float32)], i

        def match(d1, d2):
            m = 0 if d2 in bad else 20
            return np.c_[np.arange(m), np.arange(m)].astype(np.int32)

        def ransac(pts, i):
            return np.eye(3), pts

        def get_pose(F, p1, p2, s):  # the step depends on WHICH query frame the pair started from
            th = 0.01 * float(p1[0, 0])
            R = np.array([[np.cos(th), 0, np.sin(th)], [0, 1, 0], [-np.sin(th), 0, np.cos(th)]])
            return R, np.array([0, 0, s])

        return list(run_loop(frames, gt[lo:hi + 1], extract, match, ransac, get_pose)[0])

    lo1, hi1 = shard_frames(n, 0, 2)
    lo2, hi2 = shard_frames(n, 1, 2)
    boundary = hi1 - 1
    assert lo2 == boundary
    for bad, same in (({3}, True), ({boundary}, False)):
        whole = run(0, n - 1, bad)
        stitched = compose_shard_poses([run(lo1, hi1 - 1, bad), run(lo2, hi2 - 1, bad)])
        assert len(stitched) == n
        diff = max(np.abs(a - b).max() for a, b in zip(stitched, whole))
        assert (diff < 1e-12) == same, (bad, diff)
        if not same:  # identical up to the boundary frame, different after it
            assert max(np.abs(a - b).max() for a, b in zip(stitched[:boundary + 1], whole[:boundary + 1])) < 1e-12
