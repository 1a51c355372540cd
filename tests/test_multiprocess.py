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
