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
