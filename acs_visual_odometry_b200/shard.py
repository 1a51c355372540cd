"""Frame-pair sharding across GPUs (one process per GPU, no collective on the data path).

VisualOdometry::run (VisualOdometry.cpp:68-100) consumes pairs (i-1, i).  The P = n_frames-1 pairs are split
into `world` contiguous ranges; rank r extracts frames [first_pair, last_pair+1] (so neighbouring shards share
one boundary frame, which both extract) and matches its own pairs.  Nothing crosses GPUs.
"""


def shard_pairs(n_frames, rank, world):
    """Half-open pair range [lo, hi) of `rank`; pair p = frames (p, p+1)."""
    pairs = max(n_frames - 1, 0)
    lo = pairs * rank // world
    hi = pairs * (rank + 1) // world
    return lo, hi


def shard_frames(n_frames, rank, world):
    """Half-open frame range [lo, hi) a rank must extract: its pairs plus the shared right boundary frame."""
    lo, hi = shard_pairs(n_frames, rank, world)
    if hi <= lo:
        return lo, lo
    return lo, hi + 1


FLIP_Z = (1.0, 1.0, -1.0, 1.0)  # diag of flipZ (PoseUpdate.cpp:7-11); its own inverse


def _flip(T):
    import numpy as np
    return np.diag(FLIP_Z) @ T


def compose_shard_poses(shard_poses):
    """Stitch per-shard pose lists into the whole trajectory.

    shard_poses[s] is the list of 4x4 poses a shard's driver wrote for its frames [lo_s, hi_s] (hi_s = lo_{s+1}: the
    shared boundary frame), in the CSV convention of VisualOdometry::run: line 0 is the start pose as it is, every
    later line is flipZ * T_curr (VisualOdometry.cpp:70,186).  Each shard starts from the identity; the only sequential
    step of the whole pipeline is the prefix product T_curr = T_curr * T_rel (VisualOdometry.cpp:184), applied here
    once per shard: G_{s+1} = G_s * T_local_s(last).
    Restriction: every shard restarts the reference's "last valid frame" at its own first frame.  If the LAST pair of a
    shard fails (fewer than 8 matches or inliers at the boundary frame), a sequential run would match the next frame
    against the last valid frame before the boundary, the sharded run against the boundary frame itself, and the two
    trajectories differ from there on (tools/pipeline_shard_bench.py counts such boundaries as boundary_pairs_failed).
    Returns the list of poses of frames 0 .. n-1 in the same CSV
    convention (n = sum of shard lengths - (shards - 1))."""
    import numpy as np
    out, G = [], np.eye(4)
    for s, poses in enumerate(shard_poses):
        poses = [np.asarray(P, np.float64).reshape(4, 4) for P in poses]
        if not poses:
            continue
        local = [poses[0]] + [_flip(P) for P in poses[1:]]  # un-flipped T_curr per frame
        first = 0 if not out else 1  # the boundary frame was already emitted by the previous shard
        for j in range(first, len(local)):
            T = G @ local[j]
            out.append(T if (s == 0 and j == 0) or not out else _flip(T))
        G = G @ local[-1]
    return out
