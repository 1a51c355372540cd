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
