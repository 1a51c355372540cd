#!/usr/bin/env python3
"""One-line summary of a tools/pipeline_shard_bench.py result file."""
import json
import sys

d = json.load(open(sys.argv[1]))
print(d["n_gpus"], "GPUs", d["host_threads_per_rank"], "threads/rank", d["trajectory_poses"], "poses; e2e", round(d["e2e_frames_per_s"]),
      "frames/s; front-end", round(d["frontend_frames_per_s_sum_over_ranks"]), "ransac+pose", round(d["ransac_pose_pairs_per_s_sum_over_ranks"]),
      "decode", round(d["decode_frames_per_s_sum_over_ranks"]), "wall", d["per_rank_wall_s"])
