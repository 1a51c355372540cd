#!/usr/bin/env python3
"""The unchanged-loop rate: tests/cpp/adapter_latency.cpp (the reference's frame loop on include/vofront_adapter.hpp) on
synthetic KITTI-shaped frames, with the SM clock sampled during the run.  One JSON line.

    python tools/adapter_latency.py [frames] [repeats]
"""
import json
import struct
import subprocess
import sys
import tempfile
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import __graft_entry__ as g  # noqa: E402
from acs_visual_odometry_b200.synth import make_sequence  # noqa: E402
from bench import ClockSampler  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 200
reps = sys.argv[2] if len(sys.argv) > 2 else "3"
exe = g.build_adapter_driver("adapter_latency")
frames = make_sequence(n, 376, 1241, seed=0)
with tempfile.TemporaryDirectory() as tmp:
    raw = Path(tmp) / "frames.raw"
    raw.write_bytes(struct.pack("<iii", n, 376, 1241) + frames.tobytes())
    s = ClockSampler(0)
    s.start()
    r = subprocess.run([str(exe), str(raw), reps], capture_output=True, text=True, timeout=600)
    clocks = s.stop()
if r.returncode:
    raise SystemExit(r.stderr)
out = json.loads(r.stdout.strip().splitlines()[-1])
out["clocks"] = clocks
print(json.dumps(out))
