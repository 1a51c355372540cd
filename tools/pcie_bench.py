#!/usr/bin/env python3
"""Host->device and device->host copy rate of this box with pinned memory (the ceiling of bench.py's e2e number):
one 467 MB copy (the bytes of 1000 KITTI frames), best of 10, CUDA events.  Prints one JSON line."""
import json

import torch

n = 1000 * 376 * 1241
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
res = {}
for name, (src, dst) in {"h2d": (h, d), "d2h": (d, h)}.items():
    best = 1e9
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        dst.copy_(src, non_blocking=True)
        e1.record()
        e1.synchronize()
        best = min(best, e0.elapsed_time(e1))
    res[name + "_GBps"] = round(n / best / 1e6, 2)
    res[name + "_ms"] = round(best, 3)
# both directions at once, on two streams
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
d2 = torch.empty(n // 8, dtype=torch.uint8, device="cuda")
h2 = torch.empty(n // 8, dtype=torch.uint8).pin_memory()
best = 1e9
for _ in range(10):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(s1):
        e0.record()
        d.copy_(h, non_blocking=True)
        e1.record()
    with torch.cuda.stream(s2):
        h2.copy_(d2, non_blocking=True)
    torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
res["h2d_GBps_with_concurrent_d2h"] = round(n / best / 1e6, 2)
print(json.dumps(res))
