#!/usr/bin/env python3
"""bench.py — frames/s of the feature front-end hot path (detect + FREAK + match) on synthetic frames, 1..8 B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--config 1|2|3|4]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

--config picks the BASELINE.json configuration (default 1, the one the metric is quoted on):
  1  configs[1]  KITTI-shaped 1241x376, 1000 frames/GPU, VisualOdometry preset (20000/3/2000), REF32 matching
  2  configs[2]  1920x1080, 8000 keypoints/frame, FULL512 matching: matcher-bound, roofline = POPC32/s against the
                 POPC32 issue rate measured in the same process (vo_bench_popc)
  3  configs[3]  3840x2160, the stencil kernel swept over frames per launch, roofline = the stencil at batch 64
  4  configs[4]  10 000 KITTI-shaped frames sharded by frame pairs over the GPUs, each rank running the pipelined
                 VisualOdometry::run driver (decode -> front end -> host RANSAC + pose), rank 0 stitching the trajectory

A step = one pass of the whole hot path over this rank's frame sequence (every rank has its own sequence: weak scaling,
no collective on the data path; NCCL is used only for the barrier and the max-over-ranks of the timing).
  value  frames/s with the frames already resident in HBM (kernels only, CUDA events on the launch stream)
  e2e    frames/s through the C-ABI call vo_sequence_run with HOST buffers: pinned-host -> device copies of every frame
         and device -> host copies of keypoints and match lists inside the timed region
  roofline          the dominant kernel of the step as it runs in the step.  Configs 1 and 3: the stencil kernel with
                    the fused NMS: 2 B/px (u8 in, blurred u8 out) + 8 B per candidate
  roofline_unfused  the response-writing form of the same kernel (vo_response / debug path), 6 B/px, timed alone
                    (vo_bench_stencil).  The two are labelled and never mixed.
  cpu_baseline      the reference's own CPU thread-pool code (oracle/_ref) on a bounded sample, same box
`--impl reference` times that CPU path alone, in the same units.
"""
import argparse
import json
import os
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

STENCIL_BYTES_PER_PX_UNFUSED = 6  # u8 in + blurred u8 out + fp32 R out
STENCIL_BYTES_PER_PX_FUSED = 2    # u8 in + blurred u8 out; + 8 B per NMS candidate
CAND_BYTES = 8

CONFIGS = {
    1: dict(name="configs[1]", what="synthetic KITTI-shaped grayscale sequence", rows=376, cols=1241, frames=1000, batch=1000,
            e2e_batch=100, preset="vo_gpu_ref", match_bits=32, max_kp=0, threshold=0.0, rects=None, cpu_sample=96, ref_frames=24),
    2: dict(name="configs[2]", what="synthetic 1080p sequence, 8000 keypoints/frame (matcher-bound)", rows=1080, cols=1920, frames=64,
            batch=16, e2e_batch=16, preset="vo_gpu_ref", match_bits=512, max_kp=8000, threshold=6000.0, rects=6000, cpu_sample=8,
            ref_frames=4),
    3: dict(name="configs[3]", what="synthetic 4K frames, stencil sweep over frames per launch", rows=2160, cols=3840, frames=64, batch=64,
            e2e_batch=16, preset="vo_gpu_ref", match_bits=32, max_kp=0, threshold=0.0, rects=20000, cpu_sample=4, ref_frames=2,
            sweep=[1, 2, 4, 8, 16, 32, 64]),
    4: dict(name="configs[4]", what="KITTI-shaped sequence sharded by frame pairs, host RANSAC + pose", rows=376, cols=1241, frames=10000,
            batch=64, preset="vo_gpu_ref", match_bits=32),
}


def measured_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock and throttle reasons sampled through NVML every 20 ms while the GPU is under the benchmark load
    (the timed regions last tens of milliseconds, far below nvidia-smi's useful sampling period)."""

    def __init__(self, gpu_index):
        self.gpu, self.samples, self._stop, self.thread = gpu_index, [], threading.Event(), None

    def _run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.gpu)
            mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            while not self._stop.is_set():
                sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                self.samples.append((sm, mx, int(r)))
                time.sleep(0.02)
            nv.nvmlShutdown()
        except Exception as e:  # NVML unavailable: report no samples rather than guess
            self.error = repr(e)

    def start(self):
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()

    def stop(self):
        self._stop.set()
        if self.thread:
            self.thread.join(2)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "error": getattr(self, "error", None)}
        bits = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}
        seen = 0
        for _, _, r in self.samples:
            seen |= r
        return {"sm_mhz": float(np.median([x[0] for x in self.samples])), "sm_max_mhz": float(self.samples[0][1]),
                "reasons": sorted(n for b, n in bits.items() if seen & b), "samples": len(self.samples)}


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def gpu_host_link(gpu_index):
    """PCIe link of this GPU and the NUMA node it hangs off (NVML + sysfs): the end-to-end path is bound by it."""
    info = {}
    try:
        import pynvml as nv
        nv.nvmlInit()
        h = nv.nvmlDeviceGetHandleByIndex(gpu_index)
        bus = nv.nvmlDeviceGetPciInfo(h).busId
        bus = (bus.decode() if isinstance(bus, bytes) else bus).lower()
        info = {"bus": bus, "pcie_gen": int(nv.nvmlDeviceGetCurrPcieLinkGeneration(h)), "pcie_width": int(nv.nvmlDeviceGetCurrPcieLinkWidth(h)),
                "pcie_gen_max": int(nv.nvmlDeviceGetMaxPcieLinkGeneration(h)), "pcie_width_max": int(nv.nvmlDeviceGetMaxPcieLinkWidth(h))}
        nv.nvmlShutdown()
        if len(bus.split(":")[0]) == 8:  # NVML prints an 8-digit PCI domain, sysfs a 4-digit one
            bus = bus[4:]
        info["numa_node"] = int(Path(f"/sys/bus/pci/devices/{bus}/numa_node").read_text())
    except Exception as e:
        info["error"] = type(e).__name__
    return info


def bind_to_gpu_numa(link):
    """Pin this rank (and so its pinned host buffers, first touched by it) to the CPUs of the NUMA node its GPU hangs
    off: with 8 ranks streaming 50 GB/s each, frames crossing the socket interconnect would cap the end-to-end rate.
    Returns a short description for the JSON line; any failure leaves the affinity untouched."""
    try:
        node = link.get("numa_node", -1)
        if node is None or node < 0:
            return "numa: single node"
        cpus = set()
        for part in Path(f"/sys/devices/system/node/node{node}/cpulist").read_text().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return f"numa: node {node} has no allowed cpus"
        os.sched_setaffinity(0, cpus)
        return f"numa: node {node}, {len(cpus)} cpus"
    except Exception as e:
        return f"numa: unbound ({type(e).__name__})"


def workload_config(cfg, args, frames, preset_params):
    return {"workload": f"{cfg['what']} {cfg['cols']}x{cfg['rows']}, {frames} frames/GPU, consecutive-pair matching "
                        f"(BASELINE.json {cfg['name']})",
            "frames_per_gpu": frames, "rows": cfg["rows"], "cols": cfg["cols"], "preset": preset_params["name"],
            "threshold": preset_params["threshold"], "nms_window": preset_params["nms_window"], "max_keypoints": preset_params["max_keypoints"],
            "match_bits": args.match_bits, "batch": args.batch, "seed": args.seed,
            "l2_policy": f"inputs larger than L2: {frames * cfg['rows'] * cfg['cols'] / 1e6:.0f} MB of frames per pass vs 126 MB L2",
            "parallelism": f"frame-pair shards x{args.gpus}, no collective"}


def cpu_blur_figures(frames):
    """the shim's exact-integer blur is scalar; OpenCV's own is SIMD.  Time both per frame on this box so
    that the CPU figure can also be read with the blur excluded or with cv2's blur in its place."""
    from oracle import pyoracle
    out = {}
    ref = pyoracle.Reference() if pyoracle.reference_available() else pyoracle.Oracle()
    k = min(len(frames), 8)
    t0 = time.perf_counter()
    for f in frames[:k]:
        ref.blur(f)
    out["blur_ms_per_frame_shim"] = 1000.0 * (time.perf_counter() - t0) / k
    try:
        import cv2
        cv2.setNumThreads(1)
        t0 = time.perf_counter()
        for f in frames[:k]:
            cv2.GaussianBlur(f, (7, 7), 0)
        out["blur_ms_per_frame_cv2"] = 1000.0 * (time.perf_counter() - t0) / k
    except Exception:
        out["blur_ms_per_frame_cv2"] = None
    return out
