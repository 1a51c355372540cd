#!/usr/bin/env python3
"""bench.py — frames/s of the feature front-end hot path (detect + FREAK + match) on synthetic KITTI-shaped
frames (BASELINE.json config 2: 1241x376, 1000 frames, VisualOdometry preset), 1..8 B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A step = one pass of the whole hot path over this rank's frame sequence (every rank has its own
`--frames`-long sequence: weak scaling, no collective on the data path; NCCL is used only for the barrier
and the max-over-ranks of the timing).
  value  frames/s with the frames already resident in HBM (kernels only, CUDA events on the launch stream)
  e2e    frames/s through the C-ABI call vo_sequence_run with HOST buffers: pinned-host -> device copies
         of every frame and device -> host copies of keypoints and match lists inside the timed region
  roofline      the fused stencil kernel (blur+gradients+structure tensor+response): algorithmic bytes
                6 B/px (u8 in, blurred u8 out, fp32 R out) / its mean launch duration (CUDA events)
  cpu_baseline  the reference's own CPU thread-pool code (oracle/_ref) on a bounded sample, same box
`--impl reference` times that CPU path alone, in the same units.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

ROWS, COLS = 376, 1241
STENCIL_BYTES_PER_PX = 6


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


def bind_to_gpu_numa(gpu_index):
    """Pin this rank (and so its pinned host buffers, first touched by it) to the CPUs of the NUMA node its GPU hangs
    off: with 8 ranks streaming 50 GB/s each, frames crossing the socket interconnect would cap the end-to-end rate.
    Returns a short description for the JSON line; any failure leaves the affinity untouched."""
    try:
        import pynvml as nv
        nv.nvmlInit()
        bus = nv.nvmlDeviceGetPciInfo(nv.nvmlDeviceGetHandleByIndex(gpu_index)).busId
        bus = (bus.decode() if isinstance(bus, bytes) else bus).lower()
        nv.nvmlShutdown()
        if len(bus.split(":")[0]) == 8:  # NVML prints an 8-digit PCI domain, sysfs a 4-digit one
            bus = bus[4:]
        node = int(Path(f"/sys/bus/pci/devices/{bus}/numa_node").read_text())
        if node < 0:
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


def reference_arm(args, rank, world):
    """The reference's own CPU implementation (oracle/_ref/libvoref.so = its TUs compiled unmodified), all host
    threads, default BLOCK_SIZE=100; each step a bounded sample of the same workload."""
    if rank != 0:
        return
    from acs_visual_odometry_b200.synth import make_sequence
    from oracle import pyoracle
    cores = pyoracle.host_cores()
    sample = args.ref_frames
    frames = make_sequence(sample, ROWS, COLS, seed=args.seed, n_rects=args.rects)
    if pyoracle.reference_available():
        ref, kind = pyoracle.Reference(), "reference"
        run = lambda warm: ref.bench_sequence(frames, threads=cores, block_size=100, warmup_s=warm)
    else:  # the oracle always exists: single-threaded port of the same path
        o, kind, cores = pyoracle.Oracle(), "port", 1

        def run(warm):
            t0 = time.perf_counter()
            feats = [o.extract(f, 25000.0, 5, 1500) for f in frames]
            te = time.perf_counter() - t0
            nm = sum(len(o.match(feats[k][1], feats[k + 1][1], 32)) for k in range(sample - 1))
            return dict(extract_s=te, match_s=time.perf_counter() - t0 - te, keypoints=sum(len(f[0]) for f in feats), matches=nm)
    for w in range(args.warmup):
        run(2.0 if w == 0 else 0.0)
    t = []
    last = None
    for _ in range(args.steps):
        last = run(0.0 if args.warmup else 2.0)
        t.append(last["extract_s"] + last["match_s"])
    ms = 1000.0 * float(np.mean(t))
    fps = sample / (ms / 1000.0)
    line = {
        "impl": "reference", "metric": "frames/s (detect+FREAK+match)", "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8/i32/fp32", "data": "synthetic",
        "config": workload_config(args, sample),
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cores, "kind": kind,
                         "sample": f"{sample} frames / {sample - 1} pairs per step, reference defaults (threshold 25000, NMS 5, top 1500, "
                                   f"BLOCK_SIZE 100, REF32 matcher), g++ -O2; extract {1000 * last['extract_s'] / sample:.2f} ms/frame, "
                                   f"match {1000 * last['match_s'] / max(sample - 1, 1):.2f} ms/pair"},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def workload_config(args, frames):
    return {"workload": f"synthetic KITTI-shaped grayscale sequence {COLS}x{ROWS}, {frames} frames/GPU, consecutive-pair matching "
                        f"(BASELINE.json configs[1])",
            "frames_per_gpu": frames, "rows": ROWS, "cols": COLS, "preset": args.preset, "match_bits": args.match_bits,
            "batch": args.batch, "seed": args.seed,
            "l2_policy": f"inputs larger than L2: {frames * ROWS * COLS / 1e6:.0f} MB of frames per pass vs 126 MB L2",
            "parallelism": f"frame-pair shards x{args.gpus}, no collective"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--frames", type=int, default=1000)
    ap.add_argument("--batch", type=int, default=1000, help="frames per launch of the resident path (library limit 1024). Measured per 1000 frames: batch 250 (4 launches on two alternating streams) 3.19 ms, 500 3.06 ms, 1000 (one launch of each kernel, 11000 stencil tasks) 3.04 ms")
    ap.add_argument("--e2e-batch", type=int, default=100, help="frames per pipelined chunk of vo_sequence_run (copy-bound: 50 and 100 give 9.04 ms per 1000 frames, 200 gives 9.36, 250 gives 9.56 - the last chunk is not overlapped)")
    ap.add_argument("--preset", default="vo_gpu_ref", choices=["vo_gpu_ref", "cpu_ref"])
    ap.add_argument("--match-bits", type=int, default=32, choices=[32, 512])
    ap.add_argument("--rects", type=int, default=None)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--ref-frames", type=int, default=24, help="frames per step of the CPU reference arm")
    ap.add_argument("--cpu-sample", type=int, default=96, help="frames of the cpu_baseline sample (rank 0, N=1)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    rank, world, local = dist_env()

    if args.impl == "reference":
        reference_arm(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    from acs_visual_odometry_b200 import Params, VoFront
    from acs_visual_odometry_b200.api import PinnedBuffer
    from acs_visual_odometry_b200.synth import make_sequence

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: the front-end has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    numa = bind_to_gpu_numa(local)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() not in ("INFO", "TRACE"):
            os.environ["NCCL_DEBUG"] = "WARN"  # keep stdout to the one JSON line (NCCL_DEBUG=VERSION prints a banner there)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    W = max(args.warmup, 3)

    params = Params.vo_gpu_ref() if args.preset == "vo_gpu_ref" else Params.cpu_ref()
    params.match_bits = args.match_bits
    n = args.frames
    vo = VoFront(ROWS, COLS, params.max_keypoints, 2 * max(args.batch, args.e2e_batch), device=local)
    pinned = PinnedBuffer((n, ROWS, COLS), np.uint8)
    frames = make_sequence(n, ROWS, COLS, seed=args.seed + 1000 * rank, n_rects=args.rects, out=pinned.array)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- value: frames resident in HBM, kernels only, CUDA events on the library's launch stream
    vo.sequence_upload(frames)
    sampler = ClockSampler(local)
    sampler.start()
    for _ in range(W):
        vo.sequence_run_resident(params, args.batch)
    vo.sync()
    barrier()
    vo.launch_count(reset=True)
    vo.timer_start()
    for _ in range(args.steps):
        vo.sequence_run_resident(params, args.batch)
    ms_total = vo.timer_stop()
    launches = vo.launch_count()
    barrier()
    ms_step = max_over_ranks(ms_total / args.steps)
    nk, _, nm, _ = vo.sequence_fetch(want_kp=False, want_pairs=False)
    value = world * n / (ms_step / 1000.0)

    # ---- e2e: host frames -> vo_sequence_run -> host keypoints + match lists, copies inside the timed region
    mk = params.max_keypoints
    out = (PinnedBuffer((n,), np.int32), PinnedBuffer((n, mk, 2), np.float32), PinnedBuffer((max(n - 1, 1),), np.int32),
           PinnedBuffer((max(n - 1, 1), mk, 2), np.int32))
    out_arrays = tuple(b.array for b in out)
    for _ in range(W):
        vo.sequence_run(frames, params, args.e2e_batch, out=out_arrays)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        vo.sequence_run(frames, params, args.e2e_batch, out=out_arrays)   # synchronous: returns when results are on the host
    e2e_ms = max_over_ranks(1000.0 * (time.perf_counter() - t0) / args.steps)
    barrier()
    clocks = sampler.stop()   # sampled from the first warm-up pass to the end of the end-to-end passes
    # the floor of the end-to-end step: a bare copy of the same pinned frames to the device (best of 5, CUDA events)
    src = torch.from_numpy(frames.reshape(-1))
    dst = torch.empty(src.numel(), dtype=torch.uint8, device="cuda")
    bare_ms = 1e30
    for _ in range(5):
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        dst.copy_(src, non_blocking=True)
        ev1.record()
        ev1.synchronize()
        bare_ms = min(bare_ms, ev0.elapsed_time(ev1))
    del dst
    e2e_value = world * n / (e2e_ms / 1000.0)
    h2d = n * ROWS * COLS
    d2h = 4 * n + 8 * n * mk + 4 * (n - 1) + 8 * (n - 1) * mk
    assert np.array_equal(out_arrays[0], nk) and np.array_equal(out_arrays[2][:n - 1], nm), "e2e and resident paths disagree"

    # ---- per-kernel device time (separate profiled pass: events around every launch)
    vo.profile_enable(True)
    vo.profile_read(reset=True)
    for _ in range(max(1, min(args.steps, 3))):
        vo.sequence_run_resident(params, args.batch)
    prof = vo.profile_read(reset=True)
    vo.profile_enable(False)
    peak, peak_src = measured_peaks()
    st_ms, st_n = prof["stencil"]
    px_per_launch = ROWS * COLS * min(args.batch, n)
    total_frames_prof = n * max(1, min(args.steps, 3))
    st_bytes = STENCIL_BYTES_PER_PX * ROWS * COLS * total_frames_prof       # all stencil launches of the pass
    achieved = st_bytes / (st_ms / 1000.0) / 1e9 if st_ms > 0 else 0.0
    kernel_ms = {k: round(v[0] / max(1, min(args.steps, 3)), 4) for k, v in prof.items()}
    tot_ms = sum(kernel_ms.values()) or 1.0
    traffic, traffic_src = None, None   # DRAM bytes per launch from the committed ncu --set full capture of this batch size
    tp = ROOT / "profiles" / "stencil_traffic.json"
    if tp.exists():
        tj = json.loads(tp.read_text())
        if (tj.get("batch"), tj.get("rows"), tj.get("cols")) == (min(args.batch, n), ROWS, COLS):
            traffic, traffic_src = tj["dram_bytes_per_launch"], tj["source"]
    roofline = {"kernel": "k_stencil (blur+gradients+structure tensor+Shi-Tomasi response)", "bound": "hbm", "achieved": achieved,
                "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                "peak_source": peak_src,
                "algorithmic_bytes_per_launch": STENCIL_BYTES_PER_PX * px_per_launch,
                "avg_launch_ms": st_ms / max(st_n, 1), "share_of_step": kernel_ms["stencil"] / tot_ms}

    line = {
        "metric": "frames/s (detect+FREAK+match)", "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
        "warmup": W, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8/i32/fp32", "data": "synthetic", "config": workload_config(args, n),
        "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": world * h2d, "d2h_bytes_per_step": world * d2h, "ms_per_step": e2e_ms, "batch": args.e2e_batch,
                "bare_h2d_copy_ms": bare_ms, "note": "copy-bound: bare_h2d_copy_ms is one cudaMemcpyAsync of the step's frames alone",
                "timing": "wall clock around the synchronous C-ABI call vo_sequence_run, max over ranks"},
        "gpu_launches": int(launches), "clocks": clocks, "host_binding": numa, "roofline": roofline, "kernel_ms_per_step": kernel_ms,
        "keypoints_per_frame": float(np.mean(nk)), "matches_per_pair": float(np.mean(nm)) if len(nm) else 0.0,
        "compares_per_s": float(world * np.sum(nk[:-1].astype(np.float64) * nk[1:]) / (ms_step / 1000.0)),
        "match_mode": "REF32" if args.match_bits == 32 else "FULL512",
    }

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import pyoracle
        cores = pyoracle.host_cores()
        s = min(args.cpu_sample, n)
        if pyoracle.reference_available():
            r = pyoracle.Reference().bench_sequence(np.ascontiguousarray(frames[:s]), threads=cores, block_size=100, warmup_s=2.0)
            kind = "reference"
        else:
            o, kind, cores = pyoracle.Oracle(), "port", 1
            t0 = time.perf_counter()
            feats = [o.extract(f, 25000.0, 5, 1500) for f in frames[:s]]
            te = time.perf_counter() - t0
            for k in range(s - 1):
                o.match(feats[k][1], feats[k + 1][1], 32)
            r = dict(extract_s=te, match_s=time.perf_counter() - t0 - te)
        cpu_fps = s / (r["extract_s"] + r["match_s"])
        line["cpu_baseline"] = {"value": cpu_fps, "unit": "frames/s", "cores": cores, "kind": kind,
                                "sample": f"first {s} frames / {s - 1} pairs of the same sequence through the reference's own "
                                          f"feature_extraction_manager_with_points + matchCustomBinaryDescriptorsThreadPool "
                                          f"(its hard-wired 25000/5/1500 parameters, BLOCK_SIZE 100, g++ -O2); extract "
                                          f"{1000 * r['extract_s'] / s:.2f} ms/frame, match {1000 * r['match_s'] / max(s - 1, 1):.2f} ms/pair"}
    if rank == 0:
        print(json.dumps(line), flush=True)
    for b in out:
        b.free()
    vo.close()
    pinned.free()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
