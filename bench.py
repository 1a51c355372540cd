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


def cpu_reference_run(frames, warmup_s):
    """The reference's own CPU implementation of the path on all host threads (oracle/_ref/libvoref.so = its TUs compiled
    unmodified, default BLOCK_SIZE 100), else the single-threaded restatement.  Its parameters are the ones the reference
    hard-wires: threshold 25000, NMS 5, top 1500, REF32 matcher.  Returns (result dict, kind, cores)."""
    from oracle import pyoracle
    if pyoracle.reference_available():
        cores = pyoracle.host_cores()
        return pyoracle.Reference().bench_sequence(np.ascontiguousarray(frames), threads=cores, block_size=100, warmup_s=warmup_s), "reference", cores
    o = pyoracle.Oracle()
    t0 = time.perf_counter()
    feats = [o.extract(f, 25000.0, 5, 1500) for f in frames]
    te = time.perf_counter() - t0
    nm = sum(len(o.match(feats[k][1], feats[k + 1][1], 32)) for k in range(len(frames) - 1))
    return dict(extract_s=te, match_s=time.perf_counter() - t0 - te, keypoints=sum(len(f[0]) for f in feats), matches=nm), "port", 1


CPU_PARAMS = {"name": "cpu_ref (the reference's hard-wired parameters)", "threshold": 25000.0, "nms_window": 5, "max_keypoints": 1500}


def cpu_sample_text(s, r):
    return (f"{s} frames / {s - 1} pairs through the reference's own feature_extraction_manager_with_points + "
            f"matchCustomBinaryDescriptorsThreadPool (its hard-wired 25000/5/1500 parameters, BLOCK_SIZE 100, REF32 matcher, g++ -O2); "
            f"extract {1000 * r['extract_s'] / s:.2f} ms/frame, match {1000 * r['match_s'] / max(s - 1, 1):.2f} ms/pair, "
            f"{r['keypoints'] / s:.0f} keypoints/frame")


def reference_arm(cfg, args, rank, world):
    """`--impl reference`: each step a bounded sample of the config's workload on the host cores; rank 0 only."""
    if rank != 0:
        return
    from acs_visual_odometry_b200.synth import make_sequence
    sample = args.ref_frames
    frames = make_sequence(sample, cfg["rows"], cfg["cols"], seed=args.seed, n_rects=args.rects)
    kind, cores = None, None
    for w in range(args.warmup):
        cpu_reference_run(frames, 2.0 if w == 0 else 0.0)
    t, last = [], None
    for _ in range(args.steps):
        last, kind, cores = cpu_reference_run(frames, 0.0 if args.warmup else 2.0)
        t.append(last["extract_s"] + last["match_s"])
    ms = 1000.0 * float(np.mean(t))
    fps = sample / (ms / 1000.0)
    config = workload_config(cfg, args, sample, CPU_PARAMS)
    config["match_bits"] = 32
    line = {
        "impl": "reference", "metric": "frames/s (detect+FREAK+match)", "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8/i32/fp32", "data": "synthetic", "config": config,
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cores, "kind": kind, "sample": cpu_sample_text(sample, last) + " per step"},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def stencil_capture():
    """Counters of the committed ncu --set full captures of the stencil kernel (profiles/stencil_traffic.json): DRAM bytes
    and warp instructions per launch, per form, for one (batch, rows, cols)."""
    tp = ROOT / "profiles" / "stencil_traffic.json"
    return json.loads(tp.read_text()) if tp.exists() else {}


def capture_applies(cap, nb, rows, cols, code_sha1):
    """(use, stale): the committed counters are quoted only for the launch shape AND the stencil code they were captured
    from (`stencil_code_sha1`, a comment-insensitive hash of stencil.cu); same shape but other code = stale."""
    shape = (cap.get("batch"), cap.get("rows"), cap.get("cols")) == (nb, rows, cols)
    code = cap.get("stencil_code_sha1") == code_sha1
    return shape and code, shape and not code


def stencil_rooflines(vo, params, cfg, nb, fused_ms_in_step, peak, peak_src, sm_mhz):
    """The two labelled rooflines of the stencil kernel over `nb` frames per launch."""
    px = cfg["rows"] * cfg["cols"] * nb
    unf_ms = vo.bench_stencil(params, nb, fused=False, reps=5)
    try:
        fused_alone_ms = vo.bench_stencil(params, nb, fused=True, reps=5)
        cands = vo.bench_candidates(nb)
    except Exception:  # windows 7 / 9 and tiny images have no fused form
        fused_alone_ms, cands = None, 0
    from acs_visual_odometry_b200.build import source_hash
    cap = stencil_capture()
    same, stale = capture_applies(cap, nb, cfg["rows"], cfg["cols"], source_hash("stencil.cu"))
    issue_peak = 148 * 4 * 32 * (sm_mhz or 1965.0) * 1e6  # lane-instructions/s: 4 schedulers per SM, one warp instruction per clock each

    def block(kind, label, bpp, extra, ms, timed):
        if not ms:
            return None
        alg = bpp * px + extra
        ach = alg / (ms / 1000.0) / 1e9
        b = {"kernel": label, "bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
             "algorithmic_bytes_per_launch": alg, "bytes_per_px": bpp, "avg_launch_ms": ms, "timed": timed, "peak_source": peak_src,
             "traffic": None}
        c = cap.get(kind) if same else None
        if stale:
            b["traffic_stale"] = True  # the committed ncu counters belong to other stencil code: not quoted
        if c:
            b["traffic"] = c["dram_bytes_per_launch"]
            b["traffic_source"] = cap.get("source")
            li = 32.0 * c["warp_instructions_per_launch"]
            b["issue"] = {"note": "the kernel is issue-bound, not HBM-bound: lane-instructions per second against 148 SMs x 4 schedulers x 32 lanes "
                                  "x the SM clock; instruction count from the committed ncu capture, duration measured in this run",
                          "lane_instr_per_px": li / px, "achieved_Tinstr_s": li / (ms / 1000.0) / 1e12, "peak_Tinstr_s": issue_peak / 1e12,
                          "frac": li / (ms / 1000.0) / issue_peak}
        return b

    fused = block("fused", "k_stencil, fused-NMS form (blur + gradients + structure tensor + Shi-Tomasi response + NMS candidates): "
                  "2 B/px + 8 B per candidate", STENCIL_BYTES_PER_PX_FUSED, CAND_BYTES * cands, fused_ms_in_step or fused_alone_ms,
                  "CUDA events around the launches of the profiled step" if fused_ms_in_step else "alone (vo_bench_stencil mode 1)")
    if fused:
        fused["candidates_per_launch"] = cands
        fused["alone_ms"] = fused_alone_ms
    unfused = block("unfused", "k_stencil, response-writing form (blur + gradients + structure tensor + Shi-Tomasi response -> fp32 R): 6 B/px",
                    STENCIL_BYTES_PER_PX_UNFUSED, 0, unf_ms, "alone (vo_bench_stencil mode 0), not part of the step")
    return fused, unfused


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", type=int, default=1, choices=sorted(CONFIGS))
    ap.add_argument("--frames", type=int, default=None, help="frames per GPU (default: the config's)")
    ap.add_argument("--batch", type=int, default=None, help="frames per launch of the resident path (library limit 1024). KITTI, per 1000 "
                    "frames: batch 250 (4 launches on two alternating streams) 3.19 ms, 500 3.06 ms, 1000 (one launch per kernel) 3.04 ms")
    ap.add_argument("--e2e-batch", type=int, default=None, help="frames per pipelined chunk of vo_sequence_run (copy-bound: 50 and 100 give "
                    "9.04 ms per 1000 KITTI frames, 200 gives 9.36, 250 gives 9.56 - the last chunk is not overlapped)")
    ap.add_argument("--preset", default=None, choices=["vo_gpu_ref", "cpu_ref"])
    ap.add_argument("--match-bits", type=int, default=None, choices=[32, 512])
    ap.add_argument("--rects", type=int, default=None)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--ref-frames", type=int, default=None, help="frames per step of the CPU reference arm")
    ap.add_argument("--cpu-sample", type=int, default=None, help="frames of the cpu_baseline sample (rank 0, N=1)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--pipeline-fmt", default="pgm", choices=["pgm", "png"], help="config 4: frame files on disk")
    args = ap.parse_args()
    cfg = CONFIGS[args.config]
    for k in ("frames", "batch", "e2e_batch", "preset", "match_bits", "rects", "ref_frames", "cpu_sample"):
        if getattr(args, k) is None:
            setattr(args, k, cfg.get(k))
    args.warmup = max(args.warmup, 0)
    rank, world, local = dist_env()

    if args.config == 4:
        return pipeline_config(cfg, args, rank, world, local)
    if args.impl == "reference":
        return reference_arm(cfg, args, rank, world)

    import torch
    import torch.distributed as dist
    from acs_visual_odometry_b200 import Params, VoFront
    from acs_visual_odometry_b200.api import PinnedBuffer
    from acs_visual_odometry_b200.synth import make_sequence

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: the front-end has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    link = gpu_host_link(local)
    numa = bind_to_gpu_numa(link)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() not in ("INFO", "TRACE"):
            os.environ["NCCL_DEBUG"] = "WARN"  # keep stdout to the one JSON line (NCCL_DEBUG=VERSION prints a banner there)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    W = max(args.warmup, 3)
    ROWS, COLS = cfg["rows"], cfg["cols"]

    def make_params(preset):
        p = Params.vo_gpu_ref() if preset == "vo_gpu_ref" else Params.cpu_ref()
        p.match_bits = args.match_bits
        if cfg["max_kp"]:
            p.max_keypoints = cfg["max_kp"]
        if cfg["threshold"]:
            p.threshold = cfg["threshold"]
        return p

    params = make_params(args.preset)
    pdesc = {"name": args.preset, "threshold": params.threshold, "nms_window": params.nms_window, "max_keypoints": params.max_keypoints}
    n = args.frames
    vo = VoFront(ROWS, COLS, params.max_keypoints, 2 * max(args.batch, args.e2e_batch), device=local)
    pinned = PinnedBuffer((n, ROWS, COLS), np.uint8)
    frames = make_sequence(n, ROWS, COLS, seed=args.seed + 1000 * rank, n_rects=args.rects, out=pinned.array)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def all_ranks(v):
        """The value of every rank, in rank order (so the slow rank can be named)."""
        if world == 1:
            return [float(v)]
        t = torch.zeros(world, dtype=torch.float64, device="cuda")
        t[rank] = v
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return [float(x) for x in t.tolist()]

    def resident_pass(p, steps):
        for _ in range(W):
            vo.sequence_run_resident(p, args.batch)
        vo.sync()
        barrier()
        vo.launch_count(reset=True)
        vo.timer_start()
        for _ in range(steps):
            vo.sequence_run_resident(p, args.batch)
        ms = vo.timer_stop() / steps
        launches = vo.launch_count()
        barrier()
        return ms, launches

    mk = params.max_keypoints
    out = (PinnedBuffer((n,), np.int32), PinnedBuffer((n, mk, 2), np.float32), PinnedBuffer((max(n - 1, 1),), np.int32),
           PinnedBuffer((max(n - 1, 1), mk, 2), np.int32))
    out_arrays = tuple(b.array for b in out)

    def e2e_pass(p, steps):
        for _ in range(W):
            vo.sequence_run(frames, p, args.e2e_batch, out=out_arrays)
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            vo.sequence_run(frames, p, args.e2e_batch, out=out_arrays)   # synchronous: returns when results are on the host
        ms = 1000.0 * (time.perf_counter() - t0) / steps
        barrier()
        return ms

    # ---- value: frames resident in HBM, kernels only, CUDA events on the library's launch stream
    vo.sequence_upload(frames)
    sampler = ClockSampler(local)
    sampler.start()
    ms_mine, launches = resident_pass(params, args.steps)
    ms_ranks = all_ranks(ms_mine)
    ms_step = max(ms_ranks)
    nk, _, nm, _ = vo.sequence_fetch(want_kp=False, want_pairs=False)
    value = world * n / (ms_step / 1000.0)

    # ---- e2e: host frames -> vo_sequence_run -> host keypoints + match lists, copies inside the timed region
    e2e_ranks = all_ranks(e2e_pass(params, args.steps))
    e2e_ms = max(e2e_ranks)
    clocks = sampler.stop()   # sampled from the first warm-up pass to the end of the end-to-end passes
    assert np.array_equal(out_arrays[0], nk) and np.array_equal(out_arrays[2][:n - 1], nm), "e2e and resident paths disagree"
    # the floor of the end-to-end step: a bare copy of the same pinned frames to the device (best of 5, CUDA events),
    # first on every rank at the same time (what the step competes with), then - N=1 only - it is the same thing
    src = torch.from_numpy(frames.reshape(-1))
    dst = torch.empty(src.numel(), dtype=torch.uint8, device="cuda")

    def bare_copy():
        best = 1e30
        for _ in range(5):
            barrier()
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record()
            dst.copy_(src, non_blocking=True)
            ev1.record()
            ev1.synchronize()
            best = min(best, ev0.elapsed_time(ev1))
        return best
    bare_ranks = all_ranks(bare_copy())
    del dst
    e2e_value = world * n / (e2e_ms / 1000.0)
    h2d = n * ROWS * COLS
    d2h = 4 * n + 8 * n * mk + 4 * (n - 1) + 8 * (n - 1) * mk

    # ---- per-kernel device time (separate profiled pass: events around every launch)
    P = max(1, min(args.steps, 3))
    vo.profile_enable(True)
    vo.profile_read(reset=True)
    for _ in range(P):
        vo.sequence_run_resident(params, args.batch)
    prof = vo.profile_read(reset=True)
    vo.profile_enable(False)
    peak, peak_src = measured_peaks()
    kernel_ms = {k: round(v[0] / P, 4) for k, v in prof.items()}
    tot_ms = sum(kernel_ms.values()) or 1.0
    nb = min(args.batch, n)
    st_ms, st_n = prof["stencil"]
    fused, unfused = stencil_rooflines(vo, params, cfg, nb, st_ms / max(st_n, 1) if nb == n or n % nb == 0 else None, peak, peak_src,
                                       clocks.get("sm_mhz"))
    compares = float(np.sum(nk[:-1].astype(np.float64) * nk[1:]))
    popc_per_compare = 16 if args.match_bits == 512 else 1
    m_ms = prof["match"][0] / P
    popc_peak = vo.bench_popc(0, reps=3)
    matcher = {"kernel": "k_match (FULL512)" if args.match_bits == 512 else "k_match_ref32 (REF32)", "bound": "popc-pipe",
               "achieved": compares * popc_per_compare / (m_ms / 1000.0) / 1e9 if m_ms else 0.0, "peak": popc_peak / 1e9, "unit": "GPOPC32/s",
               "frac": (compares * popc_per_compare / (m_ms / 1000.0)) / popc_peak if m_ms and popc_peak else None,
               "compares_per_launch_set": compares, "popc32_per_compare": popc_per_compare, "ms_per_step": m_ms,
               "peak_source": "vo_bench_popc mode 0 in this process (independent POPC chains, best of 3 launches)",
               "compares_per_s": compares / (m_ms / 1000.0) if m_ms else 0.0}
    roofline = matcher if args.config == 2 else (fused or unfused)
    roofline = dict(roofline)
    roofline["share_of_step"] = (kernel_ms["match"] if args.config == 2 else kernel_ms["stencil"]) / tot_ms

    line = {
        "metric": "frames/s (detect+FREAK+match)", "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
        "warmup": W, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8/i32/fp32", "data": "synthetic", "config": workload_config(cfg, args, n, pdesc),
        "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": world * h2d, "d2h_bytes_per_step": world * d2h, "ms_per_step": e2e_ms,
                "batch": args.e2e_batch, "ms_per_step_by_rank": [round(x, 3) for x in e2e_ranks],
                "bare_h2d_copy_ms_by_rank": [round(x, 3) for x in bare_ranks],
                "floor_frames_per_s": world * n / (max(bare_ranks) / 1000.0), "frac_of_floor": max(bare_ranks) / e2e_ms,
                "note": "copy-bound: bare_h2d_copy_ms is one cudaMemcpyAsync of the step's pinned frames alone, all ranks copying at the same "
                        "time (barrier before every repetition, best of 5, CUDA events); floor = frames / the slowest rank's bare copy",
                "timing": "wall clock around the synchronous C-ABI call vo_sequence_run, max over ranks"},
        "gpu_launches": int(launches), "clocks": clocks, "host_binding": numa, "host_link": link,
        "ms_per_step_by_rank": [round(x, 4) for x in ms_ranks], "sm_mhz_by_rank": all_ranks(clocks.get("sm_mhz") or 0.0),
        "roofline": roofline, "roofline_unfused": unfused, "roofline_matcher": matcher if args.config != 2 else None,
        "kernel_ms_per_step": kernel_ms,
        "keypoints_per_frame": float(np.mean(nk)), "keypoints_per_frame_by_rank": [round(x, 1) for x in all_ranks(float(np.mean(nk)))],
        "compares_per_rank": [round(x) for x in all_ranks(compares)], "matches_per_pair": float(np.mean(nm)) if len(nm) else 0.0,
        "compares_per_s": matcher["compares_per_s"] * world, "compares_per_s_note": "matcher kernel time only (x n_gpus)",
        "match_mode": "REF32" if args.match_bits == 32 else "FULL512",
    }
    if fused and args.config != 2:
        line["roofline_fused_alone_ms"] = fused.get("alone_ms")
    if cfg.get("sweep"):
        sweep = []
        for b in cfg["sweep"]:
            if b > n:
                continue
            u = vo.bench_stencil(params, b, fused=False, reps=5)
            f = vo.bench_stencil(params, b, fused=True, reps=5)
            c = vo.bench_candidates(b)
            px = ROWS * COLS * b
            sweep.append({"frames_per_launch": b, "unfused_ms": round(u, 4), "unfused_GBs": round(6.0 * px / u / 1e6, 1), "unfused_frac": round(6.0 * px / u / 1e6 / peak, 4),
                          "fused_ms": round(f, 4), "fused_GBs": round((2.0 * px + 8.0 * c) / f / 1e6, 1), "fused_frac": round((2.0 * px + 8.0 * c) / f / 1e6 / peak, 4)})
        line["stencil_sweep"] = sweep

    if args.config == 1 and world == 1 and args.preset != "cpu_ref":
        # like for like with the CPU arm: the reference's hard-wired parameters (25000 / 5 / 1500) on the GPU
        pc = make_params("cpu_ref")
        ms_c, _ = resident_pass(pc, max(2, min(args.steps, 3)))
        nk_c = vo.sequence_fetch(want_kp=False, want_pairs=False)[0]
        e2e_c = e2e_pass(pc, 2)
        line["like_for_like"] = {"preset": "cpu_ref", "threshold": pc.threshold, "nms_window": pc.nms_window, "max_keypoints": pc.max_keypoints,
                                 "value": n / (ms_c / 1000.0), "ms_per_step": ms_c, "e2e_value": n / (e2e_c / 1000.0), "e2e_ms_per_step": e2e_c,
                                 "keypoints_per_frame": float(np.mean(nk_c))}

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        s = min(args.cpu_sample, n)
        r, kind, cores = cpu_reference_run(frames[:s], 2.0)
        cpu_fps = s / (r["extract_s"] + r["match_s"])
        blur = cpu_blur_figures(frames[:s])
        cb = {"value": cpu_fps, "unit": "frames/s", "cores": cores, "kind": kind, "sample": "first " + cpu_sample_text(s, r)}
        cb.update(blur)
        per_frame = (r["extract_s"] + r["match_s"]) / s
        cb["value_blur_excluded"] = 1.0 / max(per_frame - blur["blur_ms_per_frame_shim"] / 1000.0, 1e-9)
        if blur.get("blur_ms_per_frame_cv2") is not None:
            cb["value_with_cv2_blur"] = 1.0 / max(per_frame - (blur["blur_ms_per_frame_shim"] - blur["blur_ms_per_frame_cv2"]) / 1000.0, 1e-9)
        line["cpu_baseline"] = cb
        line["vs_cpu_baseline"] = {"resident": value / cpu_fps, "e2e": e2e_value / cpu_fps,
                                   "like_for_like_resident": line["like_for_like"]["value"] / cpu_fps if "like_for_like" in line else None,
                                   "like_for_like_e2e": line["like_for_like"]["e2e_value"] / cpu_fps if "like_for_like" in line else None,
                                   "e2e_blur_excluded": e2e_value / cb["value_blur_excluded"]}
    if rank == 0:
        print(json.dumps(line), flush=True)
    for b in out:
        b.free()
    vo.close()
    pinned.free()
    if world > 1:
        dist.destroy_process_group()


def pipeline_config(cfg, args, rank, world, local):
    """configs[4]: the whole pipeline, host RANSAC + pose included, sharded by frame pairs (tools/pipeline_shard_bench.py)."""
    if args.impl == "reference":
        if rank == 0:  # ransac.cpp needs Eigen and VisualOdometry.cpp OpenCL + OpenCV: neither can be built in this image
            print(json.dumps({"impl": "reference", "config": {"workload": cfg["what"]},
                              "unavailable": "the reference's VisualOdometry/RANSAC need Eigen, OpenCL and OpenCV (absent): use --config 1 for the CPU arm of the front end"}), flush=True)
        return
    sys.path.insert(0, str(ROOT / "tools"))
    import torch.distributed as dist
    from pipeline_shard_bench import run_shard_pipeline
    if world > 1:
        dist.init_process_group("gloo")
    link = gpu_host_link(local)
    sampler = ClockSampler(local)
    sampler.start()
    res = run_shard_pipeline(args.frames, args.pipeline_fmt, args.batch, 0, "", rank, world, local, dist if world > 1 else None)
    clocks = sampler.stop()
    if rank == 0:
        line = {"metric": "frames/s (decode + detect + FREAK + match + RANSAC + pose, trajectory CSV)", "value": res["e2e_frames_per_s"],
                "unit": "frames/s", "n_gpus": world, "steps": 1, "warmup": 1, "ms_per_step": 1000.0 * res["wall_s_max_over_ranks"],
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u8/i32/fp32/f64", "data": "synthetic",
                "config": {"workload": res["workload"] + " (BASELINE.json configs[4])", "frames": args.frames, "batch": args.batch,
                           "format": args.pipeline_fmt, "parallelism": f"frame-pair shards x{world}, no collective on the data path (gloo gathers the pose lists)"},
                "e2e": {"value": res["e2e_frames_per_s"], "unit": "frames/s", "h2d_bytes_per_step": args.frames * cfg["rows"] * cfg["cols"],
                        "d2h_bytes_per_step": None, "note": "files on local disk -> trajectory; the front end's share is frontend_frames_per_s"},
                "clocks": clocks, "host_link": link, "pipeline": res,
                "clocks_note": "the GPU idles most of the run (host-bound): the median SM clock is an idle clock, the throttle reasons are what matter"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
