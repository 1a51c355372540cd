#!/usr/bin/env python3
"""Generate tests/golden/*.npz by running the REFERENCE's own hot-path code (oracle/_ref/libvoref.so =
the reference translation units compiled unmodified against oracle/shim, see oracle/Makefile).

Run in the build container only (needs /root/reference):   python tests/golden/make_golden.py
The reference ships no golden vectors of its own; these are its outputs on
  * images/factory1.png + factory2.png (BASELINE.json config 1; copies kept next to this file),
  * two KITTI-shaped synthetic frames (acs_visual_odometry_b200.synth, seed 7),
  * matcher known-answer sets (random / tie-heavy / degenerate descriptor sets).
Reference settings: single strip (BLOCK_SIZE >= rows, the deterministic setting), thresholds and
NMS parameters as hard-wired in its CPU manager (25000 / 5 / 1500), ratio 0.75.
FULL512 match lists (*_full512_ref) come from the reference's own full-length matcher, matching_serial.cpp:24-77,
compiled in place into oracle/_ref/libvoref_serial.so (oracle/ref_serial_driver.cpp); the restatement's FULL512
answers (oracle/vo_oracle.c) are asserted equal to them here and in tests/test_oracle.py.
The two KITTI-shaped frames are also written as PNG (kitti0.png, kitti1.png) so that the golden comparison does not
depend on numpy's random stream.
"""
import hashlib
import sys
from pathlib import Path

import cv2
import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from acs_visual_odometry_b200.synth import make_sequence  # noqa: E402
from oracle.pyoracle import Oracle, Reference, ReferenceSerial  # noqa: E402

G = Path(__file__).resolve().parent
ref, orc, ser = Reference(), Oracle(), ReferenceSerial()


def full512(d1, d2):
    m = ser.match(d1, d2)
    assert np.array_equal(m, orc.match(d1, d2, 512)), "restatement FULL512 differs from matching_serial.cpp"
    return m


def sparse(R):
    idx = np.flatnonzero(R).astype(np.int32)
    return idx, R.reshape(-1)[idx].astype(np.float32)


def stage_dump(gray, tag, out):
    kp, desc = ref.extract(gray, threads=4)
    blurred = ref.blur(gray)
    R = ref.response(blurred)
    assert np.array_equal(kp, ref.nms(R, 5, 1500))
    assert np.array_equal(desc, ref.describe(blurred, kp))
    ang = np.array([ref.orientation(blurred, x, y) for x, y in kp], np.float32)
    ridx, rval = sparse(R)
    out[f"{tag}_kp"] = kp.astype(np.int16)
    out[f"{tag}_desc"] = orc.pack(desc)
    out[f"{tag}_angles"] = ang
    out[f"{tag}_blurred"] = blurred
    out[f"{tag}_R_idx"], out[f"{tag}_R_val"] = ridx, rval
    out[f"{tag}_kp_nms3_2000"] = ref.nms(R, 3, 2000).astype(np.int16)
    return kp, desc


def main():
    out = {}
    f1 = cv2.imread(str(G / "factory1.png"), cv2.IMREAD_GRAYSCALE)
    f2 = cv2.imread(str(G / "factory2.png"), cv2.IMREAD_GRAYSCALE)
    _, d1 = stage_dump(f1, "factory1", out)
    _, d2 = stage_dump(f2, "factory2", out)
    out["factory_matches_ref32"] = ref.match(d1, d2, threads=4)
    out["factory_matches_full512_ref"] = full512(d1, d2)
    np.savez_compressed(G / "factory.npz", **out)
    print("factory:", len(out["factory1_kp"]), len(out["factory2_kp"]), len(out["factory_matches_ref32"]))

    out = {}
    fr = make_sequence(2, seed=7)
    out["frames_sha256"] = np.frombuffer(hashlib.sha256(fr.tobytes()).digest(), np.uint8)
    for k in range(2):
        assert cv2.imwrite(str(G / f"kitti{k}.png"), fr[k], [cv2.IMWRITE_PNG_COMPRESSION, 9])
    _, d1 = stage_dump(fr[0], "kitti0", out)
    _, d2 = stage_dump(fr[1], "kitti1", out)
    del out["kitti1_blurred"]
    out["kitti_matches_ref32"] = ref.match(d1, d2, threads=4)
    out["kitti_matches_full512_ref"] = full512(d1, d2)
    np.savez_compressed(G / "kitti_synth.npz", **out)
    print("kitti:", len(out["kitti0_kp"]), len(out["kitti1_kp"]), len(out["kitti_matches_ref32"]))

    # matcher known-answer sets (descriptors as 0/1 bytes); REF32 answers from the reference matcher
    out = {}
    rng = np.random.default_rng(11)
    cases = {
        "random": (rng.integers(0, 2, (300, 512), dtype=np.uint8), rng.integers(0, 2, (257, 512), dtype=np.uint8)),
        "ties": (np.repeat(rng.integers(0, 2, (40, 512), dtype=np.uint8), 5, axis=0),
                 np.repeat(rng.integers(0, 2, (30, 512), dtype=np.uint8), 4, axis=0)),
        "n2_is_1": (rng.integers(0, 2, (17, 512), dtype=np.uint8), rng.integers(0, 2, (1, 512), dtype=np.uint8)),
        "n2_is_2": (rng.integers(0, 2, (17, 512), dtype=np.uint8), rng.integers(0, 2, (2, 512), dtype=np.uint8)),
    }
    near = rng.integers(0, 2, (200, 512), dtype=np.uint8)
    noisy = near.copy()
    flip = rng.random(noisy.shape) < 0.03
    noisy[flip] ^= 1
    cases["near_duplicates"] = (near, noisy[rng.permutation(200)])
    rng2 = np.random.default_rng(12)   # added later: its own stream, so the sets above keep their values
    noisier = near.copy()
    noisier[rng2.random(noisier.shape) < np.linspace(0.05, 0.45, 200)[:, None]] ^= 1   # d1/d2 on both sides of 0.75 in FULL512
    cases["noisy_duplicates"] = (near, noisier[rng2.permutation(200)])
    pre = rng2.integers(0, 2, (64, 512), dtype=np.uint8)
    tr = np.repeat(pre, 3, axis=0)
    tr[:, 32:] = rng2.integers(0, 2, (192, 480), dtype=np.uint8)   # equal 32-test prefixes, different tails: REF32 ties, FULL512 decides
    cases["prefix_ties"] = (pre, tr[rng2.permutation(192)])
    for name, (a, b) in cases.items():
        out[f"{name}_d1"], out[f"{name}_d2"] = orc.pack(a), orc.pack(b)
        out[f"{name}_ref32"] = ref.match(a, b, threads=3)
        out[f"{name}_full512_ref"] = full512(a, b)
        print(name, len(out[f"{name}_ref32"]), len(out[f"{name}_full512_ref"]))
    np.savez_compressed(G / "matcher_kat.npz", **out)


if __name__ == "__main__":
    main()
