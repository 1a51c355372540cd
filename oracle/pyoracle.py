"""ctypes bindings for the parity checkers — TEST INFRASTRUCTURE ONLY.

Importable only from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs.  Two libraries:

  liboracle.so        oracle/vo_oracle.c, the plain-C restatement of the reference hot path
  _ref/libvoref.so    the reference's own TUs compiled unmodified against oracle/shim
                      (built only where /root/reference exists; travels prebuilt to the GPU box)
  _ref/libvoref_serial.so  the reference's full-length matcher (matching_serial.cpp:24-77), same recipe
"""
import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")
_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")


def build(verbose=False):
    """Compile liboracle.so and (when the reference tree is present) _ref/libvoref.so."""
    r = subprocess.run(["make", "-C", str(HERE), "all"], capture_output=True, text=True)
    if verbose or r.returncode:
        print(r.stdout, r.stderr)
    if r.returncode:
        raise RuntimeError("oracle build failed")


def _load(path):
    if not path.exists():
        build()
    if not path.exists():
        return None
    return C.CDLL(str(path))


class Oracle:
    """The C restatement (vo_oracle.h)."""

    def __init__(self):
        lib = _load(HERE / "liboracle.so")
        if lib is None:
            raise RuntimeError("oracle/liboracle.so missing and could not be built")
        self.lib = lib
        lib.vo_oracle_bgr2gray.argtypes = [_u8p, C.c_int, C.c_int, _u8p]
        lib.vo_oracle_blur.argtypes = [_u8p, C.c_int, C.c_int, _u8p]
        lib.vo_oracle_gradients.argtypes = [_u8p, C.c_int, C.c_int, _f32p, _f32p, _f32p]
        lib.vo_oracle_response.argtypes = [_u8p, C.c_int, C.c_int, C.c_float, _f32p]
        lib.vo_oracle_response_harris.argtypes = [_u8p, C.c_int, C.c_int, C.c_float, C.c_float, _f32p]
        lib.vo_oracle_nms.argtypes = [_f32p, C.c_int, C.c_int, C.c_int, C.c_int, _f32p]
        lib.vo_oracle_nms.restype = C.c_int
        lib.vo_oracle_orientation.argtypes = [_u8p, C.c_int, C.c_int, C.c_float, C.c_float]
        lib.vo_oracle_orientation.restype = C.c_float
        lib.vo_oracle_describe.argtypes = [_u8p, C.c_int, C.c_int, _f32p, C.c_int, _u8p, _f32p]
        lib.vo_oracle_extract.argtypes = [_u8p, C.c_int, C.c_int, C.c_float, C.c_int, C.c_int, _f32p, _u8p, _f32p,
                                          _u8p, _f32p]
        lib.vo_oracle_extract.restype = C.c_int
        lib.vo_oracle_match.argtypes = [_u8p, C.c_int, _u8p, C.c_int, C.c_int, C.c_float, _i32p, _i32p]
        lib.vo_oracle_match.restype = C.c_int
        lib.vo_oracle_pack.argtypes = [_u8p, C.c_int, _u32p]
        lib.vo_oracle_check_recip_division.argtypes = [C.POINTER(C.c_long)]
        lib.vo_oracle_check_recip_division.restype = C.c_long
        lib.vo_oracle_check_split_weight.argtypes = [C.POINTER(C.c_long)]
        lib.vo_oracle_check_split_weight.restype = C.c_long

    def bgr2gray(self, bgr):
        bgr = np.ascontiguousarray(bgr, np.uint8)
        out = np.empty(bgr.shape[:2], np.uint8)
        self.lib.vo_oracle_bgr2gray(bgr, bgr.shape[0], bgr.shape[1], out)
        return out

    def blur(self, gray):
        gray = np.ascontiguousarray(gray, np.uint8)
        out = np.empty_like(gray)
        self.lib.vo_oracle_blur(gray, gray.shape[0], gray.shape[1], out)
        return out

    def gradients(self, blurred):
        blurred = np.ascontiguousarray(blurred, np.uint8)
        o = [np.empty(blurred.shape, np.float32) for _ in range(3)]
        self.lib.vo_oracle_gradients(blurred, blurred.shape[0], blurred.shape[1], *o)
        return o

    def response(self, blurred, threshold=25000.0):
        blurred = np.ascontiguousarray(blurred, np.uint8)
        out = np.empty(blurred.shape, np.float32)
        self.lib.vo_oracle_response(blurred, blurred.shape[0], blurred.shape[1], threshold, out)
        return out

    def response_harris(self, blurred, threshold, k):
        blurred = np.ascontiguousarray(blurred, np.uint8)
        out = np.empty(blurred.shape, np.float32)
        self.lib.vo_oracle_response_harris(blurred, blurred.shape[0], blurred.shape[1], threshold, k, out)
        return out

    def nms(self, R, nms_window=5, max_keypoints=1500):
        R = np.ascontiguousarray(R, np.float32)
        kp = np.empty((max(max_keypoints, 1), 2), np.float32)
        n = self.lib.vo_oracle_nms(R, R.shape[0], R.shape[1], nms_window, max_keypoints, kp)
        return kp[:n].copy()

    def orientation(self, blurred, x, y):
        blurred = np.ascontiguousarray(blurred, np.uint8)
        return self.lib.vo_oracle_orientation(blurred, blurred.shape[0], blurred.shape[1], x, y)

    def describe(self, blurred, kp):
        blurred = np.ascontiguousarray(blurred, np.uint8)
        kp = np.ascontiguousarray(kp, np.float32).reshape(-1, 2)
        n = kp.shape[0]
        desc = np.empty((n, 512), np.uint8)
        ang = np.empty(max(n, 1), np.float32)
        self.lib.vo_oracle_describe(blurred, blurred.shape[0], blurred.shape[1], kp, n, desc, ang)
        return desc, ang[:n]

    def extract(self, gray, threshold=25000.0, nms_window=5, max_keypoints=1500, stages=False):
        gray = np.ascontiguousarray(gray, np.uint8)
        rows, cols = gray.shape
        kp = np.empty((max(max_keypoints, 1), 2), np.float32)
        desc = np.empty((max(max_keypoints, 1), 512), np.uint8)
        ang = np.empty(max(max_keypoints, 1), np.float32)
        blurred = np.empty((rows, cols), np.uint8)
        R = np.empty((rows, cols), np.float32)
        n = self.lib.vo_oracle_extract(gray, rows, cols, threshold, nms_window, max_keypoints, kp, desc, ang,
                                       blurred, R)
        if stages:
            return kp[:n].copy(), desc[:n].copy(), ang[:n].copy(), blurred, R
        return kp[:n].copy(), desc[:n].copy()

    def match(self, d1, d2, match_bits=32, ratio=0.75, raw=False):
        d1 = np.ascontiguousarray(d1, np.uint8).reshape(-1, 512)
        d2 = np.ascontiguousarray(d2, np.uint8).reshape(-1, 512)
        n1, n2 = d1.shape[0], d2.shape[0]
        pairs = np.empty((max(n1, 1), 2), np.int32)
        rawbuf = np.empty((max(n1, 1), 3), np.int32)
        n = self.lib.vo_oracle_match(d1, n1, d2, n2, match_bits, ratio, pairs, rawbuf)
        if raw:
            return pairs[:n].copy(), rawbuf[:n1].copy() if n2 > 0 else rawbuf[:0].copy()
        return pairs[:n].copy()

    def check_recip_division(self):
        total = C.c_long()
        bad = self.lib.vo_oracle_check_recip_division(C.byref(total))
        return bad, total.value

    def check_split_weight(self):
        total = C.c_long()
        bad = self.lib.vo_oracle_check_split_weight(C.byref(total))
        return bad, total.value

    def pack(self, desc512):
        d = np.ascontiguousarray(desc512, np.uint8).reshape(-1, 512)
        out = np.empty((d.shape[0], 16), np.uint32)
        self.lib.vo_oracle_pack(d, d.shape[0], out)
        return out


class Reference:
    """The reference's own code (oracle/_ref/libvoref.so, see oracle/ref_driver.cpp)."""

    def __init__(self):
        lib = _load(HERE / "_ref" / "libvoref.so")
        if lib is None:
            raise RuntimeError("oracle/_ref/libvoref.so missing (needs /root/reference to build)")
        self.lib = lib
        lib.voref_extract.argtypes = [_u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _f32p, _u8p]
        lib.voref_extract.restype = C.c_int
        lib.voref_extract_bgr.argtypes = lib.voref_extract.argtypes
        lib.voref_extract_bgr.restype = C.c_int
        lib.voref_match.argtypes = [_u8p, C.c_int, _u8p, C.c_int, C.c_int, C.c_float, _i32p, C.c_int]
        lib.voref_match.restype = C.c_int
        lib.voref_bgr2gray.argtypes = [_u8p, C.c_int, C.c_int, _u8p]
        lib.voref_blur.argtypes = [_u8p, C.c_int, C.c_int, _u8p]
        lib.voref_gradients.argtypes = [_u8p, C.c_int, C.c_int, _f32p, _f32p, _f32p]
        lib.voref_response.argtypes = [_u8p, C.c_int, C.c_int, _f32p]
        lib.voref_nms.argtypes = [_f32p, C.c_int, C.c_int, C.c_int, C.c_int, _f32p]
        lib.voref_nms.restype = C.c_int
        if hasattr(lib, "voref_nms_ranked"):
            lib.voref_nms_ranked.argtypes = lib.voref_nms.argtypes
            lib.voref_nms_ranked.restype = C.c_int
        lib.voref_orientation.argtypes = [_u8p, C.c_int, C.c_int, C.c_float, C.c_float]
        lib.voref_orientation.restype = C.c_float
        lib.voref_describe.argtypes = [_u8p, C.c_int, C.c_int, _f32p, C.c_int, _u8p]
        lib.voref_bench_sequence.argtypes = [_u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double,
                                             C.POINTER(C.c_double), C.POINTER(C.c_double),
                                             C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        lib.voref_bench_sequence.restype = C.c_int

    def extract(self, gray, threads=1, block_size=1 << 20, cap=4096):
        """feature_extraction_manager_with_points on gray replicated to BGR (cv::imread default)."""
        gray = np.ascontiguousarray(gray, np.uint8)
        kp = np.empty((cap, 2), np.float32)
        desc = np.empty((cap, 512), np.uint8)
        n = self.lib.voref_extract(gray, gray.shape[0], gray.shape[1], threads, block_size, cap, kp, desc)
        if n < 0:
            raise RuntimeError("cap too small")
        return kp[:n].copy(), desc[:n].copy()

    def match(self, d1, d2, threads=1, ratio=0.75):
        d1 = np.ascontiguousarray(d1, np.uint8).reshape(-1, 512)
        d2 = np.ascontiguousarray(d2, np.uint8).reshape(-1, 512)
        pairs = np.empty((max(d1.shape[0], 1), 2), np.int32)
        n = self.lib.voref_match(d1, d1.shape[0], d2, d2.shape[0], threads, ratio, pairs, pairs.shape[0])
        return pairs[:n].copy()

    def bgr2gray(self, bgr):
        bgr = np.ascontiguousarray(bgr, np.uint8)
        out = np.empty(bgr.shape[:2], np.uint8)
        self.lib.voref_bgr2gray(bgr, bgr.shape[0], bgr.shape[1], out)
        return out

    def blur(self, gray):
        gray = np.ascontiguousarray(gray, np.uint8)
        out = np.empty_like(gray)
        self.lib.voref_blur(gray, gray.shape[0], gray.shape[1], out)
        return out

    def gradients(self, blurred):
        blurred = np.ascontiguousarray(blurred, np.uint8)
        o = [np.empty(blurred.shape, np.float32) for _ in range(3)]
        self.lib.voref_gradients(blurred, blurred.shape[0], blurred.shape[1], *o)
        return o

    def response(self, blurred):
        blurred = np.ascontiguousarray(blurred, np.uint8)
        out = np.empty(blurred.shape, np.float32)
        self.lib.voref_response(blurred, blurred.shape[0], blurred.shape[1], out)
        return out

    def nms(self, R, nms_window=5, max_keypoints=1500):
        R = np.ascontiguousarray(R, np.float32)
        kp = np.empty((max(max_keypoints, 1), 2), np.float32)
        n = self.lib.voref_nms(R, R.shape[0], R.shape[1], nms_window, max_keypoints, kp)
        return kp[:n].copy()

    def nms_ranked(self, R, nms_window=5, max_keypoints=1500):
        """The reference's own output order (heap order: descending (R, i, j)), before the manager's (y, x) sort."""
        R = np.ascontiguousarray(R, np.float32)
        kp = np.empty((max(max_keypoints, 1), 2), np.float32)
        n = self.lib.voref_nms_ranked(R, R.shape[0], R.shape[1], nms_window, max_keypoints, kp)
        return kp[:n].copy()

    def orientation(self, blurred, x, y):
        blurred = np.ascontiguousarray(blurred, np.uint8)
        return self.lib.voref_orientation(blurred, blurred.shape[0], blurred.shape[1], x, y)

    def describe(self, blurred, kp):
        blurred = np.ascontiguousarray(blurred, np.uint8)
        kp = np.ascontiguousarray(kp, np.float32).reshape(-1, 2)
        desc = np.empty((kp.shape[0], 512), np.uint8)
        self.lib.voref_describe(blurred, blurred.shape[0], blurred.shape[1], kp, kp.shape[0], desc)
        return desc

    def bench_sequence(self, frames, threads, block_size=100, warmup_s=2.0):
        """Time the reference thread-pool path over a (n, rows, cols) u8 sequence.
        Returns dict(extract_s, match_s, keypoints, matches)."""
        frames = np.ascontiguousarray(frames, np.uint8)
        n, rows, cols = frames.shape
        te, tm = C.c_double(), C.c_double()
        nk, nm = C.c_int64(), C.c_int64()
        rc = self.lib.voref_bench_sequence(frames.reshape(-1), n, rows, cols, threads, block_size, warmup_s,
                                           C.byref(te), C.byref(tm), C.byref(nk), C.byref(nm))
        if rc:
            raise RuntimeError("voref_bench_sequence failed")
        return dict(extract_s=te.value, match_s=tm.value, keypoints=nk.value, matches=nm.value)


class ReferenceSerial:
    """The reference's own full-length matcher (matching_serial.cpp:24-77 compiled in place into
    oracle/_ref/libvoref_serial.so, see oracle/ref_serial_driver.cpp): pins the FULL512 mode."""

    def __init__(self):
        lib = _load(HERE / "_ref" / "libvoref_serial.so")
        if lib is None:
            raise RuntimeError("oracle/_ref/libvoref_serial.so missing (needs /root/reference to build)")
        self.lib = lib
        lib.voref_serial_hamming.argtypes = [_u8p, _u8p, C.c_int]
        lib.voref_serial_hamming.restype = C.c_int
        lib.voref_serial_match.argtypes = [_u8p, C.c_int, _u8p, C.c_int, C.c_int, C.c_float, _i32p, C.c_int]
        lib.voref_serial_match.restype = C.c_int

    def hamming(self, a, b):
        a, b = np.ascontiguousarray(a, np.uint8).reshape(-1), np.ascontiguousarray(b, np.uint8).reshape(-1)
        return self.lib.voref_serial_hamming(a, b, len(a))

    def match(self, d1, d2, ratio=0.75):
        d1 = np.ascontiguousarray(d1, np.uint8).reshape(-1, 512)
        d2 = np.ascontiguousarray(d2, np.uint8).reshape(-1, 512)
        pairs = np.empty((max(d1.shape[0], 1), 2), np.int32)
        n = self.lib.voref_serial_match(d1, d1.shape[0], d2, d2.shape[0], 512, ratio, pairs, pairs.shape[0])
        return pairs[:n].copy()


def reference_serial_available():
    return (HERE / "_ref" / "libvoref_serial.so").exists() or Path("/root/reference/matching_serial.cpp").is_file()


def reference_available():
    return (HERE / "_ref" / "libvoref.so").exists() or Path("/root/reference/feature_extraction_parallel").is_dir()


def host_cores():
    return len(os.sched_getaffinity(0))
