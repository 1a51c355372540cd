"""ctypes binding of include/vofront.h plus Python-side mirrors of the two reference entry points.

Everything here forwards to libvofront.so; no image or descriptor arithmetic happens in Python.
"""
import ctypes as C
import os
from pathlib import Path

import numpy as np

MATCH_REF32 = 32
MATCH_FULL512 = 512
RESPONSE_SHI_TOMASI, RESPONSE_HARRIS = 0, 1
DBG_BLURRED, DBG_RESPONSE, DBG_ANGLES = 1, 2, 3
_ERR = {-1: "VO_ERR_BAD_ARG", -2: "VO_ERR_CUDA", -3: "VO_ERR_CAPACITY", -4: "VO_ERR_NO_DEVICE", -5: "VO_ERR_STATE"}

_PKG = Path(__file__).resolve().parent


def lib_path():
    """The in-tree library; VO_LIB names another build of the same sources (kernel tuning variants, tools/variant_bench.py)."""
    return Path(os.environ["VO_LIB"]) if os.environ.get("VO_LIB") else _PKG / "libvofront.so"


class VoError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"{_ERR.get(code, code)}: {msg}")
        self.code = code


class _CParams(C.Structure):
    _fields_ = [("threshold", C.c_float), ("nms_window", C.c_int), ("max_keypoints", C.c_int), ("ratio", C.c_float),
                ("match_bits", C.c_int), ("response", C.c_int), ("harris_k", C.c_float)]


class _CSeqResult(C.Structure):
    _fields_ = [("n_keypoints", C.c_void_p), ("kp_xy", C.c_void_p), ("n_matches", C.c_void_p), ("pairs", C.c_void_p)]


class Params:
    """vo_params: the literals the reference scatters over globals and call sites."""

    def __init__(self, threshold=25000.0, nms_window=5, max_keypoints=1500, ratio=0.75, match_bits=MATCH_REF32,
                 response=RESPONSE_SHI_TOMASI, harris_k=0.04):
        self.threshold, self.nms_window, self.max_keypoints = float(threshold), int(nms_window), int(max_keypoints)
        self.ratio, self.match_bits = float(ratio), int(match_bits)
        self.response, self.harris_k = int(response), float(harris_k)

    @classmethod
    def cpu_ref(cls, **kw):  # corner_detection_parallel.h:16, feature_extraction_parallel.cpp:274
        return cls(25000.0, 5, 1500, **kw)

    @classmethod
    def vo_gpu_ref(cls, **kw):  # corner_detection_parallel_GPU.h:25, feature_extraction_parallel_GPU.cpp:235
        return cls(20000.0, 3, 2000, **kw)

    def c(self):
        return _CParams(self.threshold, self.nms_window, self.max_keypoints, self.ratio, self.match_bits, self.response, self.harris_k)


_lib = None


def load_library():
    """dlopen libvofront.so; raises if it has not been built (no fallback of any kind)."""
    global _lib
    if _lib is not None:
        return _lib
    p = lib_path()
    if not p.exists():
        raise RuntimeError(f"{p} not found: build it with `python -m acs_visual_odometry_b200.build` "
                           "(the front-end has no CPU or PyTorch fallback)")
    lib = C.CDLL(str(p))
    vp, i, f, sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
    P = C.POINTER
    sig = {
        "vo_params_preset": (i, [P(_CParams), i]),
        "vo_create": (i, [P(vp), i, i, i, i, i]),
        "vo_create_ex": (i, [P(vp), i, i, i, i, i, i]),
        "vo_sequence_flags": (i, [vp, vp, i]),
        "vo_destroy": (None, [vp]),
        "vo_last_error": (C.c_char_p, [vp]),
        "vo_device_count": (i, []),
        "vo_sync": (i, [vp]),
        "vo_host_alloc": (vp, [sz]),
        "vo_host_free": (None, [vp]),
        "vo_extract": (i, [vp, vp, i, i, sz, P(_CParams), i]),
        "vo_extract_bgr": (i, [vp, vp, i, i, sz, P(_CParams), i]),
        "vo_extract_batch": (i, [vp, vp, i, i, i, sz, sz, P(_CParams), i]),
        "vo_frame_fetch": (i, [vp, i, P(i), vp, vp, vp]),
        "vo_debug_fetch": (i, [vp, i, i, vp]),
        "vo_blur": (i, [vp, vp, i, i, vp]),
        "vo_response": (i, [vp, vp, i, i, f, vp]),
        "vo_response_p": (i, [vp, vp, i, i, P(_CParams), vp]),
        "vo_bgr2gray": (i, [vp, vp, i, i, sz, vp]),
        "vo_gradients": (i, [vp, vp, i, i, sz, i, i, i, i, vp, vp, vp]),
        "vo_response_from_gradients": (i, [vp, vp, vp, vp, i, i, i, i, i, i, P(_CParams), vp]),
        "vo_nms": (i, [vp, vp, i, i, i, i, vp, P(i)]),
        "vo_describe": (i, [vp, vp, i, i, vp, i, vp, vp]),
        "vo_match": (i, [vp, i, i, f, i, vp, i, P(i)]),
        "vo_match_host": (i, [vp, vp, i, vp, i, i, f, i, vp, i, P(i)]),
        "vo_match_raw": (i, [vp, vp, i]),
        "vo_match_static": (i, [vp, i, vp, i, f, i, vp, vp, vp]),  # no context, no allocation: device pointers in and out
        "vo_sequence_run": (i, [vp, vp, i, i, i, sz, sz, P(_CParams), i, P(_CSeqResult)]),
        "vo_sequence_upload": (i, [vp, vp, i, i, i, sz, sz]),
        "vo_sequence_run_resident": (i, [vp, P(_CParams), i]),
        "vo_sequence_fetch": (i, [vp, P(_CSeqResult)]),
        "vo_timer_start": (i, [vp]),
        "vo_timer_stop": (i, [vp, P(f)]),
        "vo_bench_stencil": (i, [vp, P(_CParams), i, i, i, P(f)]),
        "vo_bench_popc": (i, [vp, i, i, P(C.c_double)]),
        "vo_bench_candidates": (i, [vp, i, P(C.c_longlong)]),
        "vo_launch_count": (C.c_longlong, [vp, i]),
        "vo_profile_enable": (i, [vp, i]),
        "vo_profile_read": (i, [vp, P(f), P(C.c_longlong), i]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)  # AttributeError here = the library does not export what the header declares
        fn.restype, fn.argtypes = res, args
    lib._vo_symbols = tuple(sig)
    _lib = lib
    return lib


def device_count():
    return load_library().vo_device_count()


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class PinnedBuffer:
    """numpy view over vo_host_alloc memory (page-locked, for asynchronous copies)."""

    def __init__(self, shape, dtype):
        self._lib = load_library()
        n = int(np.prod(shape)) * np.dtype(dtype).itemsize
        self._p = self._lib.vo_host_alloc(max(n, 1))
        if not self._p:
            raise MemoryError("vo_host_alloc failed")
        buf = (C.c_uint8 * max(n, 1)).from_address(self._p)
        self.array = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    def free(self):
        if self._p:
            self.array = None
            self._lib.vo_host_free(self._p)
            self._p = None


class VoFront:
    """One vo_ctx (one GPU)."""

    def __init__(self, max_rows, max_cols, max_keypoints=2000, n_slots=4, device=0, cand_cap=0):
        self.lib = load_library()
        self.max_rows, self.max_cols, self.max_keypoints, self.n_slots = max_rows, max_cols, max_keypoints, n_slots
        h = C.c_void_p()
        rc = self.lib.vo_create_ex(C.byref(h), device, max_rows, max_cols, max_keypoints, n_slots, cand_cap)
        if rc:
            raise VoError(rc, self.lib.vo_last_error(None).decode())
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.lib.vo_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _ck(self, rc):
        if rc:
            raise VoError(rc, self.lib.vo_last_error(self.h).decode())

    # ---- extraction
    def extract(self, gray, params, slot=0):
        gray = np.ascontiguousarray(gray, np.uint8)
        self._ck(self.lib.vo_extract(self.h, _ptr(gray), gray.shape[0], gray.shape[1], gray.strides[0],
                                     C.byref(params.c()), slot))

    def extract_bgr(self, bgr, params, slot=0):
        bgr = np.ascontiguousarray(bgr, np.uint8)
        self._ck(self.lib.vo_extract_bgr(self.h, _ptr(bgr), bgr.shape[0], bgr.shape[1], bgr.strides[0],
                                         C.byref(params.c()), slot))

    def extract_batch(self, frames, params, first_slot=0):
        frames = np.ascontiguousarray(frames, np.uint8)
        n, rows, cols = frames.shape
        self._ck(self.lib.vo_extract_batch(self.h, _ptr(frames), n, rows, cols, frames.strides[1], frames.strides[0],
                                           C.byref(params.c()), first_slot))

    def fetch(self, slot=0, want_desc512=True, want_packed=True):
        n = C.c_int()
        cap = self.max_keypoints
        kp = np.empty((cap, 2), np.float32)
        d512 = np.empty((cap, 512), np.uint8) if want_desc512 else None
        dp = np.empty((cap, 16), np.uint32) if want_packed else None
        self._ck(self.lib.vo_frame_fetch(self.h, slot, C.byref(n), _ptr(kp), _ptr(d512) if want_desc512 else None,
                                         _ptr(dp) if want_packed else None))
        k = n.value
        return kp[:k].copy(), (d512[:k].copy() if want_desc512 else None), (dp[:k].copy() if want_packed else None)

    def debug(self, slot, what, rows, cols, n=None):
        if what == DBG_BLURRED:
            out = np.empty((rows, cols), np.uint8)
        elif what == DBG_RESPONSE:
            out = np.empty((rows, cols), np.float32)
        else:
            out = np.empty(max(n, 1), np.float32)
        self._ck(self.lib.vo_debug_fetch(self.h, slot, what, _ptr(out)))
        return out if what != DBG_ANGLES else out[:n]

    # ---- stage entry points
    def blur(self, gray):
        gray = np.ascontiguousarray(gray, np.uint8)
        out = np.empty_like(gray)
        self._ck(self.lib.vo_blur(self.h, _ptr(gray), gray.shape[0], gray.shape[1], _ptr(out)))
        return out

    def response(self, blurred, threshold=25000.0):
        blurred = np.ascontiguousarray(blurred, np.uint8)
        out = np.empty(blurred.shape, np.float32)
        self._ck(self.lib.vo_response(self.h, _ptr(blurred), blurred.shape[0], blurred.shape[1], threshold, _ptr(out)))
        return out

    def response_p(self, blurred, params):
        blurred = np.ascontiguousarray(blurred, np.uint8)
        out = np.empty(blurred.shape, np.float32)
        self._ck(self.lib.vo_response_p(self.h, _ptr(blurred), blurred.shape[0], blurred.shape[1], C.byref(params.c()), _ptr(out)))
        return out

    def bgr2gray(self, bgr):
        bgr = np.ascontiguousarray(bgr, np.uint8)
        out = np.empty(bgr.shape[:2], np.uint8)
        self._ck(self.lib.vo_bgr2gray(self.h, _ptr(bgr), bgr.shape[0], bgr.shape[1], bgr.strides[0], _ptr(out)))
        return out

    def gradients(self, picture, interval=None, out=None):
        """direction_gradients_worker over interval = (r0, r1, c0, c1) (default: the whole image); `out` = (Jx, Jy, Jxy) to
        write into (zero matrices by default, like the manager's cv::Mat::zeros)."""
        picture = np.ascontiguousarray(picture, np.uint8)
        rows, cols = picture.shape
        r0, r1, c0, c1 = interval or (0, rows, 0, cols)
        jx, jy, jxy = out if out is not None else (np.zeros((rows, cols), np.float32) for _ in range(3))
        self._ck(self.lib.vo_gradients(self.h, _ptr(picture), rows, cols, picture.strides[0], r0, r1, c0, c1, _ptr(jx), _ptr(jy), _ptr(jxy)))
        return jx, jy, jxy

    def response_from_gradients(self, jx, jy, jxy, params, interval=None, out=None):
        jx, jy, jxy = (np.ascontiguousarray(a, np.float32) for a in (jx, jy, jxy))
        rows, cols = jx.shape
        r0, r1, c0, c1 = interval or (0, rows, 0, cols)
        R = out if out is not None else np.zeros((rows, cols), np.float32)
        self._ck(self.lib.vo_response_from_gradients(self.h, _ptr(jx), _ptr(jy), _ptr(jxy), rows, cols, r0, r1, c0, c1,
                                                     C.byref(params.c()), _ptr(R)))
        return R

    def nms(self, R, nms_window=5, max_keypoints=1500):
        R = np.ascontiguousarray(R, np.float32)
        kp = np.empty((max(max_keypoints, 1), 2), np.float32)
        n = C.c_int()
        self._ck(self.lib.vo_nms(self.h, _ptr(R), R.shape[0], R.shape[1], nms_window, max_keypoints, _ptr(kp), C.byref(n)))
        return kp[:n.value].copy()

    def describe(self, blurred, kp):
        blurred = np.ascontiguousarray(blurred, np.uint8)
        kp = np.ascontiguousarray(kp, np.float32).reshape(-1, 2)
        n = kp.shape[0]
        desc = np.empty((max(n, 1), 512), np.uint8)
        ang = np.empty(max(n, 1), np.float32)
        self._ck(self.lib.vo_describe(self.h, _ptr(blurred), blurred.shape[0], blurred.shape[1], _ptr(kp), n, _ptr(desc),
                                      _ptr(ang)))
        return desc[:n].copy(), ang[:n].copy()

    # ---- matching
    def match(self, query_slot, train_slot, ratio=0.75, match_bits=MATCH_REF32):
        cap = self.max_keypoints
        pairs = np.empty((cap, 2), np.int32)
        n = C.c_int()
        self._ck(self.lib.vo_match(self.h, query_slot, train_slot, ratio, match_bits, _ptr(pairs), cap, C.byref(n)))
        return pairs[:n.value].copy()

    def match_host(self, d1, d2, ratio=0.75, match_bits=MATCH_REF32, cap=None):
        d1 = np.ascontiguousarray(d1)
        d2 = np.ascontiguousarray(d2)
        if d1.dtype == np.uint32:
            row_bytes, d1, d2 = 64, d1.reshape(-1, 16), d2.reshape(-1, 16)
        else:
            row_bytes, d1, d2 = 512, d1.astype(np.uint8, copy=False).reshape(-1, 512), d2.astype(np.uint8, copy=False).reshape(-1, 512)
        n1, n2 = d1.shape[0], d2.shape[0]
        cap = max(n1, 1) if cap is None else cap
        pairs = np.empty((max(cap, 1), 2), np.int32)
        n = C.c_int()
        self._ck(self.lib.vo_match_host(self.h, _ptr(d1) if n1 else None, n1, _ptr(d2) if n2 else None, n2, row_bytes,
                                        ratio, match_bits, _ptr(pairs), cap, C.byref(n)))
        return pairs[:n.value].copy()

    def match_raw(self, n1):
        raw = np.empty((max(n1, 1), 3), np.int32)
        self._ck(self.lib.vo_match_raw(self.h, _ptr(raw), n1))
        return raw[:n1].copy()

    # ---- sequences
    def _seq_buffers(self, n_frames, want_kp, want_pairs, pinned=False):
        mk = self.max_keypoints
        alloc = (lambda s, d: np.empty(s, d))
        nk = alloc((n_frames,), np.int32)
        nm = alloc((max(n_frames - 1, 1),), np.int32)
        kp = alloc((n_frames, mk, 2), np.float32) if want_kp else None
        pr = alloc((max(n_frames - 1, 1), mk, 2), np.int32) if want_pairs else None
        res = _CSeqResult(_ptr(nk), _ptr(kp) if want_kp else None, _ptr(nm), _ptr(pr) if want_pairs else None)
        return nk, kp, nm, pr, res

    def sequence_run(self, frames, params, batch, want_kp=True, want_pairs=True, out=None):
        """Host frames in, host results out (pipelined H2D / kernels / D2H). frames: (n, rows, cols) u8."""
        n, rows, cols = frames.shape
        if out is None:
            nk, kp, nm, pr, res = self._seq_buffers(n, want_kp, want_pairs)
        else:
            nk, kp, nm, pr = out
            res = _CSeqResult(_ptr(nk), _ptr(kp) if kp is not None else None, _ptr(nm), _ptr(pr) if pr is not None else None)
        self._ck(self.lib.vo_sequence_run(self.h, frames.ctypes.data_as(C.c_void_p), n, rows, cols, frames.strides[1],
                                          frames.strides[0], C.byref(params.c()), batch, C.byref(res)))
        return nk, kp, nm[:max(n - 1, 0)], (pr[:max(n - 1, 0)] if pr is not None else None)

    def sequence_flags(self, n):
        """Per-frame status bits of the last sequence call (bit 0: candidate list overflow)."""
        f = np.zeros(n, np.int32)
        self._ck(self.lib.vo_sequence_flags(self.h, _ptr(f), n))
        return f

    def sequence_upload(self, frames):
        n, rows, cols = frames.shape
        self._ck(self.lib.vo_sequence_upload(self.h, frames.ctypes.data_as(C.c_void_p), n, rows, cols, frames.strides[1],
                                             frames.strides[0]))
        self._seq_n = n

    def sequence_run_resident(self, params, batch):
        self._ck(self.lib.vo_sequence_run_resident(self.h, C.byref(params.c()), batch))

    def sequence_fetch(self, want_kp=True, want_pairs=True):
        n = self._seq_n
        nk, kp, nm, pr, res = self._seq_buffers(n, want_kp, want_pairs)
        self._ck(self.lib.vo_sequence_fetch(self.h, C.byref(res)))
        return nk, kp, nm[:max(n - 1, 0)], (pr[:max(n - 1, 0)] if pr is not None else None)

    # ---- measurement
    def sync(self):
        self._ck(self.lib.vo_sync(self.h))

    def timer_start(self):
        self._ck(self.lib.vo_timer_start(self.h))

    def timer_stop(self):
        ms = C.c_float()
        self._ck(self.lib.vo_timer_stop(self.h, C.byref(ms)))
        return ms.value

    def bench_stencil(self, params, batch, fused, reps=5):
        """Mean ms per launch of the stencil kernel alone over the first `batch` resident frames (see vofront.h)."""
        ms = C.c_float()
        self._ck(self.lib.vo_bench_stencil(self.h, C.byref(params.c()), batch, int(bool(fused)), reps, C.byref(ms)))
        return ms.value

    def bench_candidates(self, batch):
        """Candidates appended by the last fused stencil launch over its first `batch` frames (see vofront.h)."""
        v = C.c_longlong()
        self._ck(self.lib.vo_bench_candidates(self.h, batch, C.byref(v)))
        return v.value

    def bench_popc(self, mode=0, reps=5):
        """POPC32 per second of this GPU (see vofront.h)."""
        v = C.c_double()
        self._ck(self.lib.vo_bench_popc(self.h, mode, reps, C.byref(v)))
        return v.value

    def launch_count(self, reset=False):
        return self.lib.vo_launch_count(self.h, int(reset))

    def profile_enable(self, on=True):
        self._ck(self.lib.vo_profile_enable(self.h, int(on)))

    def profile_read(self, reset=True):
        ms = (C.c_float * 8)()
        n = (C.c_longlong * 8)()
        self._ck(self.lib.vo_profile_read(self.h, ms, n, int(reset)))
        names = ["stencil", "nms", "select", "freak", "match", "finalize", "misc"]
        return {k: (ms[i], n[i]) for i, k in enumerate(names)}


# ---- mirrors of the reference entry points (same names, argument meaning and error behaviour) --------
def feature_extraction_manager_with_points(image, ctx, params=None, slot=0):
    """feature_extraction_manager_with_points(image, pool | GPU_settings)
    (feature_extraction_parallel.cpp:215 / feature_extraction_parallel_GPU.cpp:194).
    `ctx` (a VoFront) stands where the reference takes its thread_pool / GPU_settings.  A 3-channel image takes
    the CPU path's custom_bgr2gray first; a 1-channel image is used as is (the VisualOdometry path).
    Returns (descriptors: (n, 512) uint8 of 0/1, keypoints: (n, 2) float32 (x, y))."""
    params = params or (Params.cpu_ref() if image.ndim == 3 else Params.vo_gpu_ref())
    if image.ndim == 3:
        ctx.extract_bgr(image, params, slot)
    else:
        ctx.extract(image, params, slot)
    kp, d512, _ = ctx.fetch(slot, want_desc512=True, want_packed=False)
    return d512, kp


def matchCustomBinaryDescriptorsThreadPool(desc1, desc2, ctx, numThreads=0, ratioThreshold=0.75, match_bits=MATCH_REF32):
    """matchCustomBinaryDescriptorsThreadPool(desc1, desc2, pool, numThreads, ratioThreshold)
    (feature_matching_parallel.cpp:49-113).  numThreads is accepted and ignored.  Empty inputs give an
    empty list, like the reference (:57).  Returns (m, 2) int32 (queryIdx, trainIdx) ascending in queryIdx."""
    d1 = np.asarray(desc1, np.uint8).reshape(-1, 512)
    d2 = np.asarray(desc2, np.uint8).reshape(-1, 512)
    if d1.shape[0] == 0 or d2.shape[0] == 0:
        return np.empty((0, 2), np.int32)
    return ctx.match_host(d1, d2, ratioThreshold, match_bits)
