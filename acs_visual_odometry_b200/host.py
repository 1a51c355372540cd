"""ctypes binding of libvohost.so (include/vohost.h): the host consumers of the front end's match lists -
fundamental-matrix RANSAC, PoseUpdate::getPose, the KITTI pose / CSV formats, frame readers and the
pipelined VisualOdometry::run driver.  Used by the tests and tools; C++ callers include host/*.hpp."""
import ctypes as C
from pathlib import Path

import numpy as np

_LIB = None
_PKG = Path(__file__).resolve().parent

SYMBOLS = ["vo_host_fundamental", "vo_host_normalize_points", "vo_host_score", "vo_host_triangulate", "vo_host_sampson", "vo_host_ransac", "vo_host_get_pose", "vo_host_read_gt_line",
           "vo_host_format_pose_csv", "vo_host_euler", "vo_host_read_gray", "vo_host_write_png", "vo_host_write_pgm",
           "vo_pipeline_run", "vo_pipeline_run_ex", "vo_pipeline_release", "vo_host_last_error"]


class VoHostError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"vohost error {code}: {msg}")
        self.code = code


def load_host_library():
    global _LIB
    if _LIB is not None:
        return _LIB
    path = _PKG / "libvohost.so"
    if not path.exists():
        raise RuntimeError(f"{path} is missing: run `python -m acs_visual_odometry_b200.build`")
    C.CDLL(str(_PKG / "libvofront.so"), mode=C.RTLD_GLOBAL)
    lib = C.CDLL(str(path))
    missing = [s for s in SYMBOLS if not hasattr(lib, s)]
    if missing:
        raise RuntimeError(f"libvohost.so lacks {missing}")
    lib.vo_host_last_error.restype = C.c_char_p
    _LIB = lib
    return lib


def _ck(rc):
    if rc != 0:
        raise VoHostError(rc, load_host_library().vo_host_last_error().decode())


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _pts(p):
    p = np.ascontiguousarray(p, np.float64)
    assert p.ndim == 2 and p.shape[1] == 4
    return p


def fundamental(pts):
    pts = _pts(pts)
    F = np.empty(9)
    _ck(load_host_library().vo_host_fundamental(_dp(pts), len(pts), _dp(F)))
    return F.reshape(3, 3)


def normalize_points(pts):
    pts = _pts(pts)
    n1, n2, T1, T2 = np.empty((len(pts), 3)), np.empty((len(pts), 3)), np.empty(9), np.empty(9)
    _ck(load_host_library().vo_host_normalize_points(_dp(pts), len(pts), _dp(n1), _dp(n2), _dp(T1), _dp(T2)))
    return n1, n2, T1.reshape(3, 3), T2.reshape(3, 3)


def score(F, pts, threshold=1.0, force_scalar=False):
    pts = _pts(pts)
    F = np.ascontiguousarray(F, np.float64).reshape(9)
    flags = np.empty(len(pts), np.uint8)
    cnt = C.c_int()
    _ck(load_host_library().vo_host_score(_dp(F), _dp(pts), len(pts), C.c_double(threshold), int(force_scalar),
                                          flags.ctypes.data_as(C.POINTER(C.c_uint8)), C.byref(cnt)))
    return flags, cnt.value


def sampson(F, pts):
    pts = _pts(pts)
    F = np.ascontiguousarray(F, np.float64).reshape(9)
    err = np.empty(len(pts))
    _ck(load_host_library().vo_host_sampson(_dp(F), _dp(pts), len(pts), _dp(err)))
    return err


def ransac(pts, probability=0.99, threshold=1.0, num_threads=1, seed=1):
    pts = _pts(pts)
    F = np.empty(9)
    inl = np.empty((max(len(pts), 1), 4))
    n_in, it = C.c_int(), C.c_int()
    _ck(load_host_library().vo_host_ransac(_dp(pts), len(pts), C.c_double(probability), C.c_double(threshold), num_threads,
                                           C.c_uint64(seed), _dp(F), _dp(inl), C.byref(n_in), C.byref(it)))
    return F.reshape(3, 3), inl[:n_in.value].copy(), it.value


def get_pose(F, p1, p2, gt_step=1.0, K=None):
    F = np.ascontiguousarray(F, np.float64).reshape(9)
    p1 = np.ascontiguousarray(p1, np.float32)
    p2 = np.ascontiguousarray(p2, np.float32)
    Kp = None if K is None else _dp(np.ascontiguousarray(K, np.float64).reshape(9))
    R, t, votes = np.empty(9), np.empty(3), C.c_int()
    fp = C.POINTER(C.c_float)
    _ck(load_host_library().vo_host_get_pose(_dp(F), p1.ctypes.data_as(fp), p2.ctypes.data_as(fp), len(p1), C.c_double(gt_step), Kp,
                                             _dp(R), _dp(t), C.byref(votes)))
    return R.reshape(3, 3), t, votes.value


def triangulate(R, t, n1, n2, force_scalar=False):
    R = np.ascontiguousarray(R, np.float64).reshape(9)
    t = np.ascontiguousarray(t, np.float64).reshape(3)
    n1 = np.ascontiguousarray(n1, np.float32)
    n2 = np.ascontiguousarray(n2, np.float32)
    X = np.empty((len(n1), 4))
    fp = C.POINTER(C.c_float)
    _ck(load_host_library().vo_host_triangulate(_dp(R), _dp(t), n1.ctypes.data_as(fp), n2.ctypes.data_as(fp), len(n1), int(force_scalar), _dp(X)))
    return X


def read_gt_line(line):
    T = np.empty(16)
    _ck(load_host_library().vo_host_read_gt_line(line.encode(), _dp(T)))
    return T.reshape(4, 4)


def format_pose_csv(T):
    T = np.ascontiguousarray(T, np.float64).reshape(16)
    buf = C.create_string_buffer(512)
    _ck(load_host_library().vo_host_format_pose_csv(_dp(T), buf, 512))
    return buf.value.decode()


def euler(R):
    R = np.ascontiguousarray(R, np.float64).reshape(9)
    out = np.empty(3)
    _ck(load_host_library().vo_host_euler(_dp(R), _dp(out)))
    return out


def read_gray(path):
    lib = load_host_library()
    r, c = C.c_int(), C.c_int()
    _ck(lib.vo_host_read_gray(str(path).encode(), None, C.c_size_t(0), C.byref(r), C.byref(c)))
    img = np.empty((r.value, c.value), np.uint8)
    _ck(lib.vo_host_read_gray(str(path).encode(), img.ctypes.data_as(C.POINTER(C.c_uint8)), C.c_size_t(img.size), C.byref(r), C.byref(c)))
    return img


def write_png(path, img):
    img = np.ascontiguousarray(img, np.uint8)
    _ck(load_host_library().vo_host_write_png(str(path).encode(), img.ctypes.data_as(C.POINTER(C.c_uint8)), img.shape[0], img.shape[1]))


def write_pgm(path, img):
    img = np.ascontiguousarray(img, np.uint8)
    _ck(load_host_library().vo_host_write_pgm(str(path).encode(), img.ctypes.data_as(C.POINTER(C.c_uint8)), img.shape[0], img.shape[1]))


STAT_KEYS = ["frames", "pairs_ok", "skipped_frames", "missing_frames", "rematched_pairs", "wall_s", "decode_s", "frontend_s",
             "host_s", "compose_s", "ransac_iterations", "mean_keypoints", "mean_matches", "mean_inliers", "stale_model_pairs", "capacity_frames"]


def pipeline_run(image_dir, num_images, pose_file, output_csv, num_threads=8, device=0, preset=1, batch=0, seed=1,
                 reuse_model=False):
    """VisualOdometry::run.  image_dir must end with a path separator, like the reference's argument."""
    stats = np.zeros(16)
    _ck(load_host_library().vo_pipeline_run_ex(device, num_threads, str(image_dir).encode(), num_images, str(pose_file).encode(),
                                               str(output_csv).encode(), preset, batch, C.c_uint64(seed), int(bool(reuse_model)),
                                               _dp(stats)))
    return dict(zip(STAT_KEYS, stats.tolist()))


def pipeline_release():
    """vo_pipeline_release: drop the GPU context and the pooled page-locked buffers pipeline_run keeps between calls."""
    _ck(load_host_library().vo_pipeline_release())
