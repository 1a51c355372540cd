"""Host side of the pipeline: RANSAC, pose, formats, frame readers.
CPU tests compare libvohost.so with the numpy restatement in oracle/host_oracle.py; the GPU test runs the
pipelined VisualOdometry::run driver on a synthetic PNG sequence."""
import re
from pathlib import Path

import numpy as np
import pytest

from acs_visual_odometry_b200 import host
from oracle import host_oracle as ho

ROOT = Path(__file__).resolve().parents[1]
K = ho.K_KITTI


def rot(axis, ang):
    axis = np.asarray(axis, float) / np.linalg.norm(axis)
    Kx = np.array([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
    return np.eye(3) + np.sin(ang) * Kx + (1 - np.cos(ang)) * Kx @ Kx


def scene(n, R, t, seed=0, noise=0.0):
    """n 3-D points in front of both cameras; pixel coordinates in view 1 (P = K[I|0]) and view 2 (K[R|t])."""
    rng = np.random.default_rng(seed)
    X = np.c_[rng.uniform(-8, 8, n), rng.uniform(-2, 2, n), rng.uniform(6, 40, n)]
    x1 = X @ K.T
    x2 = (X @ R.T + t) @ K.T
    p1, p2 = x1[:, :2] / x1[:, 2:], x2[:, :2] / x2[:, 2:]
    if noise:
        p1 = p1 + rng.normal(0, noise, p1.shape)
        p2 = p2 + rng.normal(0, noise, p2.shape)
    return np.c_[p1, p2]


def same_up_to_sign(a, b, tol):
    a, b = a / np.linalg.norm(a), b / np.linalg.norm(b)
    return min(np.abs(a - b).max(), np.abs(a + b).max()) < tol


def test_host_library_exports_every_declared_symbol():
    text = (ROOT / "include" / "vohost.h").read_text()
    declared = set(re.findall(r"\b(vo_(?:host|pipeline)_[a-z_0-9]+)\s*\(", text))
    assert declared == set(host.SYMBOLS)
    host.load_host_library()


def test_pipeline_release_with_nothing_kept_is_a_no_op():
    """vo_pipeline_release before any run: no context, an empty pool, nothing to hand back - and no CUDA call."""
    host.pipeline_release()
    host.pipeline_release()


def test_fundamental_matches_restatement():
    for seed, n in ((1, 8), (2, 8), (3, 40), (4, 300)):
        pts = scene(n, rot([0.1, 1, 0.05], 0.03), np.array([0.1, -0.02, 1.0]), seed, noise=0.3)
        got, want = host.fundamental(pts), ho.fundamental(pts)
        assert same_up_to_sign(got, want, 1e-8), (seed, n)
        s = np.linalg.svd(got, compute_uv=False)
        assert s[2] < 1e-12 * s[0]                                     # rank 2
        assert np.allclose(host.sampson(got, pts), ho.sampson(got, pts), rtol=1e-9, atol=1e-12)
    with pytest.raises(host.VoHostError):
        host.fundamental(scene(7, np.eye(3), np.array([0, 0, 1.0])))


def test_sampson_degenerate_denominator_is_max():
    e = host.sampson(np.zeros((3, 3)), scene(5, np.eye(3), np.array([0, 0, 1.0])))
    assert (e == np.finfo(np.float64).max).all()


def test_ransac_recovers_consensus_set():
    # forward motion (the KITTI case): F is skew-symmetric, so the reference's transposed row layout is harmless
    R, t = np.eye(3), np.array([0.02, -0.01, 1.0])
    good = scene(400, R, t, seed=7, noise=0.2)
    rng = np.random.default_rng(8)
    bad = np.c_[rng.uniform(0, 1241, (120, 1)), rng.uniform(0, 376, (120, 1)), rng.uniform(0, 1241, (120, 1)), rng.uniform(0, 376, (120, 1))]
    data = np.r_[good, bad]
    data = data[rng.permutation(len(data))]
    F, inl, it = host.ransac(data, 0.99, 1.0, num_threads=1, seed=5)
    assert 100 <= it <= 2000                                          # the reference's clamp (ransac.cpp:186)
    is_good = (inl[:, None, :] == good[None, :, :]).all(axis=2).any(axis=1)
    assert len(inl) >= 0.75 * len(good) and is_good.mean() > 0.97
    assert np.median(ho.sampson(F, good)) < 1.0
    # same seed, same result; the final model is the fit of the consensus set
    F2, inl2, it2 = host.ransac(data, 0.99, 1.0, num_threads=1, seed=5)
    assert np.array_equal(F, F2) and np.array_equal(inl, inl2) and it == it2
    assert same_up_to_sign(F, ho.fundamental(inl), 1e-8)


def test_ransac_scores_only_whole_chunks_like_the_reference():
    # ransac.cpp:152-157: numThreads chunks of N / numThreads points; the remaining N % numThreads are never scored
    data = scene(100, np.eye(3), np.array([0.0, 0.0, 1.0]), seed=3)
    for threads, want in ((1, 100), (3, 99), (7, 98), (16, 96)):
        _, inl, _ = host.ransac(data, 0.99, 1.0, num_threads=threads, seed=2)
        assert len(inl) == want and np.array_equal(inl, data[:want]), threads
    _, inl, _ = host.ransac(data[:10], 0.99, 1.0, num_threads=16, seed=2)  # chunk size 0: nothing is scored
    assert len(inl) == 0
    F, inl, it = host.ransac(np.zeros((0, 4)), seed=2)
    assert len(inl) == 0 and it > 0


def test_get_pose_matches_restatement_and_truth():
    for seed, (axis, ang, t) in enumerate([([0, 1, 0], 0.05, [0.1, 0.0, 1.0]), ([1, 0.2, 0], -0.03, [-0.3, 0.05, 0.9]),
                                           ([0, 0, 1], 0.1, [1.0, 0.1, 0.2]), ([0.3, 1, 0.1], 0.0, [0.01, -0.02, 1.0])]):
        R, t = rot(axis, ang), np.array(t) / np.linalg.norm(t)
        pts = scene(200, R, t, seed)
        tx = np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]])
        F = np.linalg.inv(K).T @ tx @ R @ np.linalg.inv(K)             # x2^T F x1 = 0
        Rg, tg, votes = host.get_pose(F, pts[:, :2], pts[:, 2:], gt_step=2.5)
        Ro, to, votes_o = ho.get_pose(F, pts[:, :2].astype(np.float32), pts[:, 2:].astype(np.float32), 2.5)
        assert np.allclose(Rg, Ro, atol=1e-6) and np.allclose(tg, to, atol=1e-6), seed
        assert votes == max(votes_o) == 200
        assert np.allclose(Rg, R, atol=1e-5) and np.allclose(tg, 2.5 * t, atol=1e-4), seed
        assert abs(np.linalg.det(Rg) - 1) < 1e-9
    # fewer than five non-zero entries in E: "Degenerate essential matrix", as in the reference (PoseUpdate.hpp:70-72);
    # pure forward motion without rotation, E = [t]x with t = (0, 0, 1), is such a case
    with pytest.raises(host.VoHostError) as e:
        host.get_pose(np.array([[0, -1.0, 0], [1.0, 0, 0], [0, 0, 0]]), pts[:, :2], pts[:, 2:], K=np.eye(3))
    assert "Degenerate" in str(e.value)


def test_cheirality_shortcut_against_four_triangulations():
    """getPose takes the votes of (R, -t) from the triangulation of (R, t) and stops triangulating a rotation that cannot
    win (vo_pose.cpp); the reference triangulates all four candidates (PoseUpdate.hpp:98-146).  The two agree up to the
    solver's tolerance, so the place to look for a difference is data where votes hang on |w| ~ 1e-6 or depth ~ 0:
    baselines down to a millimetre, a third of the points between 2 km and 1000 km away, half-pixel noise.  Winner and
    count are compared with oracle/host_oracle.py's four brute-force triangulations: the count within 2 votes (borderline
    points may fall either way), the selected (R, t) exactly whenever the brute-force margin is larger than that slack."""
    checked = 0
    for seed in range(8):
        for base in (1.0, 0.02, 0.001):
            for ang in (0.01, 0.0003):
                rng = np.random.default_rng(100 + seed)
                n = 600
                R = rot([0.1, 1, 0.05], ang)
                t = np.array([0.2, -0.05, 1.0]); t = t / np.linalg.norm(t) * base
                X = np.c_[rng.uniform(-8, 8, n), rng.uniform(-2, 2, n), rng.uniform(6, 40, n)]
                X[:n // 3, 2] = rng.uniform(2000, 1e6, n // 3)
                X[:n // 3, :2] *= X[:n // 3, 2:3] / 20
                x1, x2 = X @ K.T, (X @ R.T + t) @ K.T
                p1 = (x1[:, :2] / x1[:, 2:] + rng.normal(0, 0.5, (n, 2))).astype(np.float32)
                p2 = (x2[:, :2] / x2[:, 2:] + rng.normal(0, 0.5, (n, 2))).astype(np.float32)
                tu = t / np.linalg.norm(t)
                tx = np.array([[0, -tu[2], tu[1]], [tu[2], 0, -tu[0]], [-tu[1], tu[0], 0]])
                F = np.linalg.inv(K).T @ tx @ R @ np.linalg.inv(K)
                Rg, tg, votes = host.get_pose(F, p1, p2, gt_step=1.0)
                Ro, to, votes_o = ho.get_pose(F, p1, p2, 1.0)
                top = sorted(votes_o)
                assert abs(votes - top[-1]) <= 2, (seed, base, ang, votes, votes_o)
                if top[-1] - top[-2] > 4:
                    assert np.allclose(Rg, Ro, atol=1e-9) and np.allclose(tg, to, atol=1e-9), (seed, base, ang, votes, votes_o)
                    checked += 1
    assert checked >= 24


def test_get_pose_against_opencv_golden():
    """tests/golden/pose_kat.npz: PoseUpdate::getPose mirrored line by line on the reference's own third-party calls
    (Python cv2 4.13: SVDecomp, undistortPoints, triangulatePoints), generated by tests/golden/make_golden_pose.py.
    The SVD's sign freedom permutes the four candidates, so the vote LIST is compared as a multiset; the selected
    (R, t) is unique.  Tolerance 1e-9 (double arithmetic on both sides; measured 1e-15)."""
    g = np.load(Path(__file__).parent / "golden" / "pose_kat.npz")
    for k in range(int(g["n_cases"])):
        F, p1, p2, step = g[f"F{k}"], g[f"p1_{k}"], g[f"p2_{k}"], float(g[f"step{k}"])
        R, t, votes = host.get_pose(F, p1, p2, gt_step=step)
        assert np.allclose(R, g[f"R{k}"], atol=1e-9) and np.allclose(t, g[f"t{k}"], atol=1e-9), k
        assert votes == int(g[f"votes{k}"].max())
        Ro, to, votes_o = ho.get_pose(F, p1, p2, step)
        assert np.allclose(Ro, g[f"R{k}"], atol=1e-9) and np.allclose(to, g[f"t{k}"], atol=1e-9), k
        assert sorted(votes_o) == sorted(g[f"votes{k}"].tolist())


def test_formats():
    line = "9.999978e-01 5.272628e-04 -2.066935e-03 -4.690294e-02 -5.296506e-04 9.999992e-01 -1.154865e-03 -2.839928e-02 2.066324e-03 1.155958e-03 9.999971e-01 8.586941e-01"
    T = host.read_gt_line(line)
    assert np.array_equal(T[:3].reshape(-1), np.array([float(v) for v in line.split()])) and np.array_equal(T[3], [0, 0, 0, 1])
    assert np.array_equal(host.read_gt_line("1 2 3")[0], [1, 2, 3, 0]) and host.read_gt_line("")[1, 1] == 1   # short lines keep the identity
    out = host.format_pose_csv(T)
    assert out == ",".join(format(v, ".9g") for v in T[:3].reshape(-1)) + "\n"                          # setprecision(9), no trailing comma
    assert host.format_pose_csv(np.diag([1, 1, -1, 1.0])) == "1,0,0,0,0,1,0,0,0,0,-1,0\n"               # flipZ
    big = np.eye(4); big[0, 3] = 123456789.123; big[1, 3] = 1e-7
    assert host.format_pose_csv(big).split(",")[3] == "123456789" and "1e-07" in host.format_pose_csv(big)
    R = rot([0.2, -1, 0.4], 0.7)
    sy = np.hypot(R[0, 0], R[1, 0])
    assert np.allclose(host.euler(R), [np.arctan2(R[2, 1], R[2, 2]), np.arctan2(-R[2, 0], sy), np.arctan2(R[1, 0], R[0, 0])])


def test_frame_readers(tmp_path):
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(0)
    gray = rng.integers(0, 256, (37, 53), dtype=np.uint8)
    cv2.imwrite(str(tmp_path / "g.png"), gray)                                        # filtered, deflated by libpng
    assert np.array_equal(host.read_gray(tmp_path / "g.png"), gray)
    smooth = (np.add.outer(np.arange(64), np.arange(80)) % 256).astype(np.uint8)      # exercises the sub/up/paeth filters
    cv2.imwrite(str(tmp_path / "s.png"), smooth, [cv2.IMWRITE_PNG_COMPRESSION, 9])
    assert np.array_equal(host.read_gray(tmp_path / "s.png"), smooth)
    bgr = rng.integers(0, 256, (21, 34, 3), dtype=np.uint8)
    cv2.imwrite(str(tmp_path / "c.png"), bgr)
    assert np.array_equal(host.read_gray(tmp_path / "c.png"), cv2.imread(str(tmp_path / "c.png"), cv2.IMREAD_GRAYSCALE))
    host.write_png(tmp_path / "w.png", gray)
    assert np.array_equal(cv2.imread(str(tmp_path / "w.png"), cv2.IMREAD_GRAYSCALE), gray)
    host.write_pgm(tmp_path / "w.pgm", gray)
    assert np.array_equal(host.read_gray(tmp_path / "w.pgm"), gray)
    cv2.imwrite(str(tmp_path / "d16.png"), gray.astype(np.uint16) * 257)
    for bad in ("missing.png", "d16.png"):
        with pytest.raises(host.VoHostError) as e:
            host.read_gray(tmp_path / bad)
        assert e.value.code == -2


@pytest.mark.gpu
def test_pipeline_run_on_png_sequence(tmp_path):
    from acs_visual_odometry_b200 import Params, VoFront
    from acs_visual_odometry_b200.synth import make_sequence
    n = 23
    frames = make_sequence(n, 200, 320, seed=4, n_rects=260)
    d = tmp_path / "seq"
    d.mkdir()
    for i, f in enumerate(frames):
        host.write_png(d / f"{i:06d}.png", f)
    gt = tmp_path / "poses.txt"
    gt.write_text("".join(f"1 0 0 {0.1 * i:.6f} 0 1 0 0 0 0 1 {0.5 * i:.6f}\n" for i in range(n)))
    image_dir = str(d) + "/"
    st = host.pipeline_run(image_dir, n, gt, tmp_path / "a.csv", num_threads=4, batch=8, seed=11)
    rows = np.loadtxt(tmp_path / "a.csv", delimiter=",")
    assert rows.shape == (n, 12) and np.array_equal(rows[0], np.eye(4)[:3].reshape(-1))
    assert st["frames"] == n and st["pairs_ok"] + st["skipped_frames"] == n - 1 and st["missing_frames"] == 0
    # the front end inside the pipeline is the same one: keypoints and matches per pair equal a direct sequence run
    p = Params.vo_gpu_ref()
    with VoFront(200, 320, p.max_keypoints, 8) as vo:
        nk, _, nm, _ = vo.sequence_run(frames, p, 4)
    assert abs(st["mean_keypoints"] - nk[1:].mean()) < 1e-9 and abs(st["mean_matches"] - nm.mean()) < 1e-9
    # poses: rotations orthonormal after the flip, step lengths equal to the ground-truth step of every solved pair
    step = np.hypot(0.1, 0.5)
    T = [np.diag([1, 1, -1.0]) @ r.reshape(3, 4) for r in rows]
    for a, b in zip(T[:-1], T[1:]):
        assert np.allclose(b[:, :3] @ b[:, :3].T, np.eye(3), atol=1e-6)
        d_ = np.linalg.norm(a[:, :3].T @ (b[:, 3] - a[:, 3]))
        assert d_ < 1e-9 or abs(d_ / step - round(d_ / step)) < 1e-6
    # a fixed seed reproduces the file; chunking does not change it (num_threads does: it sets which matches RANSAC scores)
    host.pipeline_run(image_dir, n, gt, tmp_path / "b.csv", num_threads=4, batch=5, seed=11)
    assert (tmp_path / "a.csv").read_text() == (tmp_path / "b.csv").read_text()
    # an unreadable frame repeats the pose WITHOUT the flip (VisualOdometry.cpp:78-83) and the query frame is kept
    (d / "000009.png").unlink()
    st = host.pipeline_run(image_dir, n, gt, tmp_path / "c.csv", num_threads=4, batch=8, seed=11)
    rows_c = np.loadtxt(tmp_path / "c.csv", delimiter=",")
    assert st["missing_frames"] == 1 and rows_c.shape == (n, 12)
    assert np.allclose(rows_c[9].reshape(3, 4), np.diag([1, 1, -1.0]) @ rows_c[8].reshape(3, 4), atol=1e-6)
    assert np.array_equal(rows_c[:9], rows[:9])
    with pytest.raises(host.VoHostError):
        host.pipeline_run(image_dir, n, tmp_path / "nope.txt", tmp_path / "d.csv")


# ---- N2 pinned: deterministic parts of ransac.cpp against a cv2.SVDecomp mirror (tests/golden/make_golden_ransac.py) ----
def _canon(F):
    F = np.asarray(F, np.float64)
    k = np.argmax(np.abs(F))
    return F * (1.0 if F.reshape(-1)[k] > 0 else -1.0)


def test_ransac_pieces_against_opencv_golden():
    g = np.load(ROOT / "tests" / "golden" / "ransac_kat.npz")
    assert int(g["n_cases"]) >= 7
    for k in range(int(g["n_cases"])):
        pts, F = g[f"pts{k}"], g[f"F{k}"]
        # normalizePoints: same operations in the same order - agreement to the last few ulps
        n1, n2, T1, T2 = host.normalize_points(pts)
        for got, want in ((n1, g[f"n1_{k}"]), (n2, g[f"n2_{k}"]), (T1, g[f"T1_{k}"]), (T2, g[f"T2_{k}"])):
            np.testing.assert_allclose(got, want, rtol=1e-13, atol=1e-13)
        # computeFundamentalMatrix: null vector + rank-2 projection through a different SVD: 1e-9 of the largest entry
        # (measured 5e-13 worst case, the exactly determined 8-point set)
        scale = np.abs(F).max()
        for impl in (host.fundamental, ho.fundamental):
            assert np.abs(_canon(impl(pts)) - F).max() / scale < 1e-9, (k, impl)
        # computeSampsonError: plain arithmetic in the reference's order - bit-identical
        np.testing.assert_array_equal(host.sampson(F, g[f"test{k}"]), g[f"err{k}"])
        np.testing.assert_allclose(ho.sampson(F, g[f"test{k}"]), g[f"err{k}"], rtol=1e-9)  # numpy sums in another order; cancellation in x'Fx


# ---- N1: the sequential restatement of the frame loop, exercised with scripted stages ----
def _scripted_loop(script, reuse_model):
    """script[i] for frame i >= 1: None (unreadable) or (n_matches, n_inliers).  Stages return recognisable values so the
    composition can be followed by hand."""
    from oracle.vo_loop_oracle import run_loop
    n = len(script) + 1
    frames = [np.full((4, 4), 0, np.uint8)] + [None if s is None else np.full((4, 4), i + 1, np.uint8) for i, s in enumerate(script)]
    gt = np.zeros((n, 3, 4))
    gt[:, :, :3] = np.eye(3)
    gt[:, 2, 3] = np.arange(n) * 2.0  # 2 m per frame along z
    state = {}

    def extract(img):
        i = int(img[0, 0])
        kp = np.c_[np.arange(64, dtype=np.float32) + 100 * i, np.arange(64, dtype=np.float32)]
        return kp, i

    def match(d1, d2):
        state["pair"] = (d1, d2)
        m = script[d2 - 1][0]
        return np.c_[np.arange(m), np.arange(m)].astype(np.int32)

    def ransac(pts, i):
        k = script[i - 1][1]
        return np.eye(3) * i, pts[:k]

    def get_pose(F, p1, p2, s):
        # rotation identity, translation s along z: T_curr.z accumulates the ground-truth step lengths; F tags the model
        state.setdefault("models", []).append((float(F[0, 0]), len(p1)))
        return np.eye(3), np.array([0, 0, s])

    poses, log = run_loop(frames, gt, extract, match, ransac, get_pose, reuse_model=reuse_model)
    return poses, log, state


def test_loop_oracle_skip_rules():
    script = [(20, 12), None, (3, 0), (20, 5), (20, 12), (20, 12)]
    poses, log, st = _scripted_loop(script, reuse_model=False)
    assert [e["status"] for e in log] == ["first", "ok", "missing", "few_matches", "few_inliers", "ok", "ok"]
    # the query frame only advances on success: frames 3 and 4 fail, so frame 5 is matched against frame 1
    assert [e.get("query") for e in log[3:]] == [1, 1, 1, 5]
    z = poses[:, 2, 3]
    # pose 0 is the identity unflipped; a missing frame repeats T_curr WITHOUT the flip (VisualOdometry.cpp:80), skipped
    # frames repeat it flipped; the step of frame 5 is the ground-truth distance from frame 1 (4 frames x 2 m)
    np.testing.assert_allclose(z, [0, -2, 2, -2, -2, -10, -12])
    assert poses[2, 2, 2] == 1.0 and poses[3, 2, 2] == -1.0
    # with the reference's long-lived model the low-inlier frame 4 is NOT skipped: it re-uses frame 1's model
    poses2, log2, st2 = _scripted_loop(script, reuse_model=True)
    assert [e["status"] for e in log2] == ["first", "ok", "missing", "few_matches", "ok_stale_model", "ok", "ok"]
    assert st2["models"][1] == (1.0, 12)  # F and inliers of frame 1
    assert [e.get("query") for e in log2[3:]] == [1, 1, 4, 5]
    np.testing.assert_allclose(poses2[:, 2, 3], [0, -2, 2, -2, -8, -10, -12])
    # before any pair has succeeded the guard does fire in both modes
    _, log3, _ = _scripted_loop([(20, 5), (20, 12)], reuse_model=True)
    assert [e["status"] for e in log3] == ["first", "few_inliers", "ok"]


def test_vector_scoring_equals_scalar_and_sampson():
    """Ransac::run's consensus loop: the AVX2 form (no FMA) must flag exactly what the scalar form flags, and both must
    be computeSampsonError < threshold; sizes around the vector width and a threshold that cuts through the data."""
    rng = np.random.default_rng(8)
    R, t = rot([0.2, 1, 0.1], 0.02), np.array([0.1, -0.02, 0.9])
    F = host.fundamental(scene(40, R, t, seed=1))
    for n in (0, 1, 3, 4, 5, 63, 64, 1001):
        pts = scene(max(n, 1), R, t, seed=n, noise=0.6)[:n]
        if n > 10:
            pts[::7, 2:] += rng.normal(0, 3, pts[::7, 2:].shape)
        for thr in (1.0, 0.2, 1e-6):
            fv, cv_ = host.score(F, pts, thr, force_scalar=False)
            fs, cs = host.score(F, pts, thr, force_scalar=True)
            assert cv_ == cs == int(fs.sum()) and np.array_equal(fv, fs)
            if n:
                assert np.array_equal(fs.astype(bool), host.sampson(F, pts) < thr)
    # a zero model: every denominator is below 1e-12, every error "max" (ransac.cpp:21)
    assert host.score(np.zeros((3, 3)), scene(9, R, t), 1.0)[1] == 0


def test_vector_triangulation_equals_scalar():
    """getPose's cheirality test triangulates every inlier: the four-wide AVX2 routine must return the scalar routine's
    bits, for well-conditioned points and for the ones that take the fallback (rays parallel to the baseline, points at
    infinity, identical points)."""
    rng = np.random.default_rng(12)
    R, t = rot([0.1, 1, 0.2], 0.03), np.array([0.2, -0.05, 1.0])
    t = t / np.linalg.norm(t)
    Kinv = np.linalg.inv(K)
    for n in (1, 3, 4, 5, 8, 203):
        p = scene(n, R, t, seed=40 + n, noise=0.5)
        n1 = (np.c_[p[:, :2], np.ones(n)] @ Kinv.T)[:, :2].astype(np.float32)
        n2 = (np.c_[p[:, 2:], np.ones(n)] @ Kinv.T)[:, :2].astype(np.float32)
        if n >= 8:  # degenerate lanes in the middle of a vector
            n2[1] = n1[1]                          # no parallax
            n1[2] = n2[2] = 0                      # on both optical axes
            n2[5] = (n1[5] + 1e3).astype(np.float32)
            n1[6] = np.float32(1e-30)
        for Rm, tv in ((R, t), (R, -t), (np.eye(3), np.array([0, 0, 1.0])), (np.eye(3), np.zeros(3))):
            a = host.triangulate(Rm, tv, n1, n2, force_scalar=False)
            b = host.triangulate(Rm, tv, n1, n2, force_scalar=True)
            assert np.array_equal(a.view(np.uint64), b.view(np.uint64)) or np.array_equal(np.isnan(a), np.isnan(b)) and np.array_equal(a[~np.isnan(a)], b[~np.isnan(b)])
