"""TEST INFRASTRUCTURE ONLY - numpy restatement of the reference's host-side consumers of the match lists:
ransac.cpp (normalised 8-point fit, Sampson error) and PoseUpdate.hpp (getPose).  Only tests/ may import this.

PARITY UNPINNED for these rows: the reference's own code needs Eigen (ransac.cpp) and OpenCV calib3d
(PoseUpdate.hpp); neither exists in this image, so the reference cannot be compiled here and ships no golden
vectors for them.  The restatement follows the reference line by line with numpy's LAPACK SVD standing in for
Eigen::JacobiSVD / cv::SVD; results are compared up to the sign ambiguity of singular vectors.
"""
import numpy as np

K_KITTI = np.array([[7.188560000000e+02, 0, 6.071928000000e+02],
                    [0, 7.188560000000e+02, 1.852157000000e+02],
                    [0, 0, 1.0]])


def normalize_points(pts):
    """ransac.cpp:25-61.  pts: n x 4 (x1, y1, x2, y2)."""
    out = []
    for a in (pts[:, 0:2], pts[:, 2:4]):
        mean = a.mean(axis=0)
        s = np.sqrt(2.0) / np.sqrt(((a - mean) ** 2).sum(axis=1).mean())
        T = np.array([[s, 0, -s * mean[0]], [0, s, -s * mean[1]], [0, 0, 1.0]])
        out.append((np.c_[a, np.ones(len(a))] @ T.T, T))
    return out[0][0], out[1][0], out[0][1], out[1][1]


def fundamental(pts):
    """ransac.cpp:63-93, as written: rows (p1x p2x, p1x p2y, p1x, p1y p2x, p1y p2y, p1y, p2x, p2y, 1), null vector
    reshaped row-major, F = T2^T F T1, smallest singular value zeroed."""
    n1, n2, T1, T2 = normalize_points(np.asarray(pts, np.float64))
    A = np.stack([n1[:, 0] * n2[:, 0], n1[:, 0] * n2[:, 1], n1[:, 0], n1[:, 1] * n2[:, 0], n1[:, 1] * n2[:, 1], n1[:, 1],
                  n2[:, 0], n2[:, 1], np.ones(len(n1))], axis=1)
    _, _, Vt = np.linalg.svd(A, full_matrices=True)
    F = Vt[8].reshape(3, 3)
    F = T2.T @ F @ T1
    U, s, Vt = np.linalg.svd(F)
    s[2] = 0
    return U @ np.diag(s) @ Vt


def sampson(F, pts):
    """ransac.cpp:12-23 with x = (x1, y1, 1), x' = (x2, y2, 1)."""
    pts = np.asarray(pts, np.float64)
    x = np.c_[pts[:, 0:2], np.ones(len(pts))]
    xp = np.c_[pts[:, 2:4], np.ones(len(pts))]
    Fx = x @ F.T
    Ftx = xp @ F
    num = (xp * Fx).sum(axis=1) ** 2
    den = Fx[:, 0] ** 2 + Fx[:, 1] ** 2 + Ftx[:, 0] ** 2 + Ftx[:, 1] ** 2
    return np.where(den < 1e-12, np.finfo(np.float64).max, num / np.where(den < 1e-12, 1.0, den))


def triangulate(R, t, x1, x2):
    """cv::triangulatePoints, P1 = [I|0], P2 = [R|t], normalised coordinates; returns n x 4 unit vectors."""
    P1 = np.c_[np.eye(3), np.zeros(3)]
    P2 = np.c_[R, t]
    out = np.empty((len(x1), 4))
    for i in range(len(x1)):
        A = np.stack([x1[i, 0] * P1[2] - P1[0], x1[i, 1] * P1[2] - P1[1], x2[i, 0] * P2[2] - P2[0], x2[i, 1] * P2[2] - P2[1]])
        out[i] = np.linalg.svd(A)[2][3]
    return out


def get_pose(F, p1, p2, gt_step, K=K_KITTI):
    """PoseUpdate.hpp:61-179.  Returns (R, t, votes of the four candidates)."""
    E = K.T @ F @ K
    E = E / np.linalg.norm(E)
    U, _, Vt = np.linalg.svd(E)
    if np.linalg.det(U) < 0:
        U = -U
    if np.linalg.det(Vt) < 0:
        Vt = -Vt
    W = np.array([[0, -1, 0], [1, 0, 0], [0, 0, 1.0]])
    R1, R2 = U @ W @ Vt, U @ W.T @ Vt
    if np.linalg.det(R1) < 0:
        R1 = -R1
    if np.linalg.det(R2) < 0:
        R2 = -R2
    t = U[:, 2]
    n1 = ((np.asarray(p1, np.float64) - K[:2, 2]) / np.array([K[0, 0], K[1, 1]])).astype(np.float32).astype(np.float64)
    n2 = ((np.asarray(p2, np.float64) - K[:2, 2]) / np.array([K[0, 0], K[1, 1]])).astype(np.float32).astype(np.float64)
    best, votes = None, []
    for R, tc in ((R1, t), (R1, -t), (R2, t), (R2, -t)):
        X = triangulate(R, tc, n1, n2).astype(np.float32).astype(np.float64)
        ok = np.abs(X[:, 3]) >= 1e-6
        P = X[ok, :3] / X[ok, 3:4]
        z2 = (P @ R.T + tc)[:, 2]
        v = int(((P[:, 2] > 0) & (z2 > 0)).sum())
        votes.append(v)
        if best is None or v > best[0]:
            best = (v, R, tc)
    _, R, tc = best
    if np.linalg.det(R) < 0:
        R = -R
    nt = np.linalg.norm(tc)
    if nt > 1e-6:
        tc = tc * (gt_step / nt)
    return R, tc, votes
