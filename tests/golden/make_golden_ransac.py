#!/usr/bin/env python3
"""Golden vectors for the deterministic parts of ransac.cpp: computeSampsonError (:12-23), normalizePoints (:25-61) and
FundamentalMatrix::computeFundamentalMatrix (:63-93).  The reference's file needs Eigen, which this image does not have,
so it cannot be compiled here; its only uses of Eigen are 3x3 / Nx9 products and two JacobiSVD calls.  This script mirrors
those functions LINE BY LINE with plain Python floats for the arithmetic (same operation order as the C++) and cv2's SVD
(cv2.SVDecomp, LAPACK-free Jacobi in OpenCV 4.13 for these sizes) for the two decompositions, independently of both
vo_ransac.cpp (one-sided Jacobi / Householder null vector) and oracle/host_oracle.py (numpy LAPACK).  Singular vectors
are unique up to sign and F is a null vector, unique up to sign too: the stored F is normalised to a positive entry of
largest magnitude, and consumers compare after the same normalisation.

    python tests/golden/make_golden_ransac.py          # writes tests/golden/ransac_kat.npz
"""
from pathlib import Path

import cv2
import numpy as np


def sampson_error(F, x, xp):                                     # ransac.cpp:12-23
    Fx = [sum(F[r][c] * x[c] for c in range(3)) for r in range(3)]
    Ftx = [sum(F[r][c] * xp[r] for r in range(3)) for c in range(3)]
    e = sum(xp[r] * Fx[r] for r in range(3))
    numerator = e ** 2
    denominator = Fx[0] * Fx[0] + Fx[1] * Fx[1] + Ftx[0] * Ftx[0] + Ftx[1] * Ftx[1]
    if denominator < 1e-12:
        return float(np.finfo(np.float64).max)
    return numerator / denominator


def normalize_points(points):                                    # ransac.cpp:25-61
    mx1 = my1 = mx2 = my2 = 0.0
    for x1, y1, x2, y2 in points:
        mx1 += x1; my1 += y1; mx2 += x2; my2 += y2
    n = len(points)
    mx1 /= n; my1 /= n; mx2 /= n; my2 /= n
    s1 = s2 = 0.0
    for x1, y1, x2, y2 in points:
        s1 += (x1 - mx1) ** 2 + (y1 - my1) ** 2
        s2 += (x2 - mx2) ** 2 + (y2 - my2) ** 2
    s1 = np.sqrt(2.0) / np.sqrt(s1 / n)
    s2 = np.sqrt(2.0) / np.sqrt(s2 / n)
    T1 = np.array([[s1, 0, -s1 * mx1], [0, s1, -s1 * my1], [0, 0, 1.0]])
    T2 = np.array([[s2, 0, -s2 * mx2], [0, s2, -s2 * my2], [0, 0, 1.0]])
    n1 = np.array([T1 @ np.array([x1, y1, 1.0]) for x1, y1, _, _ in points])
    n2 = np.array([T2 @ np.array([x2, y2, 1.0]) for _, _, x2, y2 in points])
    return n1, n2, T1, T2


def compute_fundamental(points):                                 # ransac.cpp:63-93
    n1, n2, T1, T2 = normalize_points(points)
    A = np.empty((len(points), 9))
    for i in range(len(points)):
        p1, p2 = n1[i], n2[i]
        A[i] = [p1[0] * p2[0], p1[0] * p2[1], p1[0], p1[1] * p2[0], p1[1] * p2[1], p1[1], p2[0], p2[1], 1.0]
    # JacobiSVD<MatrixXd>(A, ComputeFullV).matrixV().col(8): for fewer than 9 rows the full V needs the null space, so the
    # matrix is padded with zero rows (the right singular vectors are those of A^T A either way)
    Ap = A if len(points) >= 9 else np.vstack([A, np.zeros((9 - len(points), 9))])
    _, _, Vt = cv2.SVDecomp(Ap, flags=cv2.SVD_FULL_UV)
    F = Vt[8].reshape(3, 3)
    F = T2.T @ F @ T1
    S, U, Vt3 = cv2.SVDecomp(F, flags=cv2.SVD_FULL_UV)
    S = S.reshape(-1).copy()
    S[2] = 0
    return U @ np.diag(S) @ Vt3, n1, n2, T1, T2


def canon(F):
    F = np.asarray(F, np.float64)
    k = np.argmax(np.abs(F))
    return F * (1.0 if F.reshape(-1)[k] > 0 else -1.0)


def two_view_points(rng, n, noise=0.0):
    """Correspondences of a random rigid scene under the KITTI intrinsics (PoseUpdate.hpp:36-39)."""
    K = np.array([[718.856, 0, 607.1928], [0, 718.856, 185.2157], [0, 0, 1.0]])
    X = np.c_[rng.uniform(-8, 8, n), rng.uniform(-3, 2, n), rng.uniform(6, 40, n)]
    a = rng.uniform(-0.05, 0.05, 3)
    R, _ = cv2.Rodrigues(a)
    t = np.array([rng.uniform(-0.2, 0.2), rng.uniform(-0.05, 0.05), rng.uniform(0.5, 1.2)])
    x1 = X @ K.T
    x2 = (X @ R.T + t) @ K.T
    p = np.c_[x1[:, :2] / x1[:, 2:], x2[:, :2] / x2[:, 2:]]
    return p + rng.normal(0, noise, p.shape) if noise else p


def main():
    rng = np.random.default_rng(20240611)
    out = {}
    sizes = [8, 8, 9, 12, 50, 200, 777]
    for k, n in enumerate(sizes):
        pts = two_view_points(rng, n, noise=0.0 if k % 2 == 0 else 0.4)
        if k == 1:
            pts = np.round(pts)  # keypoints are integer pixel positions in the pipeline
        F, n1, n2, T1, T2 = compute_fundamental([tuple(r) for r in pts])
        test = two_view_points(rng, 64, noise=1.0)
        err = np.array([sampson_error(F.tolist(), [r[0], r[1], 1.0], [r[2], r[3], 1.0]) for r in test])
        out[f"pts{k}"], out[f"F{k}"] = pts, canon(F)
        out[f"n1_{k}"], out[f"n2_{k}"], out[f"T1_{k}"], out[f"T2_{k}"] = n1, n2, T1, T2
        out[f"test{k}"], out[f"err{k}"] = test, err
    # degenerate denominator: F = 0 gives "max" (ransac.cpp:21)
    out["n_cases"] = np.array(len(sizes))
    np.savez_compressed(Path(__file__).with_name("ransac_kat.npz"), **out)
    print("wrote", Path(__file__).with_name("ransac_kat.npz"), {k: v.shape for k, v in out.items() if k.startswith("F")})


if __name__ == "__main__":
    main()
