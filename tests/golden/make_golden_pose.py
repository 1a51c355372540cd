#!/usr/bin/env python3
"""Golden vectors for PoseUpdate::getPose (PoseUpdate.hpp:61-179), produced with the reference's own third-party
calls.  The reference's C++ cannot be compiled in the build container (no C++ OpenCV), but every numerical step of
getPose is a cv:: call, and Python cv2 (4.13.0, the same library) is present: this script mirrors getPose LINE BY LINE
on cv2 (cv2.SVDecomp, cv2.determinant, cv2.undistortPoints, cv2.triangulatePoints, cv2.norm) and stores inputs and
outputs in tests/golden/pose_kat.npz.  tests/test_host.py checks vo_host_get_pose and oracle/host_oracle.py against it.

    python tests/golden/make_golden_pose.py
"""
from pathlib import Path

import cv2
import numpy as np

K = np.array([[7.188560000000e+02, 0, 6.071928000000e+02],      # PoseUpdate.hpp:36-39
              [0, 7.188560000000e+02, 1.852157000000e+02],
              [0, 0, 1.000000000000e+00]])
W = np.array([[0, -1.0, 0], [1.0, 0, 0], [0, 0, 1.0]])           # PoseUpdate.hpp:31-34


def get_pose_cv2(F, points1, points2, gt_pos_norm):
    F = np.asarray(F, np.float64)
    E = K.T @ F @ K                                              # :67
    E = E / cv2.norm(E)                                          # :69
    if cv2.countNonZero(E) < 5:                                  # :71
        raise RuntimeError("Degenerate essential matrix")
    S, U, Vt = cv2.SVDecomp(E)                                   # :76
    if cv2.determinant(U) < 0:                                   # :81-82
        U = U * -1
    if cv2.determinant(Vt) < 0:
        Vt = Vt * -1
    R1, R2 = U @ W @ Vt, U @ W.T @ Vt                            # :85-88
    if cv2.determinant(R1) < 0:
        R1 = -R1
    if cv2.determinant(R2) < 0:
        R2 = -R2
    t = U[:, 2:3].copy()                                         # :91
    best, votes, R_final, t_final = -1, [], None, None
    for R, tc in ((R1, t), (R1, -t), (R2, t), (R2, -t)):         # :94-96
        P1 = np.eye(3, 4)
        P2 = np.hstack([R, tc])
        n1 = cv2.undistortPoints(points1.reshape(-1, 1, 2), K, None)   # :110-111 (float32 in, float32 out)
        n2 = cv2.undistortPoints(points2.reshape(-1, 1, 2), K, None)
        X4 = cv2.triangulatePoints(P1, P2, n1.reshape(-1, 2).T, n2.reshape(-1, 2).T).astype(np.float64)   # :114, :122
        cnt = 0
        for i in range(X4.shape[1]):                             # :125-137
            x = X4[:, i]
            w = x[3]
            if abs(w) < 1e-6:
                continue
            x = x / w
            z1 = x[2]
            z2 = (R @ x[:3] + tc[:, 0])[2]
            if z1 > 0 and z2 > 0:
                cnt += 1
        votes.append(cnt)
        if cnt > best:                                           # :141-145
            best, R_final, t_final = cnt, R.copy(), tc.copy()
    if cv2.determinant(R_final) < 0:                             # :155
        R_final = -R_final
    nrm = cv2.norm(t_final)                                      # :163-165
    if nrm > 1e-6:
        t_final = t_final * (gt_pos_norm / nrm)
    return R_final, t_final[:, 0], votes


def rot(axis, ang):
    r, _ = cv2.Rodrigues(np.asarray(axis, np.float64) / np.linalg.norm(axis) * ang)
    return r


def scene(n, R, t, seed, noise=0.0):
    rng = np.random.default_rng(seed)
    X = np.column_stack([rng.uniform(-8, 8, n), rng.uniform(-3, 3, n), rng.uniform(6, 40, n)])
    x1 = (K @ X.T).T
    x2 = (K @ (R @ X.T + t[:, None])).T
    p1, p2 = x1[:, :2] / x1[:, 2:], x2[:, :2] / x2[:, 2:]
    if noise:
        p1 = p1 + rng.normal(0, noise, p1.shape)
        p2 = p2 + rng.normal(0, noise, p2.shape)
    return p1.astype(np.float32), p2.astype(np.float32)


def main():
    cases = [([0, 1, 0], 0.05, [0.1, 0.0, 1.0], 0.0, 2.5), ([1, 0.2, 0], -0.03, [-0.3, 0.05, 0.9], 0.0, 1.0),
             ([0, 0, 1], 0.1, [1.0, 0.1, 0.2], 0.3, 0.8), ([0.3, 1, 0.1], 0.02, [0.01, -0.02, 1.0], 0.5, 1.3),
             ([0, 1, 0], -0.2, [0.5, 0.0, -1.0], 0.2, 3.0)]
    out = {"K": K, "cv2_version": np.array(cv2.__version__)}
    for k, (axis, ang, t, noise, step) in enumerate(cases):
        R, t = rot(axis, ang), np.asarray(t, np.float64) / np.linalg.norm(t)
        p1, p2 = scene(300, R, t, 100 + k, noise)
        tx = np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]])
        F = np.linalg.inv(K).T @ tx @ R @ np.linalg.inv(K)
        F = F / np.linalg.norm(F) * (1.0 + k)                    # any scale: getPose normalises E
        Rg, tg, votes = get_pose_cv2(F, p1, p2, step)
        out.update({f"F{k}": F, f"p1_{k}": p1, f"p2_{k}": p2, f"step{k}": np.float64(step), f"R{k}": Rg, f"t{k}": tg,
                    f"votes{k}": np.array(votes), f"Rtrue{k}": R, f"ttrue{k}": t})
    out["n_cases"] = np.int64(len(cases))
    path = Path(__file__).resolve().parent / "pose_kat.npz"
    np.savez_compressed(path, **out)
    print("wrote", path, "cv2", cv2.__version__)


if __name__ == "__main__":
    main()
