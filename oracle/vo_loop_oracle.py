"""TEST INFRASTRUCTURE - plain sequential restatement of the reference's frame loop, VisualOdometry::run
(/root/reference/VisualOdometry.cpp:38-193).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
import this module; the product (acs_visual_odometry_b200/host/vo_pipeline.cpp) must never call it.

One frame at a time, exactly in the reference's order, with the stages plugged in as functions so that the loop and
its skip rules are what is under test:

    extract(gray)            -> (kp [n,2] float32 (x, y), desc)        VisualOdometry.cpp:28-31
    match(desc1, desc2)      -> pairs [m,2] int32 (query, train)       :34-36
    ransac(pts [m,4], i)     -> (F [3,3], inliers [k,4])               :125 (i = index of the frame being added)
    get_pose(F, p1, p2, s)   -> (R [3,3], t [3])                       :172

Rules restated (line numbers of VisualOdometry.cpp):
  * frame 0 provides the first query set (:64-66); the pose list starts with the identity, NOT flipped (:58-59);
  * an unreadable frame repeats the current pose WITHOUT the z flip and moves on (:78-83);
  * fewer than 8 matches: pose repeated (flipped), frame skipped, the query frame stays (:108-115);
  * RANSAC runs on ONE model object that lives across frames (:49, :125).  Ransac::run ends with
    model.fit(bestInlierSet) (ransac.cpp:193) and fit() returns early when the set has fewer than 8 pairs
    (ransac.cpp:95-96) - so after a run that found no consensus the model still holds the PREVIOUS pair's F and inliers,
    and the < 8 inliers guard (:147) only fires while no pair has succeeded yet.  `reuse_model=True` restates that;
    `reuse_model=False` is the pipeline's documented default (a fresh model per pair: the guard fires, frame skipped);
  * on success the step length is |t| of inv(gt[i]) * gt[last_valid] (:160-161), the query frame advances (:163-165),
    T_curr = T_curr * [R t; 0 1] (:183) and flipZ * T_curr is recorded (:185-186).
"""
import numpy as np

FLIP_Z = np.diag([1.0, 1.0, -1.0, 1.0])  # PoseUpdate.cpp:7-11


def gt_to_mat4(gt_row):
    T = np.eye(4)
    T[:3, :4] = np.asarray(gt_row, np.float64).reshape(3, 4)
    return T


def run_loop(frames, gt, extract, match, ransac, get_pose, reuse_model=False):
    """frames: sequence of uint8 images or None (unreadable); gt: [n,3,4] or [n,12].  Returns (poses [n,4,4], log) where
    log[i] is one of 'first', 'missing', 'few_matches', 'few_inliers', 'ok', 'ok_stale_model' plus counters."""
    n = len(frames)
    gt4 = [gt_to_mat4(g) for g in gt]
    T_curr = np.eye(4)
    poses = [T_curr.copy()]
    log = [{"status": "first"}]
    model_F, model_inl = None, np.zeros((0, 4))  # the long-lived FundamentalMatrix object (:49)
    kp1, desc1 = extract(frames[0])
    last_valid = 0
    i = 1
    while i < n:
        if frames[i] is None:
            poses.append(T_curr.copy())
            log.append({"status": "missing"})
            i += 1
            continue
        kp2, desc2 = extract(frames[i])
        pairs = match(desc1, desc2)
        entry = {"query": last_valid, "matches": len(pairs)}
        if len(pairs) < 8:
            poses.append(FLIP_Z @ T_curr)
            entry["status"] = "few_matches"
            log.append(entry)
            i += 1
            continue
        pts = np.concatenate([kp1[pairs[:, 0]], kp2[pairs[:, 1]]], axis=1).astype(np.float64)
        F, inl = ransac(pts, i)
        # model state after Ransac::run: refitted when the consensus set has >= 8 pairs, otherwise untouched (reuse)
        # or - the pipeline's fresh model per pair - empty
        stale = False
        if len(inl) >= 8:
            model_F, model_inl = F, inl
        elif reuse_model:
            stale = len(model_inl) >= 8
        else:
            model_F, model_inl = None, np.zeros((0, 4))
        use_inl = model_inl
        entry["inliers"] = len(use_inl)
        if len(use_inl) < 8:
            poses.append(FLIP_Z @ T_curr)
            entry["status"] = "few_inliers"
            log.append(entry)
            i += 1
            continue
        T_rel_gt = np.linalg.inv(gt4[i]) @ gt4[last_valid]
        gt_scale = float(np.linalg.norm(T_rel_gt[:3, 3]))
        last_valid = i
        kp1, desc1 = kp2, desc2
        R, t = get_pose(model_F, use_inl[:, 0:2].astype(np.float32), use_inl[:, 2:4].astype(np.float32), gt_scale)
        T_rel = np.eye(4)
        T_rel[:3, :3], T_rel[:3, 3] = R, np.asarray(t).reshape(3)
        T_curr = T_curr @ T_rel
        poses.append(FLIP_Z @ T_curr)
        entry["status"] = "ok_stale_model" if stale else "ok"
        log.append(entry)
        i += 1
    return np.stack(poses), log


def relative_motion_errors(poses, gt, log=None):
    """Per successful frame i (query frame q from the log, else i-1): the angle (degrees) between the estimated and the
    ground-truth relative rotation, and between the estimated and ground-truth translation directions.
    Conventions follow the code under test: the recorded pose is flipZ * T_curr and T_curr = prod T_rel, so
    T_rel(i) = inv(T_curr(q)) * T_curr(i); its ground-truth counterpart is the motion of camera q expressed in camera i
    or its inverse - both are evaluated and the smaller error reported together with which one matched."""
    P = [FLIP_Z @ p for p in poses]  # undo the flip (flipZ is its own inverse)
    P[0] = np.asarray(poses[0])      # the first pose is recorded unflipped (VisualOdometry.cpp:58-59)
    out = []
    for i in range(1, len(P)):
        if log is not None and not str(log[i].get("status", "")).startswith("ok"):
            continue
        q = log[i]["query"] if log is not None else i - 1
        T_rel = np.linalg.inv(P[q]) @ P[i]
        G = np.linalg.inv(gt_to_mat4(gt[q])) @ gt_to_mat4(gt[i])  # camera i in camera q coordinates
        best = None
        for name, M in (("q_from_i", G), ("i_from_q", np.linalg.inv(G))):
            dR = T_rel[:3, :3].T @ M[:3, :3]
            ang = np.degrees(np.arccos(np.clip((np.trace(dR) - 1) / 2, -1, 1)))
            a, b = T_rel[:3, 3], M[:3, 3]
            na, nb = np.linalg.norm(a), np.linalg.norm(b)
            tdir = np.degrees(np.arccos(np.clip(a @ b / (na * nb), -1, 1))) if na > 0 and nb > 0 else 0.0
            if best is None or ang + tdir < best[1] + best[2]:
                best = (name, ang, tdir)
        out.append((i, q) + best)
    return out
