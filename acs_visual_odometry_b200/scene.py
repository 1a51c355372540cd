"""Seeded synthetic sequences WITH two-view geometry: a textured corridor rendered by ray casting, and its ground truth.

The shifted-canvas generator in synth.py exercises the front end but has no epipolar geometry (a 2-D shift is not a rigid
camera motion in front of a 3-D scene), so it cannot validate poses.  This one can: a pinhole camera with the KITTI
intrinsics the reference hard-wires (PoseUpdate.hpp:36-39) moves through a static scene of textured planes - ground,
two side walls, an end wall and frontal billboards at assorted depths - and every frame is rendered per pixel
(nearest ray/plane intersection, bilinear texture lookup, seeded Gaussian noise).  Pure numpy, no renderer.

Ground truth is written in the KITTI odometry format the reference reads (PoseUpdate.cpp:43-50): one line of 12 numbers
per frame, the row-major 3x4 matrix [R | p] that maps camera coordinates (x right, y down, z forward) to the world.
"""
import numpy as np

K_KITTI = np.array([[718.856, 0.0, 607.1928], [0.0, 718.856, 185.2157], [0.0, 0.0, 1.0]])
KITTI = (376, 1241)


def _texture(rng, h, w, n_rects, lo=6, hi=40):
    tex = np.full((h, w), 128, np.float32)
    ys, xs = rng.integers(0, h, n_rects), rng.integers(0, w, n_rects)
    hs, ws = rng.integers(lo, hi, n_rects), rng.integers(lo, hi, n_rects)
    gs = rng.integers(0, 256, n_rects)
    for y, x, hh, ww, g in zip(ys, xs, hs, ws, gs):
        tex[y:y + hh, x:x + ww] = g
    return tex


def _bilinear(tex, u, v):
    h, w = tex.shape
    u = np.mod(u, w - 1.0)
    v = np.mod(v, h - 1.0)
    u0, v0 = np.floor(u).astype(np.int64), np.floor(v).astype(np.int64)
    fu, fv = (u - u0).astype(np.float32), (v - v0).astype(np.float32)
    a, b = tex[v0, u0], tex[v0, u0 + 1]
    c, d = tex[v0 + 1, u0], tex[v0 + 1, u0 + 1]
    return (a * (1 - fu) + b * fu) * (1 - fv) + (c * (1 - fu) + d * fu) * fv


def camera_pose(t, step=0.8):
    """(R, p) of frame t: camera-to-world rotation and camera centre.  Forward motion of `step` metres per frame with a
    slow weave, yaw and a little pitch, so that consecutive frames differ by a general rigid motion."""
    yaw = 0.06 * np.sin(0.11 * t)
    pitch = 0.01 * np.sin(0.17 * t + 1.0)
    cy, sy, cp, sp = np.cos(yaw), np.sin(yaw), np.cos(pitch), np.sin(pitch)
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rx = np.array([[1, 0, 0], [0, cp, -sp], [0, sp, cp]])
    p = np.array([0.4 * np.sin(0.09 * t), 0.03 * np.sin(0.13 * t), step * t])
    return Ry @ Rx, p


class Scene:
    """Planes: (axis, offset, bounds on the two other axes or None, texture, pixels per metre)."""

    def __init__(self, seed=0, length=400.0):
        rng = np.random.default_rng([seed, 77])
        self.planes = []
        ppm = 40.0
        big = lambda: _texture(rng, 1024, 4096, 5000)  # noqa: E731
        self.planes.append((1, 1.65, None, big(), ppm))              # ground, y = +1.65 (y points down)
        self.planes.append((0, -6.0, None, big(), ppm))              # left wall
        self.planes.append((0, 7.0, None, big(), ppm))               # right wall
        self.planes.append((1, -4.5, None, big(), ppm))              # ceiling
        self.planes.append((2, length, None, _texture(rng, 1024, 1024, 1500), 20.0))  # end wall
        # frontal billboards standing on the ground at assorted depths, left and right of the path
        z = 12.0
        while z < length - 10:
            side = rng.integers(0, 2)
            x0 = -5.5 + rng.random() * 2.0 if side == 0 else 2.0 + rng.random() * 2.5
            w, h = 1.5 + rng.random() * 2.0, 1.5 + rng.random() * 2.5
            self.planes.append((2, z, (x0, x0 + w, 1.65 - h, 1.65), _texture(rng, 256, 256, 120, 4, 24), 48.0))
            z += 3.0 + rng.random() * 5.0

    def render(self, R, p, rows=KITTI[0], cols=KITTI[1], K=K_KITTI):
        v, u = np.mgrid[0:rows, 0:cols].astype(np.float64)
        d = np.stack([(u - K[0, 2]) / K[0, 0], (v - K[1, 2]) / K[1, 1], np.ones_like(u)], -1) @ R.T  # world directions
        best = np.full((rows, cols), np.inf)
        img = np.full((rows, cols), 128.0, np.float32)
        for axis, off, bounds, tex, ppm in self.planes:
            with np.errstate(divide="ignore", invalid="ignore"):
                s = (off - p[axis]) / d[..., axis]
            o1, o2 = [a for a in range(3) if a != axis]
            h1, h2 = p[o1] + s * d[..., o1], p[o2] + s * d[..., o2]
            ok = (s > 0.05) & (s < best) & np.isfinite(s)
            if bounds is not None:
                ok &= (h1 >= bounds[0]) & (h1 <= bounds[1]) & (h2 >= bounds[2]) & (h2 <= bounds[3])
            if not ok.any():
                continue
            # texture coordinates: the two in-plane world axes, in texture pixels (z first on the long planes)
            tu, tv = (h2, h1) if axis != 2 else (h1, h2)
            val = _bilinear(tex, tu[ok] * ppm + 4096.0, tv[ok] * ppm + 4096.0)
            img[ok] = val
            best[ok] = s[ok]
        return img


def make_scene_sequence(n_frames, rows=KITTI[0], cols=KITTI[1], seed=0, step=0.8, noise_sigma=1.5):
    """Returns (frames uint8 [n, rows, cols], gt [n, 3, 4]): gt[t] = [R_t | p_t], camera-to-world, KITTI convention."""
    scene = Scene(seed, length=step * n_frames + 60.0)
    frames = np.empty((n_frames, rows, cols), np.uint8)
    gt = np.empty((n_frames, 3, 4))
    for t in range(n_frames):
        R, p = camera_pose(t, step)
        gt[t, :, :3], gt[t, :, 3] = R, p
        img = scene.render(R, p, rows, cols)
        rng = np.random.default_rng([seed, 1000 + t])
        img = img + rng.standard_normal((rows, cols), dtype=np.float32) * np.float32(noise_sigma)
        frames[t] = np.clip(np.rint(img), 0, 255).astype(np.uint8)
    return frames, gt


def write_gt_file(path, gt):
    """KITTI poses file: 12 numbers per line (PoseUpdate.cpp:43-50 reads them back)."""
    with open(path, "w") as f:
        for T in gt:
            f.write(" ".join(f"{x:.9e}" for x in T.reshape(-1)) + "\n")
