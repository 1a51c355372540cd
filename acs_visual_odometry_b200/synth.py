"""Seeded synthetic frame sequences (SURVEY.md section 8d, config 2-5 generator).

Background 128, random axis-aligned grey rectangles on an oversized canvas, and per-frame
Gaussian noise (sigma 2).  Frame t is the canvas window shifted by (2t mod 32, t mod 16) px so
that consecutive frames overlap and matches between them are meaningful.
"""
import numpy as np

KITTI = (376, 1241)
FULL_HD = (1080, 1920)
UHD_4K = (2160, 3840)


def make_canvas(rows, cols, seed=0, n_rects=None):
    rng = np.random.default_rng(seed)
    ch, cw = rows + 16, cols + 32
    if n_rects is None:
        n_rects = int(round(1200 * (rows * cols) / float(KITTI[0] * KITTI[1])))
    canvas = np.full((ch, cw), 128, np.uint8)
    ys = rng.integers(0, ch, n_rects)
    xs = rng.integers(0, cw, n_rects)
    hs = rng.integers(6, 61, n_rects)
    ws = rng.integers(6, 61, n_rects)
    gs = rng.integers(0, 256, n_rects)
    for y, x, h, w, g in zip(ys, xs, hs, ws, gs):
        canvas[y:y + h, x:x + w] = g
    return canvas


def make_sequence(n_frames, rows=KITTI[0], cols=KITTI[1], seed=0, n_rects=None, noise_sigma=2.0, out=None, start=0):
    """Returns a (n_frames, rows, cols) uint8 array (C-contiguous): frames start .. start+n_frames-1 of the sequence
    (a shard of a longer sequence is generated with `start` = its first frame and equals the slice of the whole)."""
    canvas = make_canvas(rows, cols, seed, n_rects).astype(np.float32)
    frames = out if out is not None else np.empty((n_frames, rows, cols), np.uint8)
    for k in range(n_frames):
        t = start + k
        dx, dy = (2 * t) % 32, t % 16
        win = canvas[dy:dy + rows, dx:dx + cols]
        rng = np.random.default_rng([seed, t + 1])
        noise = rng.standard_normal((rows, cols), dtype=np.float32) * np.float32(noise_sigma)
        np.clip(np.rint(win + noise), 0, 255, out=noise)
        frames[k] = noise.astype(np.uint8)
    return frames
