"""Seeded synthetic frames for the parity tests and bench.py.

Integer-only (no cv2, no float maths) so that the same bytes come out on every box:
a canvas of random filled rectangles, a 3x3 binomial blur, then per-frame crops that
drift by a known shift plus small integer noise.  The known shift is what the matcher
workloads use to "project" keypoints from frame t-1 into frame t.
"""
import numpy as np

CONFIGS = {
    # name: (width, height, nfeatures, seed, stereo)
    "C1": (752, 480, 1000, 11, False),   # EuRoC mono shape
    "C2": (752, 480, 1200, 22, True),    # EuRoC stereo-inertial shape
    "C3": (640, 480, 1000, 33, False),   # RealSense D435 shape
    "C4": (1241, 376, 2000, 44, True),   # KITTI shape
    "C5": (3840, 2160, 5000, 55, False),  # 4K stress
}
STEREO_DISPARITY = 24
MARGIN = 128


def make_canvas(width, height, seed):
    rng = np.random.default_rng(seed)
    ch, cw = height + 2 * MARGIN, width + 2 * MARGIN
    canvas = np.full((ch, cw), 128, np.int32)
    n = int(round(600.0 * width * height / 360960.0))
    ys = rng.integers(0, ch, n)
    xs = rng.integers(0, cw, n)
    hs = rng.integers(8, 80, n)
    ws = rng.integers(8, 80, n)
    vs = rng.integers(0, 256, n)
    for y, x, h, w, v in zip(ys, xs, hs, ws, vs):
        canvas[y:y + h, x:x + w] = v
    # 3x3 binomial blur [1 2 1]x[1 2 1]/16, edges replicated, round to nearest
    p = np.pad(canvas, 1, mode="edge")
    hsum = p[:, :-2] + 2 * p[:, 1:-1] + p[:, 2:]
    canvas = (hsum[:-2] + 2 * hsum[1:-1] + hsum[2:] + 8) >> 4
    return canvas


def frame_shift(t):
    """(dx, dy) crop offset of frame t inside the canvas margin."""
    return (2 * t) % 64, t % 32


def make_frame(canvas, width, height, seed, t, right=False):
    dx, dy = frame_shift(t)
    x0 = MARGIN + dx + (STEREO_DISPARITY if right else 0)
    y0 = MARGIN + dy
    rng = np.random.default_rng(seed * 1000 + t + (500000 if right else 0))
    crop = canvas[y0:y0 + height, x0:x0 + width]
    noise = rng.integers(-2, 3, crop.shape) + rng.integers(-2, 3, crop.shape)
    return np.clip(crop + noise, 0, 255).astype(np.uint8)


def make_frames(config, n_frames, t0=0, right=False, width=None, height=None):
    """n_frames frames [n,h,w] uint8 of a named config (or explicit width/height)."""
    w, h, _, seed, _ = CONFIGS[config]
    w = width or w
    h = height or h
    canvas = make_canvas(w, h, seed)
    return np.stack([make_frame(canvas, w, h, seed, t0 + i, right) for i in range(n_frames)])
