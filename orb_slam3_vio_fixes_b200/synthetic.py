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


def make_vocabulary(k, L, seed=0, stop_fraction=0.02, desc_pool=None):
    """A synthetic DBoW2-shaped vocabulary (k-ary, depth L, random 256-bit node descriptors, idf-like positive
    weights, a few stopped words with weight 0) in the flat layout, node ids in the breadth-first creation order
    of TemplatedVocabulary::HKmeansStep.  ORBvoc.txt itself is not in the image."""
    rng = np.random.default_rng(seed)
    child_off, child_idx = [0], []
    levels = [[0]]
    nid = 1
    parents = [0]
    children_of = {0: []}
    for lvl in range(1, L + 1):
        cur = []
        for pnode in parents:
            ids = list(range(nid, nid + k))
            nid += k
            children_of[pnode] = ids
            for c in ids:
                children_of[c] = []
            cur += ids
        parents = cur
    nnodes = nid
    for i in range(nnodes):
        child_idx += children_of[i]
        child_off.append(len(child_idx))
    if desc_pool is None:
        node_desc = rng.integers(0, 256, (nnodes, 32), dtype=np.uint8)
    else:  # node centres drawn from real descriptors: features of neighbouring frames share nodes
        pool = np.ascontiguousarray(desc_pool, np.uint8).reshape(-1, 32)
        node_desc = pool[rng.integers(0, pool.shape[0], nnodes)].copy()
    node_word = np.full(nnodes, -1, np.int32)
    leaves = [i for i in range(nnodes) if not children_of[i]]
    node_word[leaves] = np.arange(len(leaves), dtype=np.int32)
    node_weight = np.zeros(nnodes, np.float64)
    w = np.log(rng.uniform(1.5, 400.0, len(leaves)))
    w[rng.random(len(leaves)) < stop_fraction] = 0.0
    node_weight[leaves] = w
    return {"child_off": np.array(child_off, np.int32), "child_idx": np.array(child_idx, np.int32),
            "node_desc": node_desc, "node_word": node_word, "node_weight": node_weight, "L": L}
