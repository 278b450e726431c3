"""CPU checks of the product's host/device-shared code (orbx_core.h, orbx_plan.h, octree_core.h)
compiled for the host by tests/cpu_harness, against the oracle."""
import ctypes as C

import numpy as np
import pytest

from orb_slam3_vio_fixes_b200 import synthetic as S


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def test_sincos_matches_libm_exhaustively(harness):
    """orbx_sincosf == glibc sincosf for every float in [0, 6.3] (all angles the descriptor forms)."""
    hi = int(np.float32(6.3).view(np.uint32))
    assert harness.h_sincos_mismatches(0, hi, 1) == 0


def test_logf_matches_libm_exhaustively(harness):
    """orbx_logf == glibc logf on the distance ratios PredictScale can see: every float in [2^-20, 2^20]
    (the sweep over ALL positive normal floats, 0 differences, takes 30 s single-threaded: stride 1 here on the
    band, stride 7 outside it)."""
    lo, hi = int(np.float32(2.0 ** -20).view(np.uint32)), int(np.float32(2.0 ** 20).view(np.uint32))
    assert harness.h_logf_mismatches(lo, hi, 1) == 0
    assert harness.h_logf_mismatches(0x00800000, lo, 7) == 0
    assert harness.h_logf_mismatches(hi, 0x7f7fffff, 7) == 0


def test_fast_atan2_matches_oracle(harness, oracle_mod):
    rng = np.random.default_rng(0)
    y = rng.integers(-3000000, 3000000, 300000).astype(np.float32)
    x = rng.integers(-3000000, 3000000, 300000).astype(np.float32)
    y[:4] = [0, 0, 7, -7]
    x[:4] = [0, -3, 0, 0]
    out = np.zeros_like(y)
    harness.h_fast_atan2(_p(y), _p(x), _p(out), y.size)
    assert np.array_equal(out, oracle_mod.fast_atan2(y, x))


def test_arc_strength_matches_oracle_fast(harness, oracle_mod):
    """orbx_arc_strength on random rings == score+1 of the oracle's FAST at threshold 0."""
    rng = np.random.default_rng(1)
    ring_xy = [(0, 3), (1, 3), (2, 2), (3, 1), (3, 0), (3, -1), (2, -2), (1, -3), (0, -3), (-1, -3), (-2, -2),
               (-3, -1), (-3, 0), (-3, 1), (-2, 2), (-1, 3)]
    n_checked = 0
    for trial in range(300):
        img = np.full((7, 7), 0, np.uint8)
        base = rng.integers(0, 256)
        vals = np.clip(base + rng.integers(-60, 60, 16), 0, 255).astype(np.int32)
        if trial % 3 == 0:  # force a clean bright/dark arc
            k0 = rng.integers(0, 16)
            for j in range(rng.integers(8, 12)):
                vals[(k0 + j) % 16] = np.clip(base + (70 if trial % 2 else -70) + rng.integers(-5, 5), 0, 255)
        for (dx, dy), v in zip(ring_xy, vals):
            img[3 + dy, 3 + dx] = v
        img[3, 3] = base
        A = harness.h_arc_strength(int(base), _p(np.ascontiguousarray(vals, np.int32)))
        got = oracle_mod.fast(img, (0, 0, 7, 7), 0)
        if A >= 2:
            assert len(got) == 1 and got[0, 2] == A - 1, (trial, A, got)
            n_checked += 1
        else:
            assert len(got) == 0
    assert n_checked > 50


@pytest.mark.parametrize("size,nf", [((752, 480), 1000), ((640, 480), 1000), ((1241, 376), 2000), ((3840, 2160), 5000),
                                     ((320, 240), 77)])
def test_plan_matches_oracle_tables(harness, oracle_mod, size, nf):
    w, h = size
    out9 = np.zeros((8, 9), np.int32)
    scales = np.zeros((8, 4), np.float32)
    umax = np.zeros(16, np.int32)
    totals = np.zeros(4, np.int32)
    harness.h_plan.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 4
    assert harness.h_plan(nf, 1.2, 8, 20, 7, w, h, _p(out9), _p(scales), _p(umax), _p(totals)) == 0
    e = oracle_mod.Extractor(nf)
    assert np.array_equal(out9[:, 6], e.features_per_level)
    assert np.array_equal(umax, e.umax)
    assert np.array_equal(scales[:, 0], e.scale_factors) and np.array_equal(scales[:, 1], e.inv_scale_factors)
    assert np.array_equal(scales[:, 2], e.level_sigma2) and np.array_equal(scales[:, 3], e.inv_level_sigma2)
    if w * h <= 1241 * 480:
        e.extract(np.zeros((h, w), np.uint8), (0, 0))
        assert [tuple(r) for r in out9[:, :2]] == [e.level_size(l) for l in range(8)]
    # max keypoints a frame can return: sum of max(N_l+3, 4*nIni_l)
    assert totals[3] == sum(max(int(n) + 3, 4 * int(r)) for n, r in zip(out9[:, 6], out9[:, 7]))
    assert totals[3] >= nf + 3 * 8 and (nf < 500 or totals[3] == nf + 3 * 8)


def test_plan_rejects_unsupported(harness):
    harness.h_plan.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 4
    bufs = [np.zeros(200, np.int32), np.zeros(200, np.float32), np.zeros(16, np.int32), np.zeros(4, np.int32)]
    args = [_p(b) for b in bufs]
    assert harness.h_plan(1000, 1.2, 8, 20, 7, 200, 150, *args) != 0   # level 7 smaller than one cell
    assert harness.h_plan(1000, 1.2, 8, 20, 7, 300, 900, *args) != 0   # portrait: nIni == 0
    assert harness.h_plan(1000, 2.0, 2, 20, 7, 640, 480, *args) != 0   # exact 2x: OpenCV INTER_AREA path
    assert harness.h_plan(1000, 1.2, 13, 20, 7, 4000, 3000, *args) != 0  # too many levels


def _distribute_both(harness, oracle_mod, xys, width, height, N):
    out = np.zeros((len(xys) + 64, 3), np.int32)
    k = harness.h_distribute(_p(xys), len(xys), width, height, N, _p(out))
    # the array ("scan") form the CUDA kernel runs must give the same list
    out2 = np.zeros((len(xys) + 64, 3), np.int32)
    k2 = harness.h_distribute_scan(_p(xys), len(xys), width, height, N, _p(out2))
    assert k2 == k and np.array_equal(out2[:max(k2, 0)], out[:max(k, 0)]), ("scan form differs", N, k, k2)
    ref = oracle_mod.distribute(xys.astype(np.float32), 16, 16 + width, 16, 16 + height, N).astype(np.int32)
    return out[:max(k, 0)], ref, k


def test_octree_rounds_on_real_candidates(harness, oracle_mod):
    e = oracle_mod.Extractor(1000)
    for img in S.make_frames("C1", 2):
        e.extract(img, (0, 0))
        for level in range(8):
            c = e.level_keypoints(level, 0)
            w, h = e.level_size(level)
            xys = np.ascontiguousarray(np.stack([c["x"], c["y"], c["response"]], 1).astype(np.int32))
            got, ref, k = _distribute_both(harness, oracle_mod, xys, w - 32, h - 32, int(e.features_per_level[level]))
            assert k == len(ref) and np.array_equal(got, ref), level


@pytest.mark.parametrize("seed", range(12))
def test_octree_rounds_random_stress(harness, oracle_mod, seed):
    """Random key sets with many response ties and clustered points: exercises the careful phase,
    the std::sort replica (ties in (count, UL.x)), single-child chains and tiny/huge N."""
    rng = np.random.default_rng(seed)
    width, height = [(720, 448), (595, 368), (1209, 344), (178, 102), (400, 400), (3808, 2128)][seed % 6]
    n = int(rng.integers(1, 4000))
    if seed % 4 == 0:   # clustered
        cx, cy = rng.integers(0, width, 6), rng.integers(0, height, 6)
        pick = rng.integers(0, 6, n)
        x = np.clip(cx[pick] + rng.integers(-25, 25, n), 0, width - 5)
        y = np.clip(cy[pick] + rng.integers(-25, 25, n), 0, height - 5)
    else:
        x, y = rng.integers(0, width - 4, n), rng.integers(0, height - 4, n)
    # distinct pixels like real FAST output
    flat = np.unique(y.astype(np.int64) * 8192 + x)
    rng.shuffle(flat)
    xys = np.ascontiguousarray(np.stack([flat % 8192, flat // 8192, rng.integers(7, 40, len(flat))], 1).astype(np.int32))
    for N in [1, 5, 60, 217, 1086, len(xys) + 10]:
        got, ref, k = _distribute_both(harness, oracle_mod, xys, width, height, N)
        assert k == len(ref), (N, k, len(ref))
        assert np.array_equal(got, ref), N


def test_octree_regular_grid_many_ties(harness, oracle_mod):
    """A regular lattice gives many nodes with equal (count, UL.x): the order among those ties is
    decided by libstdc++'s introsort, which oct_std_sort must replay (n > 16 → median-of-3 path)."""
    xs, ys = np.meshgrid(np.arange(4, 700, 9), np.arange(4, 440, 9))
    xys = np.ascontiguousarray(np.stack([xs.ravel(), ys.ravel(), np.full(xs.size, 20)], 1).astype(np.int32))
    for N in [100, 217, 500, 900, 1500]:
        got, ref, k = _distribute_both(harness, oracle_mod, xys, 720, 448, N)
        assert k == len(ref) and np.array_equal(got, ref), N
