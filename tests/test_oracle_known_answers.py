"""Known answers for the oracle that do not need cv2 (SURVEY.md §8 constants table)."""
import numpy as np
import pytest

LEVELS_752 = [(752, 480), (627, 400), (522, 333), (435, 278), (363, 231), (302, 193), (252, 161), (210, 134)]
LEVELS_640 = [(640, 480), (533, 400), (444, 333), (370, 278), (309, 231), (257, 193), (214, 161), (179, 134)]
LEVELS_1241 = [(1241, 376), (1034, 313), (862, 261), (718, 218), (598, 181), (499, 151), (416, 126), (346, 105)]


@pytest.mark.parametrize("nf,expected", [
    (1000, [217, 181, 151, 126, 105, 87, 73, 60]),
    (1200, [261, 217, 181, 151, 126, 105, 87, 72]),
    (2000, [434, 362, 302, 251, 209, 175, 145, 122]),
    (5000, [1086, 905, 754, 628, 524, 436, 364, 303]),
])
def test_features_per_level(oracle_mod, nf, expected):
    e = oracle_mod.Extractor(nf)
    assert list(e.features_per_level) == expected
    assert sum(expected) == nf


def test_umax(oracle_mod):
    e = oracle_mod.Extractor(1000)
    assert list(e.umax) == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]


@pytest.mark.parametrize("size,levels", [((752, 480), LEVELS_752), ((640, 480), LEVELS_640), ((1241, 376), LEVELS_1241)])
def test_level_sizes(oracle_mod, size, levels):
    e = oracle_mod.Extractor(300)
    rng = np.random.default_rng(0)
    e.extract(rng.integers(0, 256, (size[1], size[0]), dtype=np.uint8), (0, 0))
    assert [e.level_size(l) for l in range(8)] == levels


def test_patch_sizes_and_scale_tables(oracle_mod):
    from orb_slam3_vio_fixes_b200 import synthetic as S
    e = oracle_mod.Extractor(1000)
    k, d, m = e.extract(S.make_frames("C1", 1)[0], (0, 0))
    sizes = {int(o): float(s) for o, s in zip(k["octave"], k["size"])}
    assert sizes == {0: 31, 1: 37, 2: 44, 3: 53, 4: 64, 5: 77, 6: 92, 7: 111}
    assert e.scale_factors[0] == 1 and abs(e.scale_factors[7] - 1.2 ** 7) < 1e-5
    assert np.allclose(e.inv_scale_factors * e.scale_factors, 1, atol=1e-6)
    assert len(k) >= 1000 and m == len(k)  # lapping {0,0}: nothing at x==0, all "mono"
    assert np.all(k["class_id"] == -1) and np.all((k["angle"] >= 0) & (k["angle"] < 360))


def test_descriptor_distance(oracle_mod):
    rng = np.random.default_rng(0)
    x = rng.integers(0, 256, 32, dtype=np.uint8)
    assert oracle_mod.descriptor_distance(x, x) == 0
    assert oracle_mod.descriptor_distance(x, ~x) == 256
    y = rng.integers(0, 256, 32, dtype=np.uint8)
    assert oracle_mod.descriptor_distance(x, y) == int(np.unpackbits(x ^ y).sum())


def test_gauss_taps_impulse(oracle_mod):
    img = np.zeros((31, 31), np.uint8)
    img[15, 15] = 255
    out = oracle_mod.gaussian_blur7(img)
    taps = np.array([18, 34, 48, 56, 48, 34, 18])
    expect = (np.outer(taps, taps) * 255 + 32768) >> 16
    assert np.array_equal(out[12:19, 12:19], expect)


def test_mono_lapping_quirk(oracle_mod):
    """lapping area {0,1000} on a 752-wide image sends every keypoint to the back in reverse
    order and returns monoIndex 0 (ORBextractor.cc:1153-1162)."""
    from orb_slam3_vio_fixes_b200 import synthetic as S
    img = S.make_frames("C1", 1)[0]
    e = oracle_mod.Extractor(1000)
    k0, d0, m0 = e.extract(img, (0, 0))
    k1, d1, m1 = e.extract(img, (0, 1000))
    assert m1 == 0 and m0 == len(k0)
    assert np.array_equal(k1, k0[::-1]) and np.array_equal(d1, d0[::-1])


def test_argmin2_semantics(oracle_mod):
    """first-min / second-best / last-min rules of the Search* loops."""
    q = np.zeros((1, 32), np.uint8)
    r = np.zeros((5, 32), np.uint8)
    r[0, 0] = 0b111      # dist 3
    r[1, 0] = 0b1        # dist 1
    r[2, 0] = 0b10       # dist 1 (tie with 1)
    r[3, 0] = 0b11       # dist 2
    r[4, 0] = 0b1111     # dist 4
    lv = np.array([0, 1, 2, 3, 4], np.int8)
    off = np.array([0, 5], np.int32)
    idx = np.array([0, 1, 2, 3, 4], np.int32)
    res = oracle_mod.argmin2_csr(q, r, off, idx, lv, 0)[0]
    assert (res["best_idx"], res["best_dist"], res["second_dist"], res["best_level"], res["second_level"]) == (1, 1, 1, 1, 2)
    res = oracle_mod.argmin2_csr(q, r, off, idx, lv, 1)[0]
    assert (res["best_idx"], res["best_dist"]) == (2, 1)
    res = oracle_mod.argmin2_csr(q, r, np.array([0, 0], np.int32), idx, lv, 0)[0]
    assert (res["best_idx"], res["best_dist"], res["second_dist"], res["best_level"]) == (-1, 256, 256, -1)


def test_search_by_projection_lock_semantics(oracle_mod):
    """ORBmatcher.cc:1741-1743 / :88-90: a feature written by a MapPoint with Observations()>0 is invisible to
    later queries; one written by an unobserved point can be overwritten (both matches still count)."""
    from orb_slam3_vio_fixes_b200 import _lib
    k = np.zeros(3, _lib.KP_DTYPE)
    k["x"], k["y"], k["octave"], k["angle"] = [100, 101, 102], [50, 50, 50], 0, [10, 10, 10]
    d = np.zeros((3, 32), np.uint8)
    d[1, 0] = 0x01          # distance 1 to the zero descriptor
    d[2, 0] = 0x07          # distance 3
    qd = np.zeros((2, 32), np.uint8)
    common = dict(qx=[101, 101], qy=[50, 50], qr=[5, 5], qmin=[0, 0], qmax=[-1, -1], qdesc=qd,
                  q_angle=[10, 10], check_orientation=True)
    bounds = (0.0, 640.0, 0.0, 480.0)
    m, fq, nm = oracle_mod.search_by_projection(k, d, bounds, mode=1, q_flags=[3, 3], **common)
    assert list(m) == [0, 1] and list(fq) == [0, 1, -1] and nm == 2
    m, fq, nm = oracle_mod.search_by_projection(k, d, bounds, mode=1, q_flags=[1, 3], **common)
    assert list(m) == [0, 0] and list(fq) == [1, -1, -1] and nm == 2
    m, fq, nm = oracle_mod.search_by_projection(k, d, bounds, mode=1, q_flags=[3, 0], **common)
    assert list(m) == [0, -1] and nm == 1
    # mode 0: ratio test uses the second best AFTER the lock: q1 sees {1 (d=1), 2 (d=3)}: 1 > 0.3*3 -> rejected
    m, fq, nm = oracle_mod.search_by_projection(k, d, bounds, mode=0, nn_ratio=0.3, q_flags=[3, 3], **common)
    assert list(m) == [0, -1] and nm == 1          # q0: best 0 (d=0), 0 > 0.3*1 is false -> accepted
    m, fq, nm = oracle_mod.search_by_projection(k, d, bounds, mode=0, nn_ratio=0.4, q_flags=[3, 3], **common)
    assert list(m) == [0, 1] and nm == 2
    # rotation histogram: a lone match in another bin than the dominant ones survives only if >= 10 % of max
    m, fq, nm = oracle_mod.search_by_projection(k, d, bounds, mode=1, q_flags=[3, 3], feat_locked=[0, 0, 1],
                                                **{**common, "q_angle": [10, 100]})
    assert nm == 2 and list(m) == [0, 1]           # bins 0 and 3, one entry each: both within the top three
