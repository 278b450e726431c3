"""oracle/_ref: the REFERENCE's own src/ORBextractor.cc, compiled unmodified over the OpenCV header shim
(oracle/ref_shim, recipe `make -C oracle ref`), against the restated oracle and the committed goldens.

This is what pins oracle/orb_oracle.cpp (and through it the CUDA path) to the reference for rows a1-a11:
the cell loop, DistributeOctTree (std::list surgery + std::sort), IC_Angle, computeOrbDescriptor (with the
FMA contraction g++ actually applies), the lapping-area ordering and ComputePyramid's Mat-sharing tricks are
the reference's code, bit for bit; only the five OpenCV image primitives come from a table — the cv2-pinned
oracle primitives by default, the real cv2 binary in the `cv2` tests below.

Needs either /root/reference (build container) or the prebuilt oracle/_ref/*.so (GPU box); else skipped."""
from pathlib import Path

import numpy as np
import pytest

from orb_slam3_vio_fixes_b200 import synthetic as S

G = Path(__file__).resolve().parent / "golden"


@pytest.fixture(scope="module")
def ref(oracle_mod):
    from oracle import ref as R
    if not R.available():
        pytest.skip("neither /root/reference nor a prebuilt oracle/_ref")
    R.build()
    R.use_oracle_primitives()
    return R


def _same(oracle_mod, R, img, nf, lap, ini=20, mn=7, levels=8, sf=1.2, pyramid=True):
    o = oracle_mod.Extractor(nf, sf, levels, ini, mn)
    r = R.RefExtractor(nf, sf, levels, ini, mn)
    ok, od, om = o.extract(img, lap)
    rk, rd, rm = r.extract(img, lap)
    assert rm == om and len(rk) == len(ok)
    assert np.array_equal(rk, ok), "keypoints: reference build != oracle"
    assert np.array_equal(rd, od), "descriptors: reference build != oracle"
    if pyramid:
        for level in range(levels):
            assert r.level_size(level) == o.level_size(level)
            assert np.array_equal(r.pyramid(level, True), o.pyramid(level, True)), f"mvImagePyramid[{level}]"
    return rk, rd, rm


def test_ctor_tables(oracle_mod, ref):
    """ORBextractor.cc:409-469: scale tables, per-level quotas, umax — reference ctor vs oracle vs SURVEY §8."""
    for nf, sf, levels in ((1000, 1.2, 8), (1200, 1.2, 8), (2000, 1.2, 8), (5000, 1.2, 8), (500, 1.5, 4), (777, 1.1, 12)):
        o, r = oracle_mod.Extractor(nf, sf, levels, 20, 7), ref.RefExtractor(nf, sf, levels, 20, 7)
        for a in ("scale_factors", "inv_scale_factors", "level_sigma2", "inv_level_sigma2", "features_per_level", "umax"):
            assert np.array_equal(getattr(o, a), getattr(r, a)), (nf, sf, levels, a)
    r = ref.RefExtractor(1000, 1.2, 8, 20, 7)
    assert list(r.features_per_level) == [217, 181, 151, 126, 105, 87, 73, 60]
    assert list(r.umax) == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]


@pytest.mark.parametrize("cfg,lap", [("C1", (0, 1000)), ("C2", (0, 0)), ("C3", (0, 0)), ("C4", (0, 0)), ("C5", (0, 1000))])
def test_reference_build_equals_oracle_on_configs(oracle_mod, ref, cfg, lap):
    w, h, nf, seed, stereo = S.CONFIGS[cfg]
    n = 1 if cfg == "C5" else 2
    for right in ((False, True) if stereo else (False,)):
        for img in S.make_frames(cfg, n, right=right):
            _same(oracle_mod, ref, img, nf, lap, pyramid=(cfg != "C5"))


@pytest.mark.parametrize("ini,mn,contrast,size", [
    (20, 7, 0.25, (752, 480)), (20, 7, 0.08, (640, 480)), (7, 20, 1.0, (400, 300)), (40, 40, 1.0, (752, 480)),
    (0, 0, 0.05, (320, 240)), (130, 127, 1.0, (752, 480)), (255, 1, 1.0, (320, 240)), (20, 7, 1.0, (501, 317)),
])
def test_reference_build_equals_oracle_on_threshold_cases(oracle_mod, ref, ini, mn, contrast, size):
    """The FAST threshold / fallback cases of tests/test_gpu_extract.py::test_fast_thresholds_and_fallback."""
    w, h = size
    img = S.make_frames("C1", 1, width=w, height=h)[0].astype(np.float64)
    img = np.clip((img - 128.0) * contrast + 128.0, 0, 255).astype(np.uint8)
    _same(oracle_mod, ref, img, 800, (0, 0), ini, mn)


def test_reference_build_edge_cases(oracle_mod, ref):
    r = ref.RefExtractor(500, 1.2, 8, 20, 7)
    k, d, mono = r.extract(np.zeros((0, 0), np.uint8))
    assert mono == -1 and len(k) == 0                          # ORBextractor.cc:1090-1091
    k, d, mono = r.extract(np.full((240, 320), 77, np.uint8))
    assert mono == 0 and len(k) == 0 and d.shape == (0, 32)    # :1108-1109
    # partial lapping area, odd size, other pyramid shapes
    _same(oracle_mod, ref, S.make_frames("C1", 1, width=411, height=307)[0], 300, (100, 250))
    _same(oracle_mod, ref, S.make_frames("C3", 1, width=400, height=300)[0], 500, (0, 0), levels=4, sf=1.5)
    _same(oracle_mod, ref, S.make_frames("C3", 1, width=640, height=480)[0], 5000, (0, 1000))  # mono init extractor


@pytest.mark.parametrize("name", ["c1_mono.npz", "c3_rgbd.npz", "small_lap.npz"])
def test_reference_build_equals_goldens(ref, name):
    g = np.load(G / name, allow_pickle=False)
    r = ref.RefExtractor(int(g["nfeatures"]))
    lap = tuple(int(v) for v in g["lap"])
    for i, img in enumerate(g["frames"]):
        k, d, m = r.extract(img, lap)
        assert m == int(g[f"mono{i}"]) and np.array_equal(k, g[f"kps{i}"]) and np.array_equal(d, g[f"desc{i}"])
        if i == 0:
            assert np.array_equal(r.pyramid(3, True), g["pyr3_border"]) and np.array_equal(r.pyramid(7), g["pyr7"])


def test_reference_build_equals_stereo_golden(ref):
    g = np.load(G / "c2_stereo.npz", allow_pickle=False)
    for eye, tag in (("left", "l"), ("right", "r")):
        k, d, _ = ref.RefExtractor(1200).extract(g[eye][0], (0, 0))
        assert np.array_equal(k, g[f"kps_{tag}"]) and np.array_equal(d, g[f"desc_{tag}"])


def test_fma_contract_is_what_the_reference_build_does(oracle_mod, ref):
    """g++ contracts GET_VALUE (ORBextractor.cc:117-119) to FMA under the reference's flags.  The
    reference build must agree with the oracle's default (FMA) mode on every descriptor byte, and the strict
    (uncontracted) mode must be the one that differs wherever the two modes differ at all."""
    differ = 0
    for t0 in (0, 7, 19):
        for img in S.make_frames("C1", 2, t0=t0):
            rk, rd, _ = ref.RefExtractor(1000).extract(img, (0, 1000))
            ok, od, _ = oracle_mod.Extractor(1000).extract(img, (0, 1000))
            sk, sd, _ = oracle_mod.Extractor(1000, flags=oracle_mod.STRICT_FP).extract(img, (0, 1000))
            assert np.array_equal(rd, od) and np.array_equal(rk, ok)
            assert np.array_equal(sk, ok)
            differ += int((sd != od).any(axis=1).sum())
    print("descriptors where strict-fp differs from the FMA build:", differ)


# ---- the same reference code over the REAL OpenCV binary (cv2 callbacks) ---------------------------------------
@pytest.fixture()
def ref_cv2(ref):
    pytest.importorskip("cv2")
    ref.use_cv2_primitives()
    yield ref
    ref.use_oracle_primitives()


@pytest.mark.parametrize("cfg,lap,size", [("C1", (0, 1000), None), ("C3", (0, 0), None), ("C4", (0, 0), None),
                                          ("C1", (100, 250), (411, 307))])
def test_reference_over_real_opencv_equals_oracle(oracle_mod, ref_cv2, cfg, lap, size):
    w, h, nf, seed, _ = S.CONFIGS[cfg]
    if size:
        w, h = size
    for img in S.make_frames(cfg, 2, width=w, height=h):
        _same(oracle_mod, ref_cv2, img, nf, lap)


def test_reference_over_real_opencv_equals_goldens(ref_cv2):
    g = np.load(G / "c1_mono.npz", allow_pickle=False)
    r = ref_cv2.RefExtractor(1000)
    for i, img in enumerate(g["frames"]):
        k, d, m = r.extract(img, (0, 1000))
        assert m == int(g[f"mono{i}"]) and np.array_equal(k, g[f"kps{i}"]) and np.array_equal(d, g[f"desc{i}"])


def test_roi_input_border_comes_from_parent(ref_cv2):
    """ORBextractor.cc:1190: level 0's copyMakeBorder is not ISOLATED, so a ROI input takes its 19-px border from
    the parent Mat.  Keypoints/descriptors never read the border; only mvImagePyramid[0]'s frame differs."""
    import cv2
    parent = S.make_frames("C3", 1, width=500, height=400)[0]
    x, y, w, h = 30, 10, 400, 300           # 10 px of parent above (< 19): the rest is reflected
    r = ref_cv2.RefExtractor(500)
    kr, dr, _ = r.extract(parent, (0, 0), roi=(x, y, w, h))
    got = r.pyramid(0, True)
    # cv::copyMakeBorder on a ROI: widen the source by what the parent has (10 rows above, 19 elsewhere), then
    # extrapolate only the 9 missing rows from the widened source
    want = cv2.copyMakeBorder(parent[y - 10:y + h + 19, x - 19:x + w + 19].copy(), 9, 0, 0, 0, cv2.BORDER_REFLECT_101)
    assert np.array_equal(got, want)
    assert not np.array_equal(got, cv2.copyMakeBorder(parent[y:y + h, x:x + w].copy(), 19, 19, 19, 19, cv2.BORDER_REFLECT_101))
    ki, di, _ = ref_cv2.RefExtractor(500).extract(parent[y:y + h, x:x + w].copy(), (0, 0))
    assert np.array_equal(kr, ki) and np.array_equal(dr, di)
