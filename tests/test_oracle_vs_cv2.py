"""Pins the oracle (oracle/orb_oracle.cpp) against the REAL OpenCV binary (cv2), primitive by
primitive and end to end.  The reference ships no tests or golden vectors for this path, so this
is what anchors parity.  Skipped where cv2 is not importable."""
import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")

from orb_slam3_vio_fixes_b200 import synthetic as S  # noqa: E402


def _images(rng):
    noise = rng.integers(0, 256, (203, 311), dtype=np.uint8)
    blocks = np.kron(rng.integers(0, 256, (26, 39), dtype=np.uint8), np.ones((8, 8), np.uint8))[:203, :311]
    smooth = cv2.GaussianBlur(noise, (9, 9), 3)
    synth = S.make_frames("C3", 1, width=311, height=203)[0]
    return {"noise": noise, "blocks": np.ascontiguousarray(blocks), "smooth": smooth, "synth": synth}


@pytest.mark.parametrize("scale", [1.2, 1.1, 1.5, 1.9, 1.2 ** 3])
def test_resize_linear(oracle_mod, scale):
    rng = np.random.default_rng(1)
    for (h, w) in [(480, 752), (97, 131), (376, 1241)]:
        src = rng.integers(0, 256, (h, w), dtype=np.uint8)
        dw, dh = int(round(w / scale)), int(round(h / scale))
        ref = cv2.resize(src, (dw, dh), interpolation=cv2.INTER_LINEAR)
        assert np.array_equal(oracle_mod.resize_linear(src, dw, dh), ref)


def test_resize_chain_is_the_pyramid(oracle_mod):
    img = S.make_frames("C1", 1)[0]
    e = oracle_mod.Extractor(1000)
    e.extract(img, (0, 0))
    prev = img
    for level in range(1, 8):
        w, h = e.level_size(level)
        prev = cv2.resize(prev, (w, h), interpolation=cv2.INTER_LINEAR)
        assert np.array_equal(e.pyramid(level), prev)
        b = cv2.copyMakeBorder(prev, 19, 19, 19, 19, cv2.BORDER_REFLECT_101 | cv2.BORDER_ISOLATED)
        assert np.array_equal(e.pyramid(level, True), b)


def test_gaussian_blur(oracle_mod):
    rng = np.random.default_rng(2)
    for name, img in _images(rng).items():
        ref = cv2.GaussianBlur(img.copy(), (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
        assert np.array_equal(oracle_mod.gaussian_blur7(img), ref), name
    cv2.setUseOptimized(False)
    try:
        img = rng.integers(0, 256, (134, 210), dtype=np.uint8)
        ref = cv2.GaussianBlur(img.copy(), (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
        assert np.array_equal(oracle_mod.gaussian_blur7(img), ref)
    finally:
        cv2.setUseOptimized(True)


def test_border(oracle_mod):
    rng = np.random.default_rng(3)
    img = rng.integers(0, 256, (57, 83), dtype=np.uint8)
    ref = cv2.copyMakeBorder(img, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
    assert np.array_equal(oracle_mod.border101(img, 19), ref)


def test_fast_atan2(oracle_mod):
    rng = np.random.default_rng(4)
    y = rng.integers(-2000000, 2000000, 200000).astype(np.float32)
    x = rng.integers(-2000000, 2000000, 200000).astype(np.float32)
    y[:5] = [0, 0, 1, -1, 5]
    x[:5] = [0, 1, 0, 0, 5]
    got = oracle_mod.fast_atan2(y, x)
    ref = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in zip(y[:20000], x[:20000])], np.float32)
    assert np.array_equal(got[:20000], ref)
    mag, ang = cv2.cartToPolar(x, y, angleInDegrees=True)  # vector path, 200k samples
    assert np.max(np.abs(got - ang.ravel())) < 1e-3


@pytest.mark.parametrize("th", [20, 7, 0, 50])
def test_fast_roi(oracle_mod, th):
    rng = np.random.default_rng(5)
    det = cv2.FastFeatureDetector_create(th, True)
    for name, img in _images(rng).items():
        for roi in [(0, 0, 311, 203), (16, 16, 58, 60), (100, 50, 142, 95), (270, 160, 311, 203), (5, 5, 12, 40)]:
            x0, y0, x1, y1 = roi
            kps = det.detect(img[y0:y1, x0:x1])
            ref = np.array([[int(k.pt[0]), int(k.pt[1]), int(k.response)] for k in kps], np.int32).reshape(-1, 3)
            got = oracle_mod.fast(img, roi, th)
            assert np.array_equal(got, ref), (name, roi, len(got), len(ref))


def test_knn2_matches_bfmatcher(oracle_mod):
    rng = np.random.default_rng(6)
    q = rng.integers(0, 256, (64, 32), dtype=np.uint8)
    r = rng.integers(0, 256, (200, 32), dtype=np.uint8)
    r[10] = r[20]  # exact tie
    q[3] = r[10]
    idx, dist = oracle_mod.knn2_dense(q, r)
    m = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(q, r, k=2)
    for i, pair in enumerate(m):
        assert [p.trainIdx for p in pair] == list(idx[i]) and [int(p.distance) for p in pair] == list(dist[i])


@pytest.mark.parametrize("size,nf,lap", [((376, 280), 400, (0, 0)), ((640, 480), 1000, (0, 1000)),
                                         ((500, 300), 150, (100, 300))])
def test_end_to_end_vs_cv2_pipeline(oracle_mod, size, nf, lap):
    from oracle.cv2_pipeline import Cv2Extractor
    w, h = size
    frames = S.make_frames("C1", 2, width=w, height=h)
    e = oracle_mod.Extractor(nf)
    c = Cv2Extractor(nf)
    for img in frames:
        k, d, m = e.extract(img, lap)
        k2, d2, m2 = c.extract(img, lap)
        assert m == m2 and len(k) == len(k2)
        assert np.array_equal(k, k2)
        assert np.array_equal(d, d2)


def test_end_to_end_strict_fp_flag(oracle_mod):
    from oracle.cv2_pipeline import Cv2Extractor
    img = S.make_frames("C3", 1, width=400, height=300)[0]
    e = oracle_mod.Extractor(500, flags=oracle_mod.STRICT_FP)
    c = Cv2Extractor(500, fma=False)
    k, d, m = e.extract(img, (0, 0))
    k2, d2, m2 = c.extract(img, (0, 0))
    assert np.array_equal(k, k2) and np.array_equal(d, d2)


def test_stereo_subpixel_vs_cv2_norm(oracle_mod):
    """Frame.cc:896-981 re-expressed with cv2.norm(NORM_L1) on Mat slices and numpy float32 scalars, from the
    oracle's Hamming-stage result: pins the SAD slide, parabola, clamp and median cut of the C++ oracle."""
    f32 = np.float32
    for t0, tr in ((0, None), (2, 34)):
        left = S.make_frames("C2", 1, t0=t0)[0]
        right = S.make_frames("C2", 1, t0=t0, right=True)[0] if tr is None else S.make_frames("C2", 1, t0=tr)[0]
        el, er = oracle_mod.Extractor(1200), oracle_mod.Extractor(1200)
        kl, dl, _ = el.extract(left, (0, 0))
        kr, dr, _ = er.extract(right, (0, 0))
        pl = [np.ascontiguousarray(el.pyramid(l)) for l in range(8)]
        pr = [np.ascontiguousarray(er.pyramid(l)) for l in range(8)]
        mbf = f32(47.90639384423901)
        mb = f32(mbf / f32(458.654))
        sf, isf = el.scale_factors, el.inv_scale_factors
        bi, bd, ur, dp, kept = oracle_mod.stereo_matches(kl, dl, kr, dr, sf, isf, pl, pr, mb, mbf)
        maxD = mbf / mb
        ur2 = np.full(len(kl), -1, f32)
        dp2 = np.full(len(kl), -1, f32)
        vdist = []
        rnd = lambda v: f32(np.floor(v + f32(0.5)) if v >= 0 else np.ceil(v - f32(0.5)))  # C round(): half away
        for iL in range(len(kl)):
            if bd[iL] >= 75 or bi[iL] < 0:
                continue
            o = int(kl["octave"][iL])
            uL = f32(kl["x"][iL])
            su, sv, sr = rnd(uL * isf[o]), rnd(f32(kl["y"][iL]) * isf[o]), rnd(f32(kr["x"][bi[iL]]) * isf[o])
            w = L = 5
            if sr + L - w < 0 or sr + L + w + 1 >= pl[o].shape[1]:
                continue
            su, sv, sr = int(su), int(sv), int(sr)
            IL = pl[o][sv - w:sv + w + 1, su - w:su + w + 1]
            best, binc, dists = 2 ** 31 - 1, 0, []
            for inc in range(-L, L + 1):
                IR = pr[o][sv - w:sv + w + 1, sr + inc - w:sr + inc + w + 1]
                d = f32(cv2.norm(IL, IR, cv2.NORM_L1))
                if d < best:
                    best, binc = int(d), inc
                dists.append(d)
            if binc in (-L, L):
                continue
            d1, d2, d3 = dists[L + binc - 1], dists[L + binc], dists[L + binc + 1]
            with np.errstate(all="ignore"):
                delta = (d1 - d3) / (f32(2.0) * (d1 + d3 - f32(2.0) * d2))
            if delta < -1 or delta > 1:
                continue
            bu = sf[o] * (f32(sr) + f32(binc) + delta)
            disp = uL - bu
            if disp >= 0 and disp < maxD:
                if disp <= 0:
                    disp = f32(0.01)
                    bu = f32(np.float64(uL) - 0.01)
                dp2[iL] = mbf / disp
                ur2[iL] = bu
                vdist.append((best, iL))
        vdist.sort()
        th = f32(1.5) * f32(1.4) * f32(vdist[len(vdist) // 2][0])
        for d, i in reversed(vdist):
            if d < th:
                break
            ur2[i] = dp2[i] = -1
        assert np.array_equal(ur.view(np.uint32), ur2.view(np.uint32))
        assert np.array_equal(dp.view(np.uint32), dp2.view(np.uint32))
        assert kept == int((ur2 > 0).sum()) > 100
