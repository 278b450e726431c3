"""The C++ drop-in classes (include/orbx_compat) built against liborbx.so and run on the GPU."""
import subprocess
from pathlib import Path

import numpy as np
import pytest

from orb_slam3_vio_fixes_b200 import synthetic as S
from orb_slam3_vio_fixes_b200 import build as B

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parents[1]


def _build_exe(tmp, name="compat_main"):
    B.build()
    exe = tmp / name
    pkg = ROOT / "orb_slam3_vio_fixes_b200"
    subprocess.check_call(["nvcc", "-O2", "-std=c++17", "-DORBX_NO_OPENCV", f"-I{ROOT / 'include'}",
                           str(ROOT / "tests" / "cpp" / f"{name}.cpp"), "-o", str(exe), f"-L{pkg}", "-lorbx",
                           "-Xlinker", f"-rpath={pkg}"])
    return exe


@pytest.mark.parametrize("lap", [(0, 0), (0, 1000)])
def test_cpp_dropin_matches_oracle(oracle_mod, tmp_path, lap):
    exe = _build_exe(tmp_path)
    img = S.make_frames("C1", 1)[0]
    raw = tmp_path / "in.raw"
    img.tofile(raw)
    out = subprocess.check_output([str(exe), str(raw), "752", "480", "1000", str(lap[0]), str(lap[1]),
                                   str(tmp_path / "o")], text=True)
    mono, n, dd, levels, nsf = [int(v) for v in out.split()]
    orc = oracle_mod.Extractor(1000)
    okps, odesc, omono = orc.extract(img, lap)
    kps = np.fromfile(tmp_path / "o.kps", dtype=oracle_mod.KP_DTYPE)
    desc = np.fromfile(tmp_path / "o.desc", dtype=np.uint8).reshape(-1, 32)
    assert (mono, n, levels, nsf) == (omono, len(okps), 8, 8)
    assert np.array_equal(kps, okps) and np.array_equal(desc, odesc)
    pyr3 = np.fromfile(tmp_path / "o.pyr3", dtype=np.uint8)
    assert np.array_equal(pyr3, orc.pyramid(3, True).ravel())
    assert dd == oracle_mod.descriptor_distance(odesc[0], odesc[1])
    res = np.fromfile(tmp_path / "o.match", dtype=oracle_mod.RES_DTYPE)
    off = (np.arange(n + 1) * 8).astype(np.int32)
    idx = ((np.arange(n)[:, None] + np.arange(1, 9)[None, :]) % n).astype(np.int32).ravel()
    ref = oracle_mod.argmin2_csr(odesc, odesc, off, idx, okps["octave"].astype(np.int8), 0)
    assert np.array_equal(res, ref)


def test_cpp_stereo_dropin_matches_oracle(oracle_mod, tmp_path):
    """Frame::ComputeStereoMatches replaced by ORB_SLAM3::ComputeStereoMatches(left, right, mb, mbf, ...)."""
    exe = _build_exe(tmp_path, "compat_stereo")
    left, right = S.make_frames("C2", 1)[0], S.make_frames("C2", 1, right=True)[0]
    left.tofile(tmp_path / "l.raw")
    right.tofile(tmp_path / "r.raw")
    mbf = np.float32(47.90639384423901)
    mb = np.float32(mbf / np.float32(458.654))
    out = subprocess.check_output([str(exe), str(tmp_path / "l.raw"), str(tmp_path / "r.raw"), "752", "480", "1200",
                                   repr(float(mb)), repr(float(mbf)), str(tmp_path / "o")], text=True)
    nl, nr, kept = [int(v) for v in out.split()]
    el, er = oracle_mod.Extractor(1200), oracle_mod.Extractor(1200)
    kl, dl, _ = el.extract(left, (0, 0))
    kr, dr, _ = er.extract(right, (0, 0))
    pl = [el.pyramid(l) for l in range(8)]
    pr = [er.pyramid(l) for l in range(8)]
    bi, bd, ur, dp, rkept = oracle_mod.stereo_matches(kl, dl, kr, dr, el.scale_factors, el.inv_scale_factors, pl, pr, mb, mbf)
    assert (nl, nr, kept) == (len(kl), len(kr), rkept) and kept > 100
    assert np.array_equal(np.fromfile(tmp_path / "o.ur", dtype=np.uint32), ur.view(np.uint32))
    assert np.array_equal(np.fromfile(tmp_path / "o.depth", dtype=np.uint32), dp.view(np.uint32))


def test_cpp_bow_helpers_match_oracle(oracle_mod, tmp_path):
    """include/orbx_compat/BowMatcher.h (SearchByBoW x2, SearchForTriangulation on std:: containers) vs the oracle."""
    exe = _build_exe(tmp_path, "compat_bow")
    frames = S.make_frames("C1", 2)
    orc = oracle_mod.Extractor(1000)
    (k1, d1, _), (k2, d2, _) = [orc.extract(f, (0, 0)) for f in frames]
    voc = S.make_vocabulary(10, 4, seed=5, desc_pool=np.concatenate([d1, d2]))
    fv1 = oracle_mod.bow_transform(d1, voc, 2)["fv"]
    fv2 = oracle_mod.bow_transform(d2, voc, 2)["fv"]
    rng = np.random.default_rng(9)
    v1 = (rng.random(len(k1)) < 0.7).astype(np.uint8)
    v2 = (rng.random(len(k2)) < 0.7).astype(np.uint8)
    u1 = np.where(rng.random(len(k1)) < 0.5, 100.0, -1.0).astype(np.float32)
    u2 = np.where(rng.random(len(k2)) < 0.5, 100.0, -1.0).astype(np.float32)
    F = np.array([0, -0.01, 0.447, 0.01, 0, -0.894, -0.447, 0.894, 0], np.float32)
    ep = np.array([500.0, 300.0], np.float32)
    sf = np.array([1.2 ** i for i in range(8)], np.float32)
    sg = (sf * sf).astype(np.float32)
    with open(tmp_path / "in.bin", "wb") as f:
        f.write(np.array([len(k1), len(k2), len(fv1[0]), len(fv2[0])], np.int32).tobytes())
        for k, d, v, u, fv in ((k1, d1, v1, u1, fv1), (k2, d2, v2, u2, fv2)):
            idx = np.zeros(len(k), np.int32)
            idx[:len(fv[2])] = fv[2]
            for a in (k, d, v, u, fv[0], fv[1], idx):
                f.write(np.ascontiguousarray(a).tobytes())
        for a in (F, ep, sf, sg):
            f.write(a.tobytes())
    out = subprocess.check_output([str(exe), str(tmp_path / "in.bin"), str(tmp_path / "o")], text=True)
    n0, n1, n2, npairs = [int(v) for v in out.split()]
    m12, m21, r0 = oracle_mod.search_by_bow(0, k1, d1, fv1, k2, d2, fv2, 0.7, True, v1, v2)
    assert n0 == r0 > 50
    assert np.array_equal(np.fromfile(tmp_path / "o.bow0_m12", np.int32), m12)
    assert np.array_equal(np.fromfile(tmp_path / "o.bow0_m21", np.int32), m21)
    m12, _, r1 = oracle_mod.search_by_bow(1, k1, d1, fv1, k2, d2, fv2, 0.8, True, v1, v2)
    assert n1 == r1 > 50
    assert np.array_equal(np.fromfile(tmp_path / "o.bow1_m12", np.int32), m12)
    t12, r2 = oracle_mod.search_for_triangulation(k1, d1, fv1, k2, d2, fv2, F, ep, sf, sg, False, False, True,
                                                  1 - v1, 1 - v2, u1, u2)
    assert n2 == r2 == npairs
    pairs = np.fromfile(tmp_path / "o.tri_pairs", np.int32).reshape(-1, 2)
    ref = np.stack([np.nonzero(t12 >= 0)[0], t12[t12 >= 0]], 1)
    assert np.array_equal(pairs, ref)


def test_cpp_orbmatcher_class(oracle_mod, tmp_path):
    """include/orbx_compat/ORBmatcher.h: ORBmatcherT<Frame, KeyFrame, MapPoint> with the reference's method names and
    arguments (include/ORBmatcher.h:40-87) on stand-in object types (tests/cpp/compat_matcher.cpp), against the oracle:
    SearchByProjection (MapPoints: single camera and fisheye rig; last frame), SearchByBoW (KF-Frame, KF-rig Frame,
    KF-KF), SearchForTriangulation, SearchForInitialization, the relocalisation and Sim3 projections, Fuse x2 (with the
    reference's Replace / AddObservation rules) and SearchBySim3 — every method of include/ORBmatcher.h:40-87."""
    exe = _build_exe(tmp_path, "compat_matcher")
    rng = np.random.default_rng(101)
    f32 = np.float32
    sf = np.cumprod([1.0] + [1.2] * 7).astype(f32)   # mvScaleFactors as ORBextractor builds them is close enough here:
    orc = oracle_mod.Extractor(1000)
    sf = np.asarray(orc.scale_factors, f32)
    sig = (sf * sf).astype(f32)
    blob = []
    w = lambda a, dt: blob.append(np.ascontiguousarray(a, dt).tobytes())

    def w_points(P):
        M = len(P["inView"])
        w([M], np.int32)
        w(np.stack([P["projX"], P["projY"], P["projXR"], P["projYR"], P["viewCos"], P["viewCosR"], P["depth"],
                    P["pos"][:, 0], P["pos"][:, 1], P["pos"][:, 2]], 1), f32)
        w(np.stack([P["inView"], P["inViewR"], P["level"], P["levelR"], P["bad"], P["nObs"]], 1), np.int32)
        w(P["desc"], np.uint8)

    def w_feats(k, d):
        w([len(k)], np.int32); blob.append(np.ascontiguousarray(k).tobytes()); w(d, np.uint8)

    def w_fv(fv, n):
        idx = np.zeros(n, np.int32); idx[:len(fv[2])] = fv[2]
        w([len(fv[0])], np.int32); w(fv[0], np.int32); w(fv[1], np.int32); w(idx, np.int32)

    def points_from(kq, dq, shift, right_dx=None):
        M = len(kq)
        P = dict(projX=(kq["x"] + f32(shift[0])).astype(f32), projY=(kq["y"] + f32(shift[1])).astype(f32),
                 viewCos=np.where(rng.random(M) < 0.5, f32(0.9995), f32(0.9)).astype(f32),
                 viewCosR=np.where(rng.random(M) < 0.5, f32(0.9995), f32(0.9)).astype(f32),
                 depth=(rng.random(M) * 50).astype(f32), level=kq["octave"].astype(np.int32),
                 inView=(rng.random(M) < 0.9).astype(np.int32), bad=(rng.random(M) < 0.05).astype(np.int32),
                 nObs=rng.integers(0, 3, M).astype(np.int32), desc=dq, pos=np.zeros((M, 3), f32))
        if right_dx is None:
            P["projXR"] = (P["projX"] - f32(24.0) + rng.normal(0, 3, M)).astype(f32)
            P["projYR"] = P["projY"].copy()
            P["inViewR"] = np.zeros(M, np.int32)
            P["levelR"] = np.full(M, -1, np.int32)
        else:
            P["projXR"] = (P["projX"] + f32(right_dx) + rng.normal(0, 1.5, M)).astype(f32)
            P["projYR"] = (P["projY"] + rng.normal(0, 1.0, M)).astype(f32)
            P["inViewR"] = (rng.random(M) < 0.85).astype(np.int32)
            P["levelR"] = np.where(rng.random(M) < 0.9, kq["octave"], -1).astype(np.int32)
        return P

    def holders(n, M, frac):
        return np.where(rng.random(n) < frac, rng.integers(0, M, n), -1).astype(np.int32)

    radius = lambda vc, th: np.where(vc.astype(np.float64) > 0.998, f32(2.5), f32(4.0)).astype(f32) * f32(th)

    w(sf, f32); w(sig, f32)
    bounds = (0.0, 752.0, 0.0, 480.0)
    w(bounds, f32)
    frames = S.make_frames("C1", 4)
    feats = [oracle_mod.Extractor(1000).extract(f, (0, 0))[:2] for f in frames]
    pos = np.array([S.frame_shift(t) for t in range(4)], f32)
    exp = {}

    # 1. SearchByProjection(F, MapPoints, th=3, bFarPoints, thFar=40), single camera
    kF, dF = feats[0]
    kq = np.concatenate([feats[s][0] for s in (1, 2, 3)]); dq = np.concatenate([feats[s][1] for s in (1, 2, 3)])
    shifts = np.concatenate([np.tile(pos[s] - pos[0], (len(feats[s][0]), 1)) for s in (1, 2, 3)])
    P = points_from(kq, dq, (0, 0)); P["projX"] = (kq["x"] + shifts[:, 0]).astype(f32); P["projY"] = (kq["y"] + shifts[:, 1]).astype(f32)
    uR = np.where(rng.random(len(kF)) < 0.7, kF["x"] - f32(24.0), f32(-1)).astype(f32)
    hold = holders(len(kF), len(kq), 0.15)
    w_points(P); w_feats(kF, dF); w(uR, f32); w(hold, np.int32)
    part = (P["inView"] > 0) & ~(P["depth"] > f32(40.0)) & (P["bad"] == 0)
    q = dict(qx=P["projX"], qy=P["projY"], qr=(radius(P["viewCos"], 3.0) * sf[P["level"]]).astype(f32),
             qmin=(P["level"] - 1).astype(np.int8), qmax=P["level"].astype(np.int8))
    flags = (part.astype(np.uint8) | ((P["nObs"] > 0).astype(np.uint8) << 1)).astype(np.uint8)
    locked = ((hold >= 0) & (P["nObs"][np.maximum(hold, 0)] > 0)).astype(np.uint8)
    _, fq, nm = oracle_mod.search_by_projection(kF, dF, bounds, q["qx"], q["qy"], q["qr"], q["qmin"], q["qmax"], P["desc"], 0,
                                                0.8, False, u_right=uR, feat_locked=locked, q_ur=P["projXR"], q_flags=flags)
    exp["proj_mono"] = (np.where(fq >= 0, fq, hold), nm)

    # 2. the same on a rig frame (left / right eye of C2)
    el, er = oracle_mod.Extractor(1200), oracle_mod.Extractor(1200)
    kl, dl, _ = el.extract(S.make_frames("C2", 1)[0], (0, 0))
    kr, dr, _ = er.extract(S.make_frames("C2", 1, right=True)[0], (0, 0))
    k2, d2, _ = oracle_mod.Extractor(1200).extract(S.make_frames("C2", 2)[1], (0, 0))
    sh = np.array(S.frame_shift(1), f32) - np.array(S.frame_shift(0), f32)
    P = points_from(k2, d2, sh, right_dx=-24.0)
    nl, nr = len(kl), len(kr)
    hold = holders(nl + nr, len(k2), 0.15)
    m = int(min(nl, nr) * 0.4)
    li, ri = rng.permutation(nl)[:m], rng.permutation(nr)[:m]
    l2r, r2l = np.full(nl, -1, np.int32), np.full(nr, -1, np.int32)
    l2r[li] = ri; r2l[ri] = li
    w_points(P); w_feats(kl, dl); w_feats(kr, dr); w(hold, np.int32); w(l2r, np.int32); w(r2l, np.int32)
    part = ((P["inView"] > 0) | (P["inViewR"] > 0)) & (P["bad"] == 0)
    flags = ((part & (P["inView"] > 0)).astype(np.uint8) | ((P["nObs"] > 0).astype(np.uint8) << 1) |
             ((part & (P["inViewR"] > 0) & (P["levelR"] != -1)).astype(np.uint8) << 2)).astype(np.uint8)
    lvr = np.maximum(P["levelR"], 0)
    q = dict(qx=P["projX"], qy=P["projY"], qr=(radius(P["viewCos"], 3.0) * sf[P["level"]]).astype(f32),
             qmin=(P["level"] - 1).astype(np.int8), qmax=P["level"].astype(np.int8),
             qx_r=P["projXR"], qy_r=P["projYR"], qr_r=(radius(P["viewCosR"], 1.0) * sf[lvr]).astype(f32),
             qmin_r=(lvr - 1).astype(np.int8), qmax_r=lvr.astype(np.int8), qdesc=P["desc"], flags=flags)
    locked = ((hold >= 0) & (P["nObs"][np.maximum(hold, 0)] > 0)).astype(np.uint8)
    o = oracle_mod.search_by_projection_rig(0, kl, dl, kr, dr, bounds, q, 0.8, False, locked_l=locked[:nl],
                                            locked_r=locked[nl:], l2r=l2r, r2l=r2l)
    fq = np.concatenate([o["feat_query_l"], o["feat_query_r"]])
    exp["proj_rig"] = (np.where(fq >= 0, fq, hold), o["nmatches"])

    # 3. SearchByProjection(CurrentFrame, LastFrame, 15, bMono=true): pinhole, identity rotation, z = 1
    (kL, dL), (kC, dC) = feats[0], feats[1]
    fx, fy, cx, cy = f32(458.654), f32(457.296), f32(367.215), f32(248.375)
    M = len(kL)
    shv = pos[0] - pos[1]
    Pm = points_from(kL, dL, (0, 0))
    Pm["pos"] = np.stack([(kL["x"] - cx) / fx, (kL["y"] - cy) / fy, np.ones(M, f32)], 1).astype(f32)
    t = np.array([shv[0] / fx, shv[1] / fy, 0], f32)
    X = (Pm["pos"][:, 0] + t[0]).astype(f32); Y = (Pm["pos"][:, 1] + t[1]).astype(f32)
    u = ((fx * X).astype(f32) + cx).astype(f32); v = ((fy * Y).astype(f32) + cy).astype(f32)
    holdL = np.where(rng.random(M) < 0.8, np.arange(M), -1).astype(np.int32)
    outl = (rng.random(M) < 0.05).astype(np.int32)
    holdC = holders(len(kC), M, 0.1)
    uRC = np.where(rng.random(len(kC)) < 0.5, kC["x"] - f32(24.0), f32(-1)).astype(f32)
    mbf = f32(47.9)
    w_points(Pm); w_feats(kC, dC); w(uRC, f32); w(holdC, np.int32); w_feats(kL, dL); w(holdL, np.int32); w(outl, np.int32)
    w([fx, fy, cx, cy], f32); w(list(np.eye(3).ravel()) + list(t), f32); w(list(np.eye(3).ravel()) + [0, 0, 0], f32)
    w([0.1, mbf], f32)
    sel = np.nonzero((holdL >= 0) & (outl == 0) & (u >= f32(bounds[0])) & (u <= f32(bounds[1])) & (v >= f32(bounds[2])) &
                     (v <= f32(bounds[3])))[0]
    q = dict(qx=u[sel], qy=v[sel], qr=(f32(15.0) * sf[kL["octave"][sel]]).astype(f32),
             qmin=(kL["octave"][sel] - 1).astype(np.int8), qmax=(kL["octave"][sel] + 1).astype(np.int8))
    q_ur = (u[sel] - (mbf * f32(1.0))).astype(f32)
    flags = (1 | ((Pm["nObs"][sel] > 0).astype(np.uint8) << 1)).astype(np.uint8)
    lockedC = ((holdC >= 0) & (Pm["nObs"][np.maximum(holdC, 0)] > 0)).astype(np.uint8)
    args = (kC, dC, bounds, q["qx"], q["qy"], q["qr"], q["qmin"], q["qmax"], dL[sel], 1, 0.9)
    kw = dict(u_right=uRC, feat_locked=lockedC, q_ur=q_ur, q_flags=flags, q_angle=kL["angle"][sel])
    _, fq0, _ = oracle_mod.search_by_projection(*args, False, **kw)
    _, fq1, nm = oracle_mod.search_by_projection(*args, True, **kw)
    ids = np.where(fq0 >= 0, np.where(fq1 >= 0, sel[np.maximum(fq1, 0)], -1), holdC)
    exp["proj_last"] = (ids, nm)

    # 4.-6. SearchByBoW x3, SearchForTriangulation
    (k1, d1), (k2b, d2b) = feats[0], feats[1]
    voc = S.make_vocabulary(10, 4, seed=5, desc_pool=np.concatenate([d1, d2b]))
    fv1 = oracle_mod.bow_transform(d1, voc, 2)["fv"]; fv2 = oracle_mod.bow_transform(d2b, voc, 2)["fv"]
    Mb = 3000
    Pb = points_from(np.concatenate([k1, k2b, k1])[:Mb], np.concatenate([d1, d2b, d1])[:Mb], (0, 0))
    h1, h2 = holders(len(k1), Mb, 0.7), holders(len(k2b), Mb, 0.7)
    u1 = np.where(rng.random(len(k1)) < 0.5, 100.0, -1.0).astype(f32)
    u2 = np.where(rng.random(len(k2b)) < 0.5, 100.0, -1.0).astype(f32)
    Fm = np.array([0, -0.01, 0.447, 0.01, 0, -0.894, -0.447, 0.894, 0], f32); ep = np.array([500.0, 300.0], f32)
    w_points(Pb)
    for k_, d_, fv_, h_, u_ in ((k1, d1, fv1, h1, u1), (k2b, d2b, fv2, h2, u2)):
        w_feats(k_, d_); w_fv(fv_, len(k_)); w(h_, np.int32); w(u_, f32)
    w(np.concatenate([Fm, ep]), f32)
    v1 = ((h1 >= 0) & (Pb["bad"][np.maximum(h1, 0)] == 0)).astype(np.uint8)
    v2 = ((h2 >= 0) & (Pb["bad"][np.maximum(h2, 0)] == 0)).astype(np.uint8)
    _, m21, nm = oracle_mod.search_by_bow(0, k1, d1, fv1, k2b, d2b, fv2, 0.7, True, v1, None)
    exp["bow_kf_f"] = (np.where(m21 >= 0, h1[np.maximum(m21, 0)], -1), nm)
    m12, _, nm = oracle_mod.search_by_bow(1, k1, d1, fv1, k2b, d2b, fv2, 0.7, True, v1, v2)
    exp["bow_kf_kf"] = (np.where(m12 >= 0, h2[np.maximum(m12, 0)], -1), nm)
    t12, nm = oracle_mod.search_for_triangulation(k1, d1, fv1, k2b, d2b, fv2, Fm, ep, sf, sig, False, False, True,
                                                  (h1 >= 0).astype(np.uint8), (h2 >= 0).astype(np.uint8), u1, u2)
    exp["tri_pairs"] = (np.stack([np.nonzero(t12 >= 0)[0], t12[t12 >= 0]], 1).ravel(), nm)
    dcat = np.concatenate([dl, dr]); kcat = np.concatenate([kl, kr])
    voc2 = S.make_vocabulary(10, 4, seed=6, desc_pool=np.concatenate([d1, dcat]))
    fvk = oracle_mod.bow_transform(d1, voc2, 2)["fv"]
    # (the keyframe's FeatureVector was written for `voc`; the rig frame must share the keyframe's tree, so the rig
    # frame is transformed with `voc` as well)
    fvr = oracle_mod.bow_transform(dcat, voc, 2)["fv"]
    del fvk, voc2
    w_feats(kl, dl); w_feats(kr, dr); w_fv(fvr, len(kcat))
    _, _, m21, nm = oracle_mod.search_by_bow_rig(k1, d1, fv1, kcat, dcat, nl, fvr, 0.7, True, valid1=v1)
    exp["bow_kf_rig"] = (np.where(m21 >= 0, h1[np.maximum(m21, 0)], -1), nm)

    # 7. SearchForInitialization
    w_feats(k1, d1); w_feats(k2b, d2b)
    prev = np.stack([k1["x"], k1["y"]], 1)
    m12, nm, _ = oracle_mod.search_for_initialization(k1, d1, k2b, d2b, bounds, prev, 100, 0.9, True)
    exp["init_m12"] = (m12, nm)

    # 8.-11. relocalisation, Sim3 projections x2, Fuse, Fuse(Sim3): keyframe / current frame = frame 1, query points =
    # frame 0's keypoints at world (u, v, 1) (identity poses, unit pinhole); PredictScale returns the keypoint's octave
    (kQ, dQ), (kT, dT) = feats[0], feats[1]
    shv = pos[0] - pos[1]
    Mq, nT = len(kQ), len(kT)
    uvq = np.stack([(kQ["x"] + shv[0]).astype(f32), (kQ["y"] + shv[1]).astype(f32)], 1)
    distq = np.sqrt(uvq[:, 0].astype(np.float64) ** 2 + uvq[:, 1].astype(np.float64) ** 2 + 1.0)
    lvq = kQ["octave"].astype(np.int64)
    mxq, mnq = (distq * 1.2 ** (lvq - 0.5)).astype(f32), (distq * 0.1).astype(f32)
    qf = ((rng.random(Mq) < 0.9).astype(np.int32) | ((rng.random(Mq) < 0.05).astype(np.int32) << 1) |
          ((rng.random(Mq) < 0.1).astype(np.int32) << 2)).astype(np.int32)
    hfT = ((rng.random(nT) < 0.5).astype(np.int32) | ((rng.random(nT) < 0.1).astype(np.int32) << 1)).astype(np.int32)
    uRT = np.where(rng.random(nT) < 0.5, kT["x"] - f32(40.0), f32(-1)).astype(f32)
    w([Mq], np.int32); w(qf, np.int32); w(uvq, f32); w(mxq, f32); w(mnq, f32); w(kQ["angle"], f32); w(dQ, np.uint8)
    w_feats(kT, dT); w(hfT, np.int32); w(uRT, f32)
    held = (hfT & 1) > 0
    own_ids = np.where(held, -2 - np.arange(nT), -1)
    inb = (uvq[:, 0] >= 0) & (uvq[:, 0] <= 752) & (uvq[:, 1] >= 0) & (uvq[:, 1] <= 480)
    inimg = (uvq[:, 0] >= 0) & (uvq[:, 0] < 752) & (uvq[:, 1] >= 0) & (uvq[:, 1] < 480)
    bad_q, found_q, has_q = (qf & 2) > 0, (qf & 4) > 0, (qf & 1) > 0
    # 8. relocalisation (th = 10, ORBdist = 100, rotation check)
    sel = np.nonzero(has_q & ~bad_q & ~found_q & inb)[0]
    args = (kT, dT, bounds, uvq[sel, 0], uvq[sel, 1], (f32(10.0) * sf[lvq[sel]]).astype(f32), lvq[sel] - 1, lvq[sel] + 1,
            dQ[sel], 1, 0.9)
    kw = dict(feat_locked=held.astype(np.uint8), q_flags=np.full(len(sel), 3, np.uint8), q_angle=kQ["angle"][sel], th_dist=100)
    _, fq0, _ = oracle_mod.search_by_projection(*args, False, **kw)
    _, fq1, nm = oracle_mod.search_by_projection(*args, True, **kw)
    exp["reloc"] = (np.where(fq0 >= 0, np.where(fq1 >= 0, sel[np.maximum(fq1, 0)], -1), own_ids), nm)
    # 9. Sim3 projections: vpMatched = the keyframe's own points, "already found" query i sits at feature i
    minit = own_ids.copy()
    placed = np.nonzero(found_q & (np.arange(Mq) < nT))[0]
    minit[placed] = placed
    already = np.zeros(Mq, bool); already[placed] = True
    sel = np.nonzero(~bad_q & ~already & inimg)[0]
    _, fq, nm = oracle_mod.search_by_projection(kT, dT, bounds, uvq[sel, 0], uvq[sel, 1], (f32(8.0) * sf[lvq[sel]]).astype(f32),
                                                lvq[sel] - 1, lvq[sel], dQ[sel], 1, 0.75, False,
                                                feat_locked=(minit != -1).astype(np.uint8),
                                                q_flags=np.full(len(sel), 3, np.uint8), th_dist=50)
    exp["sim3_a"] = exp["sim3_b"] = (np.where(fq >= 0, sel[np.maximum(fq, 0)], minit), nm)
    # 10. Fuse(pKF, vpMapPoints, 3.0): candidate scoring from the oracle, then the reference's Replace / Add rules
    inv_sig = (f32(1.0) / sig).astype(f32)
    sel = np.nonzero(has_q & ~bad_q & inimg)[0]
    bi, bd = oracle_mod.fuse_candidates(kT, dT, bounds, inv_sig, uvq[sel, 0], uvq[sel, 1], (f32(3.0) * sf[lvq[sel]]).astype(f32),
                                        lvq[sel], dQ[sel], u_right=uRT, q_ur=(uvq[sel, 0] - f32(40.0)).astype(f32))
    state = {j: ("own", 1, bool(hfT[j] & 2)) for j in range(nT) if held[j]}   # feature -> (who, nObs, bad)
    where = np.full(Mq, -1, np.int32)
    nf = 0
    for k_, i in enumerate(sel):
        if bi[k_] < 0 or bd[k_] > 50:
            continue
        j = int(bi[k_]); nf += 1
        if j in state:
            who, nobs, isbad = state[j]
            if not isbad:
                if nobs > 1:                     # pMPinKF->Observations() > pMP->Observations(): the new point gives way
                    where[i] = -3
                else:                            # the held point is replaced by the new one, which takes its observation
                    if who != "own":
                        where[who] = -1          # (cannot happen: a query that holds a feature has 2 observations)
                    state[j] = (int(i), 2, False); where[i] = j
        else:
            state[j] = (int(i), 2, False); where[i] = j
    exp["fuse"] = (where, nf)
    # 11. Fuse(pKF, Scw, vpPoints, 4.0, vpReplacePoint)
    sel = np.nonzero(~bad_q & inimg)[0]
    bi, bd = oracle_mod.fuse_candidates(kT, dT, bounds, np.zeros(8, f32), uvq[sel, 0], uvq[sel, 1],
                                        (f32(4.0) * sf[lvq[sel]]).astype(f32), lvq[sel], dQ[sel])
    holder = {j: (-2 - j, bool(hfT[j] & 2)) for j in range(nT) if held[j]}    # feature -> (id, bad)
    where = np.full(Mq, -1, np.int32)
    nf = 0
    for k_, i in enumerate(sel):
        if bi[k_] < 0 or bd[k_] > 50:
            continue
        j = int(bi[k_]); nf += 1
        if j in holder:
            hid, hbad = holder[j]
            if not hbad:
                where[i] = -2 - hid              # vpReplacePoint[i] = the holder (own point: its feature; query k: -2-k)
        else:
            holder[j] = (int(i), False); where[i] = j
    exp["fuse_sim3"] = (where, nf)
    # 12. SearchBySim3: both keyframes' points, scored both ways, agreement
    (kA, dA), (kB, dB) = feats[0], feats[1]
    sAB = pos[0] - pos[1]
    uvA = np.stack([(kA["x"] + sAB[0]).astype(f32), (kA["y"] + sAB[1]).astype(f32)], 1)
    uvB = np.stack([(kB["x"] - sAB[0]).astype(f32), (kB["y"] - sAB[1]).astype(f32)], 1)

    def rng_of(uv, k):
        dd = np.sqrt(uv[:, 0].astype(np.float64) ** 2 + uv[:, 1].astype(np.float64) ** 2 + 1.0)
        return (dd * 1.2 ** (k["octave"].astype(np.int64) - 0.5)).astype(f32), (dd * 0.1).astype(f32)

    mxA, mnA = rng_of(uvA, kA); mxB, mnB = rng_of(uvB, kB)
    hfA = ((rng.random(len(kA)) < 0.85).astype(np.int32) | ((rng.random(len(kA)) < 0.05).astype(np.int32) << 1)).astype(np.int32)
    hfB = ((rng.random(len(kB)) < 0.85).astype(np.int32) | ((rng.random(len(kB)) < 0.05).astype(np.int32) << 1)).astype(np.int32)
    init12 = np.full(len(kA), -1, np.int32)
    pre = rng.permutation(len(kA))[:60]
    init12[pre] = rng.permutation(np.nonzero(hfB & 1)[0])[:60]
    for k_, d_, hf_, uv_, mx_, mn_ in ((kA, dA, hfA, uvA, mxA, mnA), (kB, dB, hfB, uvB, mxB, mnB)):
        w_feats(k_, d_); w(hf_, np.int32); w(uv_, f32); w(mx_, f32); w(mn_, f32)
    w(init12, np.int32)

    def direction(kq, dq, hfq, uv, skip, kt, dt):
        ii = (uv[:, 0] >= 0) & (uv[:, 0] < 752) & (uv[:, 1] >= 0) & (uv[:, 1] < 480)
        s_ = np.nonzero(((hfq & 1) > 0) & ((hfq & 2) == 0) & ~skip & ii)[0]
        lv = kq["octave"][s_].astype(np.int64)
        b_i, b_d = oracle_mod.fuse_candidates(kt, dt, bounds, np.zeros(8, f32), uv[s_, 0], uv[s_, 1],
                                              (f32(7.5) * sf[lv]).astype(f32), lv, dq[s_])
        o_ = np.full(len(kq), -1, np.int64)
        ok_ = (b_i >= 0) & (b_d <= 100)
        o_[s_[ok_]] = b_i[ok_]
        return o_

    skipA = init12 >= 0
    skipB = np.zeros(len(kB), bool); skipB[init12[init12 >= 0]] = True
    m1 = direction(kA, dA, hfA, uvA, skipA, kB, dB)
    m2 = direction(kB, dB, hfB, uvB, skipB, kA, dA)
    want = init12.astype(np.int64).copy()
    agree = (m1 >= 0) & (m2[np.maximum(m1, 0)] == np.arange(len(kA)))
    want[agree] = m1[agree]
    exp["sim3_search"] = (want.astype(np.int32), int(agree.sum()))

    (tmp_path / "in.bin").write_bytes(b"".join(blob))
    out = subprocess.check_output([str(exe), str(tmp_path / "in.bin"), str(tmp_path / "o")], text=True)
    rets = [int(v) for v in out.split()]
    order = ["proj_mono", "proj_rig", "proj_last", "bow_kf_f", "bow_kf_kf", "tri_pairs", "bow_kf_rig", "init_m12", "reloc",
             "sim3_a", "sim3_b", "fuse", "fuse_sim3", "sim3_search"]
    assert len(rets) == len(order)
    for name, ret in zip(order, rets):
        want, nm = exp[name]
        got = np.fromfile(tmp_path / f"o.{name}", np.int32)
        assert ret == nm, (name, ret, nm)
        assert np.array_equal(got, want), (name, np.nonzero(got != want)[0][:10])
        assert nm > 50, (name, nm)
