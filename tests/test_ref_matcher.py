"""oracle/_ref/libref_orbmatcher.so: the REFERENCE's own src/ORBmatcher.cc, compiled unmodified in front of the
stand-in prelude (oracle/ref_shim_matcher, recipe `make -C oracle ref_matcher`), against the oracle's restatements.

This is what pins the matcher rows to the reference: every candidate loop, gate, tie rule,
`continue`, rotation histogram and pointer write runs as the reference wrote and compiled it; the oracle (and through
it the CUDA path, tests/test_gpu_*.py) must return the same matches on the same inputs.  Geometry the matcher computes
inside (Tcw * x, project) is driven with poses that make it exact (identity rotation, unit pinhole, z = 1).

Needs either /root/reference (build container) or the prebuilt oracle/_ref/*.so (GPU box); else skipped."""
import numpy as np
import pytest

from orb_slam3_vio_fixes_b200 import synthetic as S

f32 = np.float32
BOUNDS = (0.0, 752.0, 0.0, 480.0)


@pytest.fixture(scope="module")
def RM(oracle_mod):
    from oracle import ref_matcher as R
    if not R.available():
        pytest.skip("neither /root/reference nor a prebuilt oracle/_ref")
    R.build()
    return R


@pytest.fixture(scope="module")
def data(oracle_mod):
    ex = oracle_mod.Extractor(1000)
    sf = np.asarray(ex.scale_factors, f32)
    mono = [ex.extract(f, (0, 0))[:2] for f in S.make_frames("C1", 4)]
    el = oracle_mod.Extractor(1200)
    left = [el.extract(f, (0, 0))[:2] for f in S.make_frames("C2", 2)]
    right = [el.extract(f, (0, 0))[:2] for f in S.make_frames("C2", 2, right=True)]
    pos = np.array([S.frame_shift(t) for t in range(4)], f32)
    return dict(sf=sf, sig=(sf * sf).astype(f32), mono=mono, left=left, right=right, pos=pos)


def radius(view_cos, th):
    r = np.where(view_cos.astype(np.float64) > 0.998, f32(2.5), f32(4.0)).astype(f32)
    return r if th == 1.0 else (r * f32(th)).astype(f32)


def points(rng, kq, dq, shift, right_dx=None):
    M = len(kq)
    P = dict(view_cos=np.where(rng.random(M) < 0.5, f32(0.9995), f32(0.9)).astype(f32),
             view_cos_r=np.where(rng.random(M) < 0.5, f32(0.9995), f32(0.9)).astype(f32),
             depth=(rng.random(M) * 50).astype(f32), level=kq["octave"].astype(np.int32), desc=dq)
    x, y = (kq["x"] + f32(shift[0])).astype(f32), (kq["y"] + f32(shift[1])).astype(f32)
    in_view, bad, obs = rng.random(M) < 0.9, rng.random(M) < 0.05, rng.random(M) < 0.7
    if right_dx is None:
        xr, yr = (x - f32(24.0) + rng.normal(0, 3, M)).astype(f32), y.copy()
        in_view_r = np.zeros(M, bool)
        P["level_r"] = np.full(M, -1, np.int32)
    else:
        xr, yr = (x + f32(right_dx) + rng.normal(0, 1.5, M)).astype(f32), (y + rng.normal(0, 1.0, M)).astype(f32)
        in_view_r = rng.random(M) < 0.85
        P["level_r"] = np.where(rng.random(M) < 0.9, kq["octave"], -1).astype(np.int32)
    P["proj"] = np.stack([x, y, xr, yr], 1).astype(f32)
    P["flags"] = (in_view.astype(np.uint8) | (obs.astype(np.uint8) << 1) | (in_view_r.astype(np.uint8) << 2) |
                  (bad.astype(np.uint8) << 3)).astype(np.uint8)
    return P


@pytest.mark.parametrize("th,far", [(3.0, True), (1.0, False)])
def test_search_by_projection_mappoints_mono(oracle_mod, RM, data, th, far):
    """ORBmatcher.cc:43-142: radius by viewing cosine (x th unless th == 1), bFarPoints, isBad, the mvuRight gate, the
    ratio test with levels, features locked on entry and by earlier MapPoints."""
    rng = np.random.default_rng(11)
    sf, pos = data["sf"], data["pos"]
    kF, dF = data["mono"][0]
    kq = np.concatenate([data["mono"][s][0] for s in (1, 2, 3)])
    dq = np.concatenate([data["mono"][s][1] for s in (1, 2, 3)])
    P = points(rng, kq, dq, (0, 0))
    sh = np.concatenate([np.tile(pos[s] - pos[0], (len(data["mono"][s][0]), 1)) for s in (1, 2, 3)])
    P["proj"][:, 0] = kq["x"] + sh[:, 0]; P["proj"][:, 1] = kq["y"] + sh[:, 1]
    uR = np.where(rng.random(len(kF)) < 0.7, kF["x"] - f32(24.0), f32(-1)).astype(f32)
    locked = (rng.random(len(kF)) < 0.15).astype(np.uint8)
    holder, nm = RM.search_by_projection_mappoints(kF, dF, BOUNDS, sf, P, 0.8, th, u_right=uR, locked=locked,
                                                   far_points=far, th_far=40.0)
    fl = P["flags"]
    part = ((fl & 1) > 0) & ~(far & (P["depth"] > f32(40.0))) & ((fl & 8) == 0)
    q_flags = (part.astype(np.uint8) | (fl & 2)).astype(np.uint8)
    qr = (radius(P["view_cos"], th) * sf[P["level"]]).astype(f32)
    _, fq, onm = oracle_mod.search_by_projection(kF, dF, BOUNDS, P["proj"][:, 0], P["proj"][:, 1], qr, P["level"] - 1,
                                                 P["level"], P["desc"], 0, 0.8, False, u_right=uR, feat_locked=locked,
                                                 q_ur=P["proj"][:, 2], q_flags=q_flags)
    assert nm == onm > 300
    assert np.array_equal(holder, np.where(fq >= 0, fq, np.where(locked > 0, -2, -1)))


def test_search_by_projection_mappoints_rig(oracle_mod, RM, data):
    """ORBmatcher.cc:43-213 with F.Nleft != -1: right block, pair propagation, the skipped right search."""
    rng = np.random.default_rng(12)
    sf = data["sf"]
    (kl, dl), (kr, dr) = data["left"][0], data["right"][0]
    k2, d2 = data["left"][1]
    sh = np.array(S.frame_shift(1), f32) - np.array(S.frame_shift(0), f32)
    P = points(rng, k2, d2, sh, right_dx=-24.0)
    nl, nr = len(kl), len(kr)
    locked = (rng.random(nl + nr) < 0.15).astype(np.uint8)
    m = int(min(nl, nr) * 0.4)
    li, ri = rng.permutation(nl)[:m], rng.permutation(nr)[:m]
    l2r, r2l = np.full(nl, -1, np.int32), np.full(nr, -1, np.int32)
    l2r[li] = ri; r2l[ri] = li
    for th in (3.0, 1.0):
        holder, nm = RM.search_by_projection_mappoints(kl, dl, BOUNDS, sf, P, 0.8, th, kps_r=kr, desc_r=dr, locked=locked,
                                                       l2r=l2r, r2l=r2l)
        fl = P["flags"]
        part = (((fl & 1) > 0) | ((fl & 4) > 0)) & ((fl & 8) == 0)
        q_flags = ((part & ((fl & 1) > 0)).astype(np.uint8) | (fl & 2) |
                   ((part & ((fl & 4) > 0) & (P["level_r"] != -1)).astype(np.uint8) << 2)).astype(np.uint8)
        lvr = np.maximum(P["level_r"], 0)
        q = dict(qx=P["proj"][:, 0], qy=P["proj"][:, 1], qr=(radius(P["view_cos"], th) * sf[P["level"]]).astype(f32),
                 qmin=P["level"] - 1, qmax=P["level"], qx_r=P["proj"][:, 2], qy_r=P["proj"][:, 3],
                 qr_r=(radius(P["view_cos_r"], 1.0) * sf[lvr]).astype(f32), qmin_r=lvr - 1, qmax_r=lvr, qdesc=P["desc"],
                 flags=q_flags)
        o = oracle_mod.search_by_projection_rig(0, kl, dl, kr, dr, BOUNDS, q, 0.8, False, locked_l=locked[:nl],
                                                locked_r=locked[nl:], l2r=l2r, r2l=r2l)
        fq = np.concatenate([o["feat_query_l"], o["feat_query_r"]])
        assert nm == o["nmatches"] > 400
        assert np.array_equal(holder, np.where(fq >= 0, fq, np.where(locked > 0, -2, -1)))


def _lastframe_case(rng, data, rig):
    sf, pos = data["sf"], data["pos"]
    if rig:
        (kc, dc), (kr, dr) = data["left"][1], data["right"][1]
        kL, dL = data["left"][0]
        shv = np.array(S.frame_shift(0), f32) - np.array(S.frame_shift(1), f32)
    else:
        kc, dc = data["mono"][1]
        kr = dr = None
        kL, dL = data["mono"][0]
        shv = pos[0] - pos[1]
    M = len(kL)
    uv = np.stack([(kL["x"] + shv[0]).astype(f32), (kL["y"] + shv[1]).astype(f32)], 1)
    has = rng.random(M) < 0.85
    obs = rng.random(M) < 0.7
    outl = rng.random(M) < 0.05
    last_flags = (has.astype(np.uint8) | (obs.astype(np.uint8) << 1) | (outl.astype(np.uint8) << 2)).astype(np.uint8)
    angle = kL["angle"].astype(f32).copy()
    rot = rng.random(M) < 0.05       # a few rotated keypoints: minor histogram bins
    angle[rot] = (angle[rot] + f32(95.0)) % f32(360.0)
    return kc, dc, kr, dr, kL, dL, uv, last_flags, angle


@pytest.mark.parametrize("mono,tlc_z,check_ori", [(True, 0.0, True), (False, 1.0, True), (False, -1.0, False),
                                                    (False, 0.0, True)])
def test_search_by_projection_lastframe_mono_camera(oracle_mod, RM, data, mono, tlc_z, check_ori):
    """ORBmatcher.cc:1676-1792, 1864-1887: forward / backward / default octave bands, the mvuRight gate, best only,
    rotation histogram + ComputeThreeMaxima."""
    rng = np.random.default_rng(13)
    sf = data["sf"]
    kc, dc, _, _, kL, dL, uv, last_flags, angle = _lastframe_case(rng, data, False)
    uR = np.where(rng.random(len(kc)) < 0.6, kc["x"] - f32(30.0), f32(-1)).astype(f32)
    locked = (rng.random(len(kc)) < 0.1).astype(np.uint8)
    mbf = f32(30.0)
    th = 15.0 if mono else 7.0
    holder, nm = RM.search_by_projection_lastframe(kc, dc, BOUNDS, sf, last_flags, uv, kL["octave"], angle, dL, th, mono,
                                                   check_ori, u_right=uR, locked=locked, mb=0.1, mbf=mbf, tlc_z=tlc_z)
    inb = (uv[:, 0] >= BOUNDS[0]) & (uv[:, 0] <= BOUNDS[1]) & (uv[:, 1] >= BOUNDS[2]) & (uv[:, 1] <= BOUNDS[3])
    sel = np.nonzero(((last_flags & 1) > 0) & ((last_flags & 4) == 0) & inb)[0]
    octv = kL["octave"][sel]
    fwd, bwd = (tlc_z > 0.1) and not mono, (-tlc_z > 0.1) and not mono
    qmin = octv if fwd else (np.zeros_like(octv) if bwd else octv - 1)
    qmax = np.full_like(octv, -1) if fwd else (octv if bwd else octv + 1)
    args = (kc, dc, BOUNDS, uv[sel, 0], uv[sel, 1], (f32(th) * sf[octv]).astype(f32), qmin, qmax, dL[sel], 1, 0.9)
    kw = dict(u_right=uR, feat_locked=locked, q_ur=(uv[sel, 0] - mbf).astype(f32),
              q_flags=(1 | (last_flags[sel] & 2)).astype(np.uint8), q_angle=angle[sel])
    _, fq0, _ = oracle_mod.search_by_projection(*args, False, **kw)
    _, fq1, onm = oracle_mod.search_by_projection(*args, check_ori, **kw)
    want = np.where(fq0 >= 0, np.where(fq1 >= 0, sel[np.maximum(fq1, 0)], -1), np.where(locked > 0, -2, -1))
    assert nm == onm > 300
    assert np.array_equal(holder, want)
    if check_ori:
        assert (fq0 >= 0).sum() > (fq1 >= 0).sum()      # the rotation check did erase something


@pytest.mark.parametrize("check_ori", [True, False])
def test_search_by_projection_lastframe_rig(oracle_mod, RM, data, check_ori):
    """ORBmatcher.cc:1676-1887 with Nleft != -1: the right block :1794-1862 and the shared rotation histogram."""
    rng = np.random.default_rng(14)
    sf = data["sf"]
    kc, dc, kr, dr, kL, dL, uv, last_flags, angle = _lastframe_case(rng, data, True)
    nl, nr = len(kc), len(kr)
    locked = (rng.random(nl + nr) < 0.1).astype(np.uint8)
    trl_x = f32(-24.0)
    holder, nm = RM.search_by_projection_lastframe(kc, dc, BOUNDS, sf, last_flags, uv, kL["octave"], angle, dL, 15.0,
                                                   True, check_ori, kps_r=kr, desc_r=dr, locked=locked, trl_x=trl_x)
    inb = (uv[:, 0] >= BOUNDS[0]) & (uv[:, 0] <= BOUNDS[1]) & (uv[:, 1] >= BOUNDS[2]) & (uv[:, 1] <= BOUNDS[3])
    sel = np.nonzero(((last_flags & 1) > 0) & ((last_flags & 4) == 0) & inb)[0]
    octv = kL["octave"][sel]
    qr = (f32(15.0) * sf[octv]).astype(f32)
    q = dict(qx=uv[sel, 0], qy=uv[sel, 1], qr=qr, qmin=octv - 1, qmax=octv + 1, qx_r=(uv[sel, 0] + trl_x).astype(f32),
             qy_r=uv[sel, 1], qr_r=qr, qmin_r=octv - 1, qmax_r=octv + 1, qdesc=dL[sel],
             flags=(1 | (last_flags[sel] & 2)).astype(np.uint8), angle=angle[sel])
    kw = dict(locked_l=locked[:nl], locked_r=locked[nl:])
    o0 = oracle_mod.search_by_projection_rig(1, kc, dc, kr, dr, BOUNDS, q, 0.9, False, **kw)
    o1 = oracle_mod.search_by_projection_rig(1, kc, dc, kr, dr, BOUNDS, q, 0.9, check_ori, **kw)
    fq0 = np.concatenate([o0["feat_query_l"], o0["feat_query_r"]])
    fq1 = np.concatenate([o1["feat_query_l"], o1["feat_query_r"]])
    want = np.where(fq0 >= 0, np.where(fq1 >= 0, sel[np.maximum(fq1, 0)], -1), np.where(locked > 0, -2, -1))
    assert nm == o1["nmatches"] > 600
    assert np.array_equal(holder, want)
    assert (o1["feat_query_r"] >= 0).sum() > 200


def _bow_case(oracle_mod, data, rng):
    (k1, d1), (k2, d2) = data["mono"][0], data["mono"][1]
    voc = S.make_vocabulary(10, 4, seed=5, desc_pool=np.concatenate([d1, d2]))
    fv1 = oracle_mod.bow_transform(d1, voc, 2)["fv"]
    fv2 = oracle_mod.bow_transform(d2, voc, 2)["fv"]
    hf1 = ((rng.random(len(k1)) < 0.75).astype(np.uint8) | ((rng.random(len(k1)) < 0.1).astype(np.uint8) << 1)).astype(np.uint8)
    hf2 = ((rng.random(len(k2)) < 0.75).astype(np.uint8) | ((rng.random(len(k2)) < 0.1).astype(np.uint8) << 1)).astype(np.uint8)
    return k1, d1, fv1, hf1, k2, d2, fv2, hf2, voc


@pytest.mark.parametrize("nn_ratio,check_ori", [(0.7, True), (0.9, False)])
def test_search_by_bow(oracle_mod, RM, data, nn_ratio, check_ori):
    """ORBmatcher.cc:223-425 (KeyFrame x Frame, single camera) and :765-905 (KeyFrame x KeyFrame)."""
    rng = np.random.default_rng(15)
    k1, d1, fv1, hf1, k2, d2, fv2, hf2, _ = _bow_case(oracle_mod, data, rng)
    v1, v2 = (hf1 == 1).astype(np.uint8), (hf2 == 1).astype(np.uint8)    # pMP && !pMP->isBad()
    m21, nm = RM.search_by_bow_kf_frame(k1, d1, hf1, fv1, k2, d2, fv2, data["sf"], nn_ratio, check_ori)
    _, o21, onm = oracle_mod.search_by_bow(0, k1, d1, fv1, k2, d2, fv2, nn_ratio, check_ori, v1, None)
    assert nm == onm > 80 and np.array_equal(m21, o21)
    m12, nm = RM.search_by_bow_kf_kf(k1, d1, hf1, fv1, k2, d2, hf2, fv2, data["sf"], nn_ratio, check_ori)
    o12, _, onm = oracle_mod.search_by_bow(1, k1, d1, fv1, k2, d2, fv2, nn_ratio, check_ori, v1, v2)
    assert nm == onm > 50 and np.array_equal(m12, o12)


def test_search_by_bow_rig_frame(oracle_mod, RM, data):
    """ORBmatcher.cc:295-323, 356-380: separate left / right bests, right match without ratio test."""
    rng = np.random.default_rng(16)
    k1, d1, fv1, hf1, _, _, _, _, voc = _bow_case(oracle_mod, data, rng)
    (kl, dl), (kr, dr) = data["left"][0], data["right"][0]
    kcat, dcat = np.concatenate([kl, kr]), np.concatenate([dl, dr])
    fvr = oracle_mod.bow_transform(dcat, voc, 2)["fv"]
    v1 = (hf1 == 1).astype(np.uint8)
    for nn_ratio, ori in ((0.7, True), (0.9, False)):
        m21, nm = RM.search_by_bow_kf_frame(k1, d1, hf1, fv1, kcat, dcat, fvr, data["sf"], nn_ratio, ori, nleft2=len(kl))
        _, _, o21, onm = oracle_mod.search_by_bow_rig(k1, d1, fv1, kcat, dcat, len(kl), fvr, nn_ratio, ori, valid1=v1)
        assert nm == onm > 60 and np.array_equal(m21, o21)
        assert (o21[len(kl):] >= 0).sum() > 20


@pytest.mark.parametrize("only_stereo,coarse", [(False, False), (True, False), (False, True)])
def test_search_for_triangulation(oracle_mod, RM, data, only_stereo, coarse):
    """ORBmatcher.cc:907-1146 (pinhole): skips, last minimum wins, epipole gate, Pinhole::epipolarConstrain with the
    F12 / epipole the reference's own pose algebra produced."""
    rng = np.random.default_rng(17)
    k1, d1, fv1, hf1, k2, d2, fv2, hf2, _ = _bow_case(oracle_mod, data, rng)
    hf1 = (rng.random(len(k1)) < 0.3).astype(np.uint8)
    hf2 = (rng.random(len(k2)) < 0.3).astype(np.uint8)
    u1 = np.where(rng.random(len(k1)) < 0.5, 100.0, -1.0).astype(f32)
    u2 = np.where(rng.random(len(k2)) < 0.5, 100.0, -1.0).astype(f32)
    ang = 0.02
    R2 = np.array([[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]])
    pose1 = np.concatenate([np.eye(3).ravel(), [0, 0, 0]]).astype(f32)
    pose2 = np.concatenate([R2.ravel(), [-0.012, 0.003, 0.05]]).astype(f32)   # mostly forward motion: epipole in the image
    cam = np.array([458.654, 457.296, 367.215, 248.375], f32)
    m12, nm, F, ep = RM.search_for_triangulation(k1, d1, hf1, u1, fv1, k2, d2, hf2, u2, fv2, data["sf"], data["sig"], pose1,
                                                 pose2, cam, only_stereo, coarse, True)
    o12, onm = oracle_mod.search_for_triangulation(k1, d1, fv1, k2, d2, fv2, F, ep, data["sf"], data["sig"], only_stereo,
                                                   coarse, True, hf1, hf2, u1, u2)
    assert 0 < ep[0] < 752 and 0 < ep[1] < 480
    assert nm == onm and np.array_equal(m12, o12)
    assert nm > (20 if not coarse else 100)


def test_search_for_initialization(oracle_mod, RM, data):
    """ORBmatcher.cc:648-763: level-0 keypoints, window, stealing rule, stale histogram entries."""
    (k1, d1), (k2, d2) = data["mono"][0], data["mono"][1]
    prev = np.stack([k1["x"], k1["y"]], 1).astype(f32)
    for win, ratio in ((100, 0.9), (30, 0.7)):
        m12, nm, pm = RM.search_for_initialization(k1, d1, k2, d2, BOUNDS, data["sf"], prev, win, ratio, True)
        o12, onm, opm = oracle_mod.search_for_initialization(k1, d1, k2, d2, BOUNDS, prev, win, ratio, True)
        assert nm == onm > 80 and np.array_equal(m12, o12) and np.array_equal(pm, opm)


@pytest.mark.parametrize("orb_dist,th", [(100, 10.0), (64, 3.0)])
def test_search_by_projection_relocalisation(oracle_mod, RM, data, orb_dist, th):
    """ORBmatcher.cc:1889-2010: levels from PredictScale, band [l-1, l+1], every non-null mvpMapPoints entry blocks
    its feature, best <= ORBdist, rotation histogram."""
    rng = np.random.default_rng(18)
    sf, pos = data["sf"], data["pos"]
    (kK, dK), (kC, dC) = data["mono"][0], data["mono"][1]
    shv = pos[0] - pos[1]
    M = len(kK)
    uv = np.stack([(kK["x"] + shv[0]).astype(f32), (kK["y"] + shv[1]).astype(f32)], 1)
    dist = np.sqrt(uv[:, 0].astype(np.float64) ** 2 + uv[:, 1].astype(np.float64) ** 2 + 1.0)
    level = kK["octave"].astype(np.int64)
    max_dist = (dist * 1.2 ** (level - 0.5)).astype(f32)
    min_dist = (dist * 0.1).astype(f32)
    kf_flags = ((rng.random(M) < 0.85).astype(np.uint8) | ((rng.random(M) < 0.05).astype(np.uint8) << 1) |
                ((rng.random(M) < 0.2).astype(np.uint8) << 2)).astype(np.uint8)
    holds = (rng.random(len(kC)) < 0.2).astype(np.uint8)
    holder, nm = RM.search_by_projection_reloc(kC, dC, BOUNDS, sf, holds, kf_flags, uv, max_dist, min_dist, kK["angle"], dK,
                                               th, orb_dist, True)
    inb = (uv[:, 0] >= BOUNDS[0]) & (uv[:, 0] <= BOUNDS[1]) & (uv[:, 1] >= BOUNDS[2]) & (uv[:, 1] <= BOUNDS[3])
    sel = np.nonzero(((kf_flags & 1) > 0) & ((kf_flags & 2) == 0) & ((kf_flags & 4) == 0) & inb)[0]
    lv = level[sel]
    args = (kC, dC, BOUNDS, uv[sel, 0], uv[sel, 1], (f32(th) * sf[lv]).astype(f32), lv - 1, lv + 1, dK[sel], 1, 0.9)
    kw = dict(feat_locked=holds, q_flags=np.full(len(sel), 3, np.uint8), q_angle=kK["angle"][sel], th_dist=orb_dist)
    _, fq0, _ = oracle_mod.search_by_projection(*args, False, **kw)
    _, fq1, onm = oracle_mod.search_by_projection(*args, True, **kw)
    want = np.where(fq0 >= 0, np.where(fq1 >= 0, sel[np.maximum(fq1, 0)], -1), np.where(holds > 0, -2, -1))
    assert nm == onm > 100
    assert np.array_equal(holder, want)


def test_fuse_candidates(oracle_mod, RM, data):
    """ORBmatcher::Fuse (ORBmatcher.cc:1148-1338): the feature each MapPoint is fused into — the reference's
    Replace / AddObservation decisions — equals the oracle's candidate scoring (best <= TH_LOW)."""
    rng = np.random.default_rng(19)
    sf, sig, pos = data["sf"], data["sig"], data["pos"]
    (kK, dK), (kQ, dQ) = data["mono"][1], data["mono"][0]
    shv = pos[0] - pos[1]
    M = len(kQ)
    uv = np.stack([(kQ["x"] + shv[0]).astype(f32), (kQ["y"] + shv[1]).astype(f32)], 1)
    dist = np.sqrt(uv[:, 0].astype(np.float64) ** 2 + uv[:, 1].astype(np.float64) ** 2 + 1.0)
    level = kQ["octave"].astype(np.int64)
    max_dist, min_dist = (dist * 1.2 ** (level - 0.5)).astype(f32), (dist * 0.1).astype(f32)
    mp_flags = ((rng.random(M) < 0.9).astype(np.uint8) | ((rng.random(M) < 0.05).astype(np.uint8) << 1)).astype(np.uint8)
    bf = f32(40.0)
    uR = np.where(rng.random(len(kK)) < 0.5, kK["x"] - bf, f32(-1)).astype(f32)
    holder_flags = (rng.random(len(kK)) < 0.5).astype(np.uint8)
    kb = (0.0, 752.0, 0.0, 480.0)
    best, nf = RM.fuse(kK, dK, holder_flags, uR, kb, sf, sig, bf, mp_flags, uv, max_dist, min_dist, dQ, 3.0)
    inimg = (uv[:, 0] >= kb[0]) & (uv[:, 0] < kb[1]) & (uv[:, 1] >= kb[2]) & (uv[:, 1] < kb[3])   # KeyFrame::IsInImage
    sel = np.nonzero(((mp_flags & 1) > 0) & ((mp_flags & 2) == 0) & inimg)[0]
    lv = level[sel]
    inv_sig = (f32(1.0) / sig).astype(f32)
    bi, bd = oracle_mod.fuse_candidates(kK, dK, kb, inv_sig, uv[sel, 0], uv[sel, 1], (f32(3.0) * sf[lv]).astype(f32), lv,
                                        dQ[sel], u_right=uR, q_ur=(uv[sel, 0] - bf).astype(f32))
    want = np.full(M, -1, np.int32)
    ok = (bi >= 0) & (bd <= 50)
    want[sel[ok]] = bi[ok]
    assert nf == int(ok.sum()) > 300
    assert np.array_equal(best, want)


@pytest.mark.parametrize("th", [3.0, 6.0])
def test_fuse_sim3(oracle_mod, RM, data, th):
    """ORBmatcher::Fuse(KeyFrame*, Sim3f&, vpPoints, th, vpReplacePoint) (ORBmatcher.cc:1340-1455): no chi-square gate,
    band [l-1, l], best <= TH_LOW; a feature that holds a point answers through vpReplacePoint (nothing recorded when
    that point is bad, still counted), a free one is taken by the fused point — and later points see it as held."""
    rng = np.random.default_rng(23)
    sf, pos = data["sf"], data["pos"]
    (kK, dK), (kQ, dQ) = data["mono"][1], data["mono"][0]
    shv = pos[0] - pos[1]
    M, n = len(kQ), len(kK)
    # every point twice: the second copy finds the feature its twin has just taken
    kQ, dQ = np.concatenate([kQ, kQ]), np.concatenate([dQ, dQ])
    M *= 2
    uv = np.stack([(kQ["x"] + shv[0]).astype(f32), (kQ["y"] + shv[1]).astype(f32)], 1)
    dist = np.sqrt(uv[:, 0].astype(np.float64) ** 2 + uv[:, 1].astype(np.float64) ** 2 + 1.0)
    level = kQ["octave"].astype(np.int64)
    max_dist, min_dist = (dist * 1.2 ** (level - 0.5)).astype(f32), (dist * 0.1).astype(f32)
    mp_flags = ((rng.random(M) < 0.05).astype(np.uint8) << 1).astype(np.uint8)
    own = rng.random(M) < 0.1
    own_feat = np.full(M, -1, np.int32)
    own_feat[own] = rng.permutation(n)[:int(own.sum())]
    mp_flags[own] |= 4
    holder_flags = (rng.random(n) < 0.4).astype(np.uint8)
    holder_flags[own_feat[own]] = 0
    held_bad = ((rng.random(n) < 0.15) & (holder_flags > 0)).astype(np.uint8)
    best, repl, nf = RM.fuse_sim3(kK, dK, holder_flags, held_bad, BOUNDS, sf, mp_flags, own_feat, uv, max_dist, min_dist,
                                  dQ, th)
    bad = (mp_flags & 2) > 0
    inimg = (uv[:, 0] >= BOUNDS[0]) & (uv[:, 0] < BOUNDS[1]) & (uv[:, 1] >= BOUNDS[2]) & (uv[:, 1] < BOUNDS[3])
    sel = np.nonzero(~bad & ~(own & ~bad) & inimg)[0]      # spAlreadyFound holds the keyframe's own non-bad points
    lv = level[sel]
    bi, bd = oracle_mod.fuse_candidates(kK, dK, BOUNDS, np.zeros(8, f32), uv[sel, 0], uv[sel, 1],
                                        (f32(th) * sf[lv]).astype(f32), lv, dQ[sel])
    holder = np.full(n, -1, np.int64)                      # -1 free, -2 held by a foreign point, i >= 0 by query point i
    holder[holder_flags > 0] = -2
    holder[own_feat[own]] = np.nonzero(own)[0]
    want, wrepl, wnf = np.full(M, -1, np.int32), np.zeros(M, np.uint8), 0
    for i, j, d in zip(sel, bi, bd):
        if j < 0 or d > 50:
            continue
        wnf += 1
        if holder[j] == -1:
            holder[j] = i
            want[i] = j
        elif holder[j] == -2:
            if not held_bad[j]:
                want[i], wrepl[i] = j, 1
        elif not bad[holder[j]]:
            want[i], wrepl[i] = -3 - holder[j], 1
    assert nf == wnf > 400
    assert np.array_equal(best, want) and np.array_equal(repl, wrepl)
    assert (wrepl > 0).sum() > 100 and (want <= -3).sum() > 50


@pytest.mark.parametrize("variant,th,ratio", [(0, 8, 1.0), (1, 4, 1.5), (0, 5, 0.8)])
def test_search_by_projection_sim3(oracle_mod, RM, data, variant, th, ratio):
    """ORBmatcher.cc:427-532 / :534-646 (loop closing / merging): KeyFrame grid without level filter, band [l-1, l]
    inside the loop, vpMatched blocks its features, best <= TH_LOW * ratioHamming, already-found points skipped."""
    rng = np.random.default_rng(20 + variant)
    sf, pos = data["sf"], data["pos"]
    (kK, dK), (kQ, dQ) = data["mono"][1], data["mono"][0]
    shv = pos[0] - pos[1]
    M, n = len(kQ), len(kK)
    uv = np.stack([(kQ["x"] + shv[0]).astype(f32), (kQ["y"] + shv[1]).astype(f32)], 1)
    dist = np.sqrt(uv[:, 0].astype(np.float64) ** 2 + uv[:, 1].astype(np.float64) ** 2 + 1.0)
    level = kQ["octave"].astype(np.int64)
    max_dist, min_dist = (dist * 1.2 ** (level - 0.5)).astype(f32), (dist * 0.1).astype(f32)
    mp_flags = (1 | ((rng.random(M) < 0.05).astype(np.uint8) << 1)).astype(np.uint8)
    matched_init = np.full(n, -1, np.int32)
    held = rng.permutation(n)[:n // 4]
    matched_init[held[:len(held) // 2]] = -2                                  # foreign points
    found = rng.permutation(M)[:len(held) - len(held) // 2]                   # query points already found
    matched_init[held[len(held) // 2:]] = found
    matched, nm = RM.search_by_projection_sim3(variant, kK, dK, BOUNDS, sf, matched_init, mp_flags, uv, max_dist, min_dist,
                                               dQ, th, ratio)
    inimg = (uv[:, 0] >= BOUNDS[0]) & (uv[:, 0] < BOUNDS[1]) & (uv[:, 1] >= BOUNDS[2]) & (uv[:, 1] < BOUNDS[3])
    already = np.zeros(M, bool); already[found] = True
    sel = np.nonzero(((mp_flags & 2) == 0) & ~already & inimg)[0]
    lv = level[sel]
    th_dist = int(np.floor(f32(50) * f32(ratio)))
    _, fq, onm = oracle_mod.search_by_projection(kK, dK, BOUNDS, uv[sel, 0], uv[sel, 1], (f32(th) * sf[lv]).astype(f32),
                                                 lv - 1, lv, dQ[sel], 1, 0.75, False,
                                                 feat_locked=(matched_init != -1).astype(np.uint8),
                                                 q_flags=np.full(len(sel), 3, np.uint8), th_dist=th_dist)
    want = np.where(fq >= 0, sel[np.maximum(fq, 0)], matched_init)
    assert nm == onm > 100
    assert np.array_equal(matched, want)


def test_search_by_sim3(oracle_mod, RM, data):
    """ORBmatcher::SearchBySim3 (ORBmatcher.cc:1457-1674): both directions scored like Fuse without the chi-square gate
    (best <= TH_HIGH, band [l-1, l]), then the two-way agreement; points matched on entry are skipped on both sides."""
    rng = np.random.default_rng(23)
    sf, pos = data["sf"], data["pos"]
    (k1, d1), (k2, d2) = data["mono"][0], data["mono"][1]
    n1, n2 = len(k1), len(k2)
    s12 = pos[0] - pos[1]          # keyframe-1 points seen in keyframe 2
    uv1 = np.stack([(k1["x"] + s12[0]).astype(f32), (k1["y"] + s12[1]).astype(f32)], 1)
    uv2 = np.stack([(k2["x"] - s12[0]).astype(f32), (k2["y"] - s12[1]).astype(f32)], 1)

    def ranges(uv, k):
        dist = np.sqrt(uv[:, 0].astype(np.float64) ** 2 + uv[:, 1].astype(np.float64) ** 2 + 1.0)
        return (dist * 1.2 ** (k["octave"].astype(np.int64) - 0.5)).astype(f32), (dist * 0.1).astype(f32)

    mx1, mn1 = ranges(uv1, k1)
    mx2, mn2 = ranges(uv2, k2)
    hf1 = ((rng.random(n1) < 0.85).astype(np.uint8) | ((rng.random(n1) < 0.05).astype(np.uint8) << 1)).astype(np.uint8)
    hf2 = ((rng.random(n2) < 0.85).astype(np.uint8) | ((rng.random(n2) < 0.05).astype(np.uint8) << 1)).astype(np.uint8)
    init12 = np.full(n1, -1, np.int32)
    pre = rng.permutation(n1)[:60]
    init12[pre] = rng.permutation(np.nonzero(hf2 & 1)[0])[:60]
    th = 7.5
    match12, nf = RM.search_by_sim3(k1, d1, hf1, uv1, mx1, mn1, k2, d2, hf2, uv2, mx2, mn2, BOUNDS, sf, init12, th)

    def direction(kq, dq, hfq, uv, skip, kt, dt):
        inimg = (uv[:, 0] >= BOUNDS[0]) & (uv[:, 0] < BOUNDS[1]) & (uv[:, 1] >= BOUNDS[2]) & (uv[:, 1] < BOUNDS[3])
        sel = np.nonzero(((hfq & 1) > 0) & ((hfq & 2) == 0) & ~skip & inimg)[0]
        lv = kq["octave"][sel].astype(np.int64)
        bi, bd = oracle_mod.fuse_candidates(kt, dt, BOUNDS, np.zeros(8, f32), uv[sel, 0], uv[sel, 1],
                                            (f32(th) * sf[lv]).astype(f32), lv, dq[sel])
        out = np.full(len(kq), -1, np.int64)
        ok = (bi >= 0) & (bd <= 100)
        out[sel[ok]] = bi[ok]
        return out

    skip1 = init12 >= 0
    skip2 = np.zeros(n2, bool); skip2[init12[init12 >= 0]] = True
    m1 = direction(k1, d1, hf1, uv1, skip1, k2, d2)
    m2 = direction(k2, d2, hf2, uv2, skip2, k1, d1)
    want = init12.astype(np.int64).copy()
    agree = (m1 >= 0) & (m2[np.maximum(m1, 0)] == np.arange(n1))
    want[agree] = m1[agree]
    assert nf == int(agree.sum()) > 300
    assert np.array_equal(match12, want)


@pytest.mark.parametrize("only_stereo,coarse", [(False, False), (False, True), (True, False)])
def test_search_for_triangulation_rig_loop(oracle_mod, RM, data, only_stereo, coarse):
    """ORBmatcher.cc:907-1146 with mpCamera2 (rig keyframes): skips, no stereo / epipole gates, the camera and
    (R12, t12) selection :1036-1062 (verified call by call by the table cameras), last minimum wins.  The outcome of
    KannalaBrandt8::epipolarConstrain is an input (random table) on both sides."""
    rng = np.random.default_rng(24)
    (kl1, dl1), (kr1, dr1) = data["left"][0], data["right"][0]
    (kl2, dl2), (kr2, dr2) = data["left"][1], data["right"][1]
    k1, d1 = np.concatenate([kl1, kr1]), np.concatenate([dl1, dr1])
    k2, d2 = np.concatenate([kl2, kr2]), np.concatenate([dl2, dr2])
    voc = S.make_vocabulary(10, 4, seed=7, desc_pool=np.concatenate([d1, d2]))
    fv1 = oracle_mod.bow_transform(d1, voc, 2)["fv"]
    fv2 = oracle_mod.bow_transform(d2, voc, 2)["fv"]
    hf1 = (rng.random(len(k1)) < 0.3).astype(np.uint8)
    hf2 = (rng.random(len(k2)) < 0.3).astype(np.uint8)
    gate = (rng.random((len(k1), len(k2))) < 0.5).astype(np.uint8)
    ang = 0.03
    R2 = np.array([[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]])
    pose1 = np.concatenate([np.eye(3).ravel(), [0.01, 0.02, 0.03]]).astype(f32)
    pose2 = np.concatenate([R2.ravel(), [-0.3, 0.04, 0.08]]).astype(f32)
    pose_rl = np.concatenate([np.eye(3).ravel(), [-0.101, 0.002, 0.001]]).astype(f32)
    m12, nm, bad = RM.search_for_triangulation_rig(k1, d1, len(kl1), hf1, fv1, k2, d2, len(kl2), hf2, fv2, data["sf"], pose1,
                                                   pose2, pose_rl, gate, only_stereo, coarse, True)
    o12, onm = oracle_mod.search_for_triangulation_rig(k1, d1, fv1, k2, d2, fv2, gate, only_stereo, coarse, True, hf1, hf2)
    assert bad == 0
    assert nm == onm and np.array_equal(m12, o12)
    assert nm == 0 if only_stereo else nm > 100
