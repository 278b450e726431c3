"""oracle/_ref/libref_frame.so: Frame::ComputeStereoMatches, ComputeStereoFromRGBD, isInFrustum and isInFrustumChecks
cut verbatim out of the REFERENCE's src/Frame.cc and compiled against the stand-in prelude (recipe `make -C oracle
ref_frame`, oracle/ref_shim_matcher/frame_ref_api.cpp), against the oracle's restatements.

This pins SURVEY §8 rows m11 (stereo matching: row table, Hamming gate, 11x11 SAD search, parabola, disparity gates,
median cut), (f).1 (RGB-D depth) and (f).4 (frustum test: gate order, the early mTrackProjX/Y write, the fields left
untouched on failure, PredictScale clamps, the two-camera composition) to the reference's own text.  The 3-vector
arithmetic inside isInFrustum runs in the stand-in Eigen (plain fp32, uncontracted), so that part pins the logic under
the fp contract of DESIGN.md §2, not Eigen's code generation.

Needs either /root/reference (build container) or the prebuilt oracle/_ref/*.so (GPU box); else skipped."""
import numpy as np
import pytest

from orb_slam3_vio_fixes_b200 import synthetic as S
from test_frame_geometry import frustum_case
import rig_geometry as G

f32 = np.float32


@pytest.fixture(scope="module")
def RF(oracle_mod):
    from oracle import ref_frame as R
    if not R.available():
        pytest.skip("neither /root/reference nor a prebuilt oracle/_ref")
    R.build()
    return R


@pytest.mark.parametrize("cfg,nfeat,mb,mbf", [("C2", 1200, 0.11, 47.9), ("C4", 2000, 0.537, 386.1448)])
def test_compute_stereo_matches(oracle_mod, RF, cfg, nfeat, mb, mbf):
    """Frame.cc:811-981 on extractor output of both eyes (C2: EuRoC-like baseline; C4: KITTI, maxD = fx)."""
    for t in range(2):
        el, er = oracle_mod.Extractor(nfeat), oracle_mod.Extractor(nfeat)
        fl, fr = S.make_frames(cfg, t + 1)[t], S.make_frames(cfg, t + 1, right=True)[t]
        kl, dl, _ = el.extract(fl, (0, 0))
        kr, dr, _ = er.extract(fr, (0, 0))
        pl = [el.pyramid(l) for l in range(8)]
        pr = [er.pyramid(l) for l in range(8)]
        bi, bd, ur, dp, kept = oracle_mod.stereo_matches(kl, dl, kr, dr, el.scale_factors, el.inv_scale_factors, pl, pr,
                                                         mb, mbf)
        rur, rdp, rkept = RF.compute_stereo_matches(kl, dl, kr, dr, el.scale_factors, el.inv_scale_factors, pl, pr, mb,
                                                    mbf)
        assert rkept == kept > 100
        assert np.array_equal(rur, ur)
        assert np.array_equal(rdp, dp)
        assert (ur[dp > 0] >= 0).all() and (dp <= 0).sum() > 50      # both outcomes occur


def test_rgbd_depth(oracle_mod, RF):
    """Frame.cc:984-1005: depth read at the distorted keypoint (float pixel truncated), uRight from the undistorted."""
    rng = np.random.default_rng(5)
    ex = oracle_mod.Extractor(1000)
    k, _, _ = ex.extract(S.make_frames("C3", 1)[0], (0, 0))
    ku = k.copy()
    ku["x"] = (k["x"] + rng.normal(0, 0.7, len(k))).astype(f32)
    depth = (rng.random((480, 640)) * 6).astype(f32)
    depth[rng.random(depth.shape) < 0.3] = 0
    ur, dp = oracle_mod.rgbd_depth(k, ku, depth, 40.0)
    rur, rdp = RF.rgbd_depth(k, ku, depth, 40.0)
    assert np.array_equal(rur, ur) and np.array_equal(rdp, dp)
    assert 0.5 * len(k) < (dp > 0).sum() < len(k)


FIELDS = ("in_view", "proj_x", "proj_y", "proj_xr", "depth", "view_cos", "level")


def _pose18(pose15):
    R = pose15[:9].reshape(3, 3)
    return np.concatenate([pose15, R.T.reshape(-1)]).astype(f32)


@pytest.mark.parametrize("seed", [31, 32, 33])
def test_is_in_frustum_single_camera(oracle_mod, RF, seed):
    """Frame.cc:512-574 (Nleft == -1): every tracking field, including the mTrackProjX/Y written before the distance
    gate and everything a failed gate leaves untouched."""
    n = 40000
    pos, normal, mn, mx, pose, cam = frustum_case(n, seed)
    bounds = (0.0, 752.0, 0.0, 480.0)
    log_sf = float(np.log(f32(1.2)).astype(f32))
    want = oracle_mod.is_in_frustum(pos, normal, mn, mx, pose, cam, 47.9, bounds, 0.5, log_sf, 8)
    got = RF.is_in_frustum(pos, normal, mn, mx, _pose18(pose), cam[1:5], 47.9, bounds, 0.5, log_sf, 8)
    assert got["count"] == int(want["in_view"].sum()) > 0.1 * n
    for k in FIELDS:
        assert np.array_equal(got[k], want[k]), k
    early = (want["in_view"] == 0) & (want["proj_x"] >= 0)
    assert early.sum() > 100                                         # projected inside the image, rejected later


def test_is_in_frustum_two_camera_rig(oracle_mod, RF):
    """Frame.cc:575-585 + :1168-1242 (Nleft != -1): left and right checks, the right pose composed from mTrl / mTlr."""
    n = 40000
    pos, normal, mn, mx, pose, cam = frustum_case(n, 41)
    rng = np.random.default_rng(42)
    ang = 0.05
    Rrl = np.array([[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]])
    trl_t = np.array([-0.11, 0.002, 0.001])
    trl = np.concatenate([Rrl.reshape(-1), trl_t]).astype(f32)
    tlr = (-Rrl.T @ trl_t).astype(f32)
    cam_r = (457.587, 456.134, 379.999, 255.238)
    bounds = (0.0, 752.0, 0.0, 480.0)
    log_sf = float(np.log(f32(1.2)).astype(f32))
    got = RF.is_in_frustum(pos, normal, mn, mx, _pose18(pose), cam[1:5], 47.9, bounds, 0.5, log_sf, 8, cam_r=cam_r, trl=trl,
                           tlr=tlr)
    left = oracle_mod.is_in_frustum(pos, normal, mn, mx, pose, cam, 47.9, bounds, 0.5, log_sf, 8, fisheye_rig=True)
    right = oracle_mod.is_in_frustum(pos, normal, mn, mx, got["right_pose"], (0,) + cam_r, 47.9, bounds, 0.5, log_sf, 8,
                                     fisheye_rig=True)
    # the composed right pose is the textbook one (within fp32 rounding of the two products)
    R, t = pose[:9].reshape(3, 3).astype(np.float64), pose[9:12].astype(np.float64)
    assert np.allclose(got["right_pose"][:9].reshape(3, 3), Rrl @ R, atol=1e-6)
    assert np.allclose(got["right_pose"][9:12], Rrl @ t + trl_t, atol=1e-6)
    assert np.allclose(got["right_pose"][12:], -(Rrl @ R).T @ (Rrl @ t + trl_t), atol=1e-5)
    for k in ("in_view", "proj_x", "proj_y", "depth", "view_cos", "level"):
        assert np.array_equal(got[k], left[k]), k
    for k, kr in (("in_view", "in_view_r"), ("proj_x", "proj_xr_r"), ("proj_y", "proj_yr"), ("depth", "depth_r"),
                  ("view_cos", "view_cos_r"), ("level", "level_r")):
        assert np.array_equal(got[kr], right[k]), kr
    both = (got["in_view"] > 0) & (got["in_view_r"] > 0)
    only = (got["in_view"] > 0) ^ (got["in_view_r"] > 0)
    assert both.sum() > 0.1 * n and only.sum() > 100
    assert got["count"] == int(((got["in_view"] > 0) | (got["in_view_r"] > 0)).sum())


@pytest.mark.parametrize("seed,accept", [(51, 0.8), (52, 0.3)])
def test_stereo_fisheye_index_logic(oracle_mod, RF, seed, accept):
    """Frame.cc:1126-1166 with TriangulateMatches answered from a table: which pairs reach the triangulation (knn-2,
    ratio 0.7), the depth > 0.0001 gate, the left / right tables with "last left index wins", mvDepth.  The numpy
    restatement checked here (tests/rig_geometry.py) is what tests/test_tri_rig.py holds the CUDA function to."""
    rng = np.random.default_rng(seed)
    n_pts = 700
    base = rng.integers(0, 256, (n_pts, 32), dtype=np.uint8)

    def feats(frac, dup):
        seen = rng.permutation(np.nonzero(rng.random(n_pts) < frac)[0])
        seen = np.concatenate([seen, seen[len(seen) - dup:]])        # some 3-D points twice: several lefts want one right
        k = np.zeros(len(seen), oracle_mod.KP_DTYPE)
        k["x"], k["y"], k["size"] = rng.random(len(seen)) * 500, rng.random(len(seen)) * 500, 31
        k["octave"] = rng.integers(0, 8, len(seen))
        flips = rng.random((len(seen), 256)) < 0.05
        return k, base[seen] ^ np.packbits(flips, axis=1, bitorder="little")

    kl, dl = feats(0.8, 60)
    kr, dr = feats(0.8, 0)
    mono_l, mono_r = 90, 60
    sig = (np.cumprod([1.0] + [1.2] * 7) ** 2).astype(f32)
    tri = np.where(rng.random((len(kl), len(kr))) < accept, rng.random((len(kl), len(kr))) * 20 + 0.5, -1.0).astype(f32)
    tri[rng.random(tri.shape) < 0.02] = f32(0.00005)                 # below the 0.0001 gate
    l2r, r2l, depth = RF.stereo_fisheye_matches(kl, dl, mono_l, kr, dr, mono_r, sig, tri)
    qi, ti = G.fisheye_ratio_candidates(dl, dr, mono_l, mono_r)
    ok = tri[qi, ti] > f32(0.0001)
    wl2r, wr2l = G.fisheye_tables(len(kl), len(kr), qi, ti, ok)
    assert np.array_equal(l2r, wl2r) and np.array_equal(r2l, wr2l)
    wdepth = np.full(len(kl), -1, f32)
    wdepth[qi[ok]] = tri[qi[ok], ti[ok]]
    assert np.array_equal(depth, wdepth)
    assert ok.sum() > 60 and (~ok).sum() > 20 and (l2r[:mono_l] == -1).all()
    stolen = (wl2r >= 0) & (wr2l[np.maximum(wl2r, 0)] != np.arange(len(kl)))
    assert stolen.sum() > 3                                          # a right feature claimed by two lefts


@pytest.mark.parametrize("cfg,bounds", [("C1", (0.0, 752.0, 0.0, 480.0)), ("C1", (-13.4, 771.2, -9.7, 493.1)),
                                        ("C5", (0.0, 3840.0, 0.0, 2160.0))])
def test_grid_and_area_query(oracle_mod, RF, cfg, bounds):
    """Row m10: Frame::AssignFeaturesToGrid / PosInGrid / GetFeaturesInArea (Frame.cc:385-416, 657-735) — the candidates
    of every window and the ORDER they come in (it decides the ties of every matcher loop): levels open / lower bound
    only / closed band, windows that leave the image, undistorted bounds that are not the image rectangle."""
    rng = np.random.default_rng(61)
    ex = oracle_mod.Extractor(1000 if cfg == "C1" else 5000)
    k, _, _ = ex.extract(S.make_frames(cfg, 1)[0], (0, 0))
    w, h = bounds[1] - bounds[0], bounds[3] - bounds[2]
    nq = 6000
    qx = (bounds[0] + rng.random(nq) * w * 1.1 - 0.05 * w).astype(f32)
    qy = (bounds[2] + rng.random(nq) * h * 1.1 - 0.05 * h).astype(f32)
    qr = (rng.choice([4.0, 12.0, 15.0, 40.0, 120.0], nq) * rng.choice([1.0, 1.2, 1.728], nq)).astype(f32)
    kind = rng.integers(0, 4, nq)
    lv = rng.integers(0, 8, nq)
    qmin = np.where(kind == 0, -1, np.where(kind == 1, lv, lv - 1)).astype(np.int32)
    qmax = np.where(kind <= 1, -1, np.where(kind == 2, lv, lv + 1)).astype(np.int32)
    off, idx = oracle_mod.grid_queries(k, bounds, qx, qy, qr, qmin, qmax)
    roff, ridx = RF.grid_queries(k, bounds, qx, qy, qr, qmin, qmax)
    assert np.array_equal(roff, off) and np.array_equal(ridx, idx)
    assert len(idx) > 20 * nq // 10 and (np.diff(off) == 0).sum() > 50


def test_compute_distinctive_descriptors(oracle_mod, RF):
    """MapPoint.cc:328-402: the observed descriptor with the least median distance to the others (median = element
    0.5*(N-1) of the sorted row, first minimum wins), bad keyframes skipped, left and right index of a rig keyframe."""
    rng = np.random.default_rng(71)
    base = rng.integers(0, 256, (40, 32), dtype=np.uint8)
    for trial in range(300):
        n_obs = int(rng.integers(1, 14))
        rows = int(rng.integers(n_obs, 3 * n_obs + 1))
        flips = rng.random((rows, 256)) < rng.choice([0.02, 0.1, 0.3])
        desc = base[trial % 40][None, :] ^ np.packbits(flips, axis=1, bitorder="little")
        if trial % 5 == 0:
            desc[rng.integers(0, rows)] = desc[0]                    # ties between identical descriptors
        first = rng.integers(0, rows, n_obs).astype(np.int32)
        second = np.where(rng.random(n_obs) < 0.3, rng.integers(0, rows, n_obs), -1).astype(np.int32)
        bad = (rng.random(n_obs) < 0.15).astype(np.uint8)
        got = RF.distinctive_descriptor(desc, first, second, bad)
        order = []
        for o in range(n_obs):
            if bad[o]:
                continue
            order.append(first[o])
            if second[o] >= 0:
                order.append(second[o])
        if not order:
            assert got is None
            continue
        best = oracle_mod.distinctive_descriptor(desc[order])
        assert np.array_equal(got, desc[order][best]), trial


def _camera_coords(pose, pos):
    """Rcw * P + tcw the way the oracle / the product evaluate it (3-term dot = c0 + (c1 + c2), uncontracted fp32)."""
    R, t = pose[:9].reshape(3, 3), pose[9:12]
    d3 = lambda i: (R[i, 0] * pos[:, 0] + (R[i, 1] * pos[:, 1] + R[i, 2] * pos[:, 2])).astype(f32)
    return np.stack([(d3(i) + t[i]).astype(f32) for i in range(3)], 1)


@pytest.mark.parametrize("model", [0, 1])
def test_camera_project_is_the_reference_text(oracle_mod, RF, model):
    """Pinhole::project (Pinhole.cpp:49-55) and KannalaBrandt8::project (KannalaBrandt8.cpp:67-85), cut out of the
    reference and compiled with the name lookup of its translation units (`cos` / `sin` on a float are the double
    functions there): bit-exact with the projection inside the oracle's frustum test."""
    n = 200000
    pos, normal, mn, mx, pose, cam = frustum_case(n, 5, model=model)
    bounds = (0.0, 752.0, 0.0, 480.0) if model == 0 else (0.0, 512.0, 0.0, 512.0)
    log_sf = float(np.log(f32(1.2)).astype(f32))
    o = oracle_mod.is_in_frustum(pos, normal, mn, mx, pose, cam, 47.9, bounds, 0.5, log_sf, 8, fisheye_rig=(model == 1))
    iv = o["in_view"] > 0
    uv = RF.camera_project(model, cam[1:], _camera_coords(pose, pos)[iv])
    assert iv.sum() > 50000
    assert np.array_equal(uv[:, 0], o["proj_x"][iv]) and np.array_equal(uv[:, 1], o["proj_y"][iv])


def test_kb8_unproject_is_the_reference_text(harness, RF):
    """KannalaBrandt8::unproject / unprojectEig (KannalaBrandt8.cpp:111-146: 10 Newton steps, precision 1e-6,
    std::tan) against the product's orbx_kb8_unproject compiled for the host (tests/cpu_harness): bit-exact."""
    import ctypes as C
    rng = np.random.default_rng(81)
    cam = np.array([190.978, 190.973, 254.932, 256.897, 0.0034, 0.00071, -0.0020, 0.00020], f32)   # TUM-VI 512 fisheye
    uv = (rng.random((100000, 2)) * 512).astype(f32)
    uv[:50] = cam[2:4]                                              # the principal point: theta_d = 0
    want = RF.kb8_unproject(cam, uv)
    got = np.zeros_like(want)
    harness.h_kb8_unproject.restype = None
    harness.h_kb8_unproject.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    harness.h_kb8_unproject(cam.ctypes.data, uv.ctypes.data, len(uv), got.ctypes.data)
    assert np.array_equal(got, want)
    assert np.isfinite(want).all() and (want[:, 2] == 1).all()
