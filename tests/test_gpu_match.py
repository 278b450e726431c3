"""GPU parity of the matcher kernels (through the C ABI) against the oracle's restated loops."""
import numpy as np
import pytest

from orb_slam3_vio_fixes_b200 import synthetic as S
from orb_slam3_vio_fixes_b200.extractor import ORBextractor
from orb_slam3_vio_fixes_b200.matcher import ORBmatcher, results_to_numpy, TIE_FIRST_MIN, TIE_LAST_MIN

pytestmark = pytest.mark.gpu
ORB = dict(scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7)


def _t(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _same_results(got, ref, with_second=True):
    assert np.array_equal(got["best_idx"], ref["best_idx"])
    assert np.array_equal(got["best_dist"], ref["best_dist"])
    assert np.array_equal(got["best_level"], ref["best_level"])
    if with_second:
        assert np.array_equal(got["second_dist"], ref["second_dist"])
        assert np.array_equal(got["second_level"], ref["second_level"])


@pytest.mark.parametrize("tie_rule", [TIE_FIRST_MIN, TIE_LAST_MIN])
def test_argmin2_csr_random_with_ties(oracle_mod, tie_rule):
    rng = np.random.default_rng(7)
    nq, nr = 700, 900
    base = rng.integers(0, 256, (40, 32), dtype=np.uint8)
    # few distinct descriptors with 0-2 flipped bits => plenty of exact distance ties
    r = base[rng.integers(0, 40, nr)].copy()
    q = base[rng.integers(0, 40, nq)].copy()
    for a in (r, q):
        flip = rng.integers(0, 3, len(a))
        for i, k in enumerate(flip):
            for _ in range(k):
                a[i, rng.integers(0, 32)] ^= np.uint8(1 << rng.integers(0, 8))
    lens = rng.integers(0, 90, nq)
    lens[:5] = [0, 1, 2, 33, 64]
    off = np.zeros(nq + 1, np.int32)
    off[1:] = np.cumsum(lens)
    idx = rng.integers(0, nr, off[-1]).astype(np.int32)
    lv = rng.integers(0, 8, nr).astype(np.int8)
    ref = oracle_mod.argmin2_csr(q, r, off, idx, lv, tie_rule)
    got = results_to_numpy(ORBmatcher.argmin2_csr(_t(q), _t(r), _t(off), _t(idx), _t(lv), tie_rule))
    _same_results(got, ref, with_second=(tie_rule == TIE_FIRST_MIN))


def test_descriptor_distance_pairs(oracle_mod):
    rng = np.random.default_rng(12)
    a = rng.integers(0, 256, (500, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (700, 32), dtype=np.uint8)
    b[3] = a[9]
    ia = rng.integers(0, 500, 4000).astype(np.int32)
    ib = rng.integers(0, 700, 4000).astype(np.int32)
    ia[0], ib[0] = 9, 3
    got = ORBmatcher.descriptor_distance_pairs(_t(a), _t(b), _t(ia), _t(ib)).cpu().numpy()
    ref = np.array([oracle_mod.descriptor_distance(a[i], b[j]) for i, j in zip(ia, ib)], np.int32)
    assert np.array_equal(got, ref) and got[0] == 0
    assert all(ORBmatcher.DescriptorDistance(a[i], b[j]) == r for i, j, r in zip(ia[:50], ib[:50], ref[:50]))
    got = ORBmatcher.descriptor_distance_pairs(_t(a), _t(b[:500])).cpu().numpy()      # row i against row i
    assert np.array_equal(got, [oracle_mod.descriptor_distance(a[i], b[i]) for i in range(500)])


def test_knn2_dense(oracle_mod):
    rng = np.random.default_rng(8)
    q = rng.integers(0, 256, (333, 32), dtype=np.uint8)
    r = rng.integers(0, 256, (1201, 32), dtype=np.uint8)
    r[100] = r[7]
    q[5] = r[7]
    idx, dist = ORBmatcher.knn2_dense(_t(q), _t(r))
    ridx, rdist = oracle_mod.knn2_dense(q, r)
    assert np.array_equal(idx.cpu().numpy(), ridx) and np.array_equal(dist.cpu().numpy(), rdist)
    idx, dist = ORBmatcher.knn2_dense(_t(q[:3]), _t(r[:1]))   # a single reference: no second
    assert np.array_equal(idx.cpu().numpy()[:, 1], [-1, -1, -1])


def _extract_batch(ex, frames, lap):
    import torch
    out = ex.extract_batch_device(_t(frames), lap)
    torch.cuda.synchronize()
    return out


def test_grid_search_area_matches_reference_loops(oracle_mod):
    """C1 match shape: frame t-1 keypoints shifted into frame t, window 15*scale, band [o-1,o+1]."""
    import torch
    from bench import shifts_for
    B = 4
    frames = S.make_frames("C1", B)
    ex = ORBextractor(1000, max_batch=B, **ORB)
    out = _extract_batch(ex, frames, (0, 1000))
    sf = ex.GetScaleFactors()
    bounds = (0.0, 752.0, 0.0, 480.0)
    shift = shifts_for(0, B)
    src = ((np.arange(B) - 1) % B).astype(np.int32)
    grid = ORBmatcher.grid_build(out["kps"], out["n"], bounds)
    qb = ORBmatcher.queries_shift(out["kps"], out["desc"], out["n"], _t(src), _t(shift), 15.0, sf, (-1, 1))
    kps = ORBextractor.kps_from_tensor(out["kps"])
    desc = out["desc"].cpu().numpy()
    n = out["n"].cpu().numpy()
    for tie in (TIE_FIRST_MIN, TIE_LAST_MIN):
        res = results_to_numpy(ORBmatcher.search_area(out["kps"], out["desc"], out["n"], grid, bounds, qb["qx"],
                                                      qb["qy"], qb["qr"], qb["qmin"], qb["qmax"], qb["qdesc"],
                                                      qb["nq"], tie))
        total_cand = 0
        for f in range(B):
            s = src[f]
            k_ref, k_q = kps[f, :n[f]], kps[s, :n[s]]
            qx = k_q["x"] + shift[f, 0]
            qy = k_q["y"] + shift[f, 1]
            qr = (np.float32(15.0) * sf[k_q["octave"]]).astype(np.float32)
            assert np.array_equal(qb["qx"][f, :n[s]].cpu().numpy(), qx)
            assert np.array_equal(qb["qr"][f, :n[s]].cpu().numpy(), qr)
            off, idx = oracle_mod.grid_queries(k_ref, bounds, qx, qy, qr, k_q["octave"] - 1, k_q["octave"] + 1)
            total_cand += len(idx)
            ref = oracle_mod.argmin2_csr(desc[s, :n[s]], desc[f, :n[f]], off, idx, k_ref["octave"].astype(np.int8), tie)
            _same_results(res[f, :n[s]], ref, with_second=(tie == TIE_FIRST_MIN))
            # and the explicit-CSR kernel on the same lists
            got2 = results_to_numpy(ORBmatcher.argmin2_csr(_t(desc[s, :n[s]]), _t(desc[f, :n[f]]), _t(off), _t(idx),
                                                           _t(k_ref["octave"].astype(np.int8)), tie))
            _same_results(got2, ref, with_second=(tie == TIE_FIRST_MIN))
        assert total_cand > 3 * B * 900  # the windows really contain candidates
        good = (res[1, :n[0]]["best_dist"] <= 100).mean()
        assert good > 0.7  # consecutive synthetic frames do match


def test_grid_build_is_the_reference_grid(oracle_mod):
    """CSR cells == mGrid[ix][iy] push_back order: queries that cover one whole cell column."""
    B = 2
    frames = S.make_frames("C3", B)
    ex = ORBextractor(1000, max_batch=B, **ORB)
    out = _extract_batch(ex, frames, (0, 0))
    bounds = (0.0, 640.0, 0.0, 480.0)
    cell_off, cell_idx = ORBmatcher.grid_build(out["kps"], out["n"], bounds)
    kps = ORBextractor.kps_from_tensor(out["kps"])
    n = out["n"].cpu().numpy()
    co, ci = cell_off.cpu().numpy(), cell_idx.cpu().numpy()
    for f in range(B):
        k = kps[f, :n[f]]
        import ctypes
        libm = ctypes.CDLL("libm.so.6")
        libm.roundf.restype = ctypes.c_float
        libm.roundf.argtypes = [ctypes.c_float]
        rnd = np.vectorize(lambda v: int(libm.roundf(float(v))))  # PosInGrid uses C round()
        px = rnd((k["x"] - np.float32(0)) * (np.float32(64) / np.float32(640)))
        py = rnd((k["y"] - np.float32(0)) * (np.float32(48) / np.float32(480)))
        ok = (px >= 0) & (px < 64) & (py >= 0) & (py < 48)
        cell = px * 48 + py
        assert co[f, -1] == ok.sum()
        for c in np.unique(cell[ok])[:400]:
            members = np.nonzero(ok & (cell == c))[0]
            assert np.array_equal(ci[f, co[f, c]:co[f, c + 1]], members)


def test_stereo_rowband(oracle_mod):
    """C2 shape: left/right extraction, then the Hamming stage of ComputeStereoMatches."""
    B = 2
    left = S.make_frames("C2", B)
    right = S.make_frames("C2", B, right=True)
    exl = ORBextractor(1200, max_batch=B, **ORB)
    exr = ORBextractor(1200, max_batch=B, **ORB)
    ol = _extract_batch(exl, left, (0, 0))
    orr = _extract_batch(exr, right, (0, 0))
    sf, isf = exl.GetScaleFactors(), exl.GetInverseScaleFactors()
    fx, bf = 458.654, 47.90639384423901  # Stereo-Inertial/EuRoC.yaml: maxD = mbf/mb = fx
    mb = bf / fx
    bi, bd = ORBmatcher.stereo_rowband(ol["kps"], ol["desc"], ol["n"], orr["kps"], orr["desc"], orr["n"], sf,
                                       np.float32(bf) / np.float32(mb))
    kl, kr = ORBextractor.kps_from_tensor(ol["kps"]), ORBextractor.kps_from_tensor(orr["kps"])
    dl, dr = ol["desc"].cpu().numpy(), orr["desc"].cpu().numpy()
    nl, nr = ol["n"].cpu().numpy(), orr["n"].cpu().numpy()
    for f in range(B):
        pl = [exl.GetPyramid(l, frame=f) for l in range(8)]
        pr = [exr.GetPyramid(l, frame=f) for l in range(8)]
        rbi, rbd, ur, dp, kept = oracle_mod.stereo_matches(kl[f, :nl[f]], dl[f, :nl[f]], kr[f, :nr[f]], dr[f, :nr[f]],
                                                           sf, isf, pl, pr, np.float32(mb), np.float32(bf))
        assert np.array_equal(bi[f, :nl[f]].cpu().numpy(), rbi)
        assert np.array_equal(bd[f, :nl[f]].cpu().numpy(), rbd)
        assert (rbi >= 0).sum() > 300 and kept > 100
        # constant 24 px disparity in the synthetic pair
        ok = ur > 0
        assert np.median(kl[f, :nl[f]]["x"][ok] - ur[ok]) == pytest.approx(24.0, abs=0.5)


def _stereo_case(oracle_mod, left, right, nfeat=1200, fx=458.654, bf=47.90639384423901):
    """Full ComputeStereoMatches on the device (rowband -> refine -> median cut) vs the oracle, bit for bit."""
    B = left.shape[0]
    exl = ORBextractor(nfeat, max_batch=B, **ORB)
    exr = ORBextractor(nfeat, max_batch=B, **ORB)
    ol = _extract_batch(exl, left, (0, 0))
    orr = _extract_batch(exr, right, (0, 0))
    sf, isf = exl.GetScaleFactors(), exl.GetInverseScaleFactors()
    mb = np.float32(bf / fx)
    mbf = np.float32(bf)
    max_d = mbf / mb
    bi, bd = ORBmatcher.stereo_rowband(ol["kps"], ol["desc"], ol["n"], orr["kps"], orr["desc"], orr["n"], sf, max_d)
    res = ORBmatcher.stereo_refine(exl, exr, ol["kps"], ol["n"], orr["kps"], bi, bd, mbf, max_d)
    kl, kr = ORBextractor.kps_from_tensor(ol["kps"]), ORBextractor.kps_from_tensor(orr["kps"])
    dl, dr = ol["desc"].cpu().numpy(), orr["desc"].cpu().numpy()
    nl, nr = ol["n"].cpu().numpy(), orr["n"].cpu().numpy()
    ur_g, dp_g, kept_g = res["u_right"].cpu().numpy(), res["depth"].cpu().numpy(), res["kept"].cpu().numpy()
    out = []
    for f in range(B):
        pl = [exl.GetPyramid(l, frame=f) for l in range(8)]
        pr = [exr.GetPyramid(l, frame=f) for l in range(8)]
        rbi, rbd, ur, dp, kept = oracle_mod.stereo_matches(kl[f, :nl[f]], dl[f, :nl[f]], kr[f, :nr[f]], dr[f, :nr[f]],
                                                           sf, isf, pl, pr, mb, mbf)
        assert np.array_equal(ur_g[f, :nl[f]].view(np.uint32), ur.view(np.uint32))
        assert np.array_equal(dp_g[f, :nl[f]].view(np.uint32), dp.view(np.uint32))
        assert kept_g[f] == kept == int((ur > 0).sum())
        out.append((kl[f, :nl[f]], ur, dp, kept))
    return out


def test_stereo_refine_constant_disparity(oracle_mod):
    left = S.make_frames("C2", 3)
    right = S.make_frames("C2", 3, right=True)
    for kl, ur, dp, kept in _stereo_case(oracle_mod, left, right):
        ok = ur > 0
        assert kept > 100
        d = kl["x"][ok] - ur[ok]
        assert np.median(d) == pytest.approx(24.0, abs=0.5)
        assert len(np.unique(ur[ok] - np.floor(ur[ok]))) > 20  # the parabola really moves things off the grid


def test_stereo_c4_kitti_shape(oracle_mod):
    """BASELINE config C4: 1241x376 pairs, K=2000/eye, maxD = mbf/mb = fx = 718.856 (Stereo/KITTI00-02.yaml:
    Camera1.fx 718.856, Stereo.b 0.53716): both eyes + the whole ComputeStereoMatches against the oracle."""
    left = S.make_frames("C4", 2)
    right = S.make_frames("C4", 2, right=True)
    fx = 718.856
    for kl, ur, dp, kept in _stereo_case(oracle_mod, left, right, nfeat=2000, fx=fx, bf=fx * 0.53716):
        ok = ur > 0
        assert len(kl) >= 2000 and kept > 300
        assert np.median(kl["x"][ok] - ur[ok]) == pytest.approx(24.0, abs=0.5)


def test_stereo_refine_zero_and_mixed_disparity(oracle_mod):
    """Edge cases of Frame.cc:951-980: identical eyes (median SAD 0 -> every match is cut), the same view with
    independent noise (disparity ~0: negative ones rejected, exact 0 clamped to 0.01 with uL-0.01 in double),
    and a (2,1) px offset pair."""
    left = np.stack([S.make_frames("C2", 1, t0=0)[0], S.make_frames("C2", 1, t0=2)[0],
                     S.make_frames("C2", 1, t0=100)[0]])
    right = np.stack([left[0], S.make_frames("C2", 1, t0=34)[0],   # t and t+32 share the crop, not the noise
                      S.make_frames("C2", 1, t0=101)[0]])
    res = _stereo_case(oracle_mod, left, right)
    assert res[0][3] == 0
    kl, ur, dp, kept = res[1]
    assert kept > 200
    clamped = ur == (kl["x"].astype(np.float64) - 0.01).astype(np.float32)
    assert (clamped & (dp == np.float32(47.90639384423901) / np.float32(0.01))).sum() >= 3
    assert res[2][3] > 500


def test_rgbd_depth(oracle_mod):
    import torch
    B = 2
    frames = S.make_frames("C3", B)
    ex = ORBextractor(1000, max_batch=B, **ORB)
    o = _extract_batch(ex, frames, (0, 0))
    rng = np.random.default_rng(5)
    depth = rng.uniform(-0.5, 6.0, (B, 480, 640)).astype(np.float32)
    depth[rng.random(depth.shape) < 0.2] = 0.0
    mbf = np.float32(40.0)
    d_t = torch.from_numpy(depth).cuda()
    ur, dp = ORBmatcher.rgbd_depth(o["kps"], o["n"], d_t, mbf)
    k = ORBextractor.kps_from_tensor(o["kps"])
    n = o["n"].cpu().numpy()
    for f in range(B):
        rur, rdp = oracle_mod.rgbd_depth(k[f, :n[f]], k[f, :n[f]], depth[f], mbf)
        assert np.array_equal(ur[f, :n[f]].cpu().numpy().view(np.uint32), rur.view(np.uint32))
        assert np.array_equal(dp[f, :n[f]].cpu().numpy().view(np.uint32), rdp.view(np.uint32))
        assert (rdp > 0).sum() > 300 and (rdp < 0).sum() > 50
    # pitched view + separate undistorted keypoints
    big = torch.zeros((B, 480, 704), dtype=torch.float32, device="cuda")
    big[:, :, :640] = d_t
    kun = o["kps"].clone()
    kun[:, :, 0] += 0.25
    kun_np = ORBextractor.kps_from_tensor(kun)
    ur2, dp2 = ORBmatcher.rgbd_depth(o["kps"], o["n"], big[:, :, :640], mbf, kps_un=kun)
    for f in range(B):
        rur, rdp = oracle_mod.rgbd_depth(k[f, :n[f]], kun_np[f, :n[f]], depth[f], mbf)
        assert np.array_equal(ur2[f, :n[f]].cpu().numpy().view(np.uint32), rur.view(np.uint32))
        assert np.array_equal(dp2[f, :n[f]].cpu().numpy().view(np.uint32), rdp.view(np.uint32))


def _proj_queries(kq, dq, shift, th, sf, band, rng, dup=1, drop=0.0, unlocked=0.0, cap=None):
    """Host-built SearchByProjection queries: keypoints kq of another frame moved by `shift`."""
    k = np.concatenate([kq] * dup)
    d = np.concatenate([dq] * dup)
    nq = len(k)
    q = {"qx": (k["x"] + np.float32(shift[0])).astype(np.float32), "qy": (k["y"] + np.float32(shift[1])).astype(np.float32),
         "qr": (np.float32(th) * sf[k["octave"]]).astype(np.float32),
         "qmin": (k["octave"] + band[0]).astype(np.int8), "qmax": (k["octave"] + band[1]).astype(np.int8),
         "qdesc": d, "angle": k["angle"].astype(np.float32)}
    fl = np.full(nq, 3, np.uint8)
    fl[rng.random(nq) < unlocked] = 1
    fl[rng.random(nq) < drop] &= 2
    q["flags"] = fl
    q["q_ur"] = (q["qx"] - np.float32(24.0) + rng.normal(0, 3, nq)).astype(np.float32)
    return q


def _proj_run(oracle_mod, out, q_list, bounds, mode, nn_ratio, check_ori, u_right=None, feat_locked=None, th_dist=100):
    import torch
    B, cap = out["kps"].shape[0], out["kps"].shape[1]
    Q = max(len(q["qx"]) for q in q_list)
    dev = out["kps"].device
    qt = {"qx": torch.zeros((B, Q), dtype=torch.float32), "qy": torch.zeros((B, Q), dtype=torch.float32),
          "qr": torch.zeros((B, Q), dtype=torch.float32), "qmin": torch.zeros((B, Q), dtype=torch.int8),
          "qmax": torch.zeros((B, Q), dtype=torch.int8), "qdesc": torch.zeros((B, Q, 32), dtype=torch.uint8),
          "angle": torch.zeros((B, Q), dtype=torch.float32), "flags": torch.zeros((B, Q), dtype=torch.uint8),
          "q_ur": torch.zeros((B, Q), dtype=torch.float32), "nq": torch.zeros(B, dtype=torch.int32)}
    for f, q in enumerate(q_list):
        m = len(q["qx"])
        for key in ("qx", "qy", "qr", "qmin", "qmax", "qdesc", "angle", "flags", "q_ur"):
            qt[key][f, :m] = torch.from_numpy(np.ascontiguousarray(q[key]))
        qt["nq"][f] = m
    qt = {k_: v.to(dev) for k_, v in qt.items()}
    if u_right is None:
        qt.pop("q_ur")
    grid = ORBmatcher.grid_build(out["kps"], out["n"], bounds)
    res = ORBmatcher.search_by_projection(out["kps"], out["desc"], out["n"], grid, bounds, qt, mode, nn_ratio, check_ori,
                                          u_right=None if u_right is None else _t(u_right),
                                          feat_locked=None if feat_locked is None else _t(feat_locked), th_dist=th_dist)
    kps = ORBextractor.kps_from_tensor(out["kps"])
    desc = out["desc"].cpu().numpy()
    n = out["n"].cpu().numpy()
    mt, fq, nm = res["match"].cpu().numpy(), res["feat_query"].cpu().numpy(), res["nmatches"].cpu().numpy()
    tot = 0
    for f, q in enumerate(q_list):
        rm, rfq, rnm = oracle_mod.search_by_projection(
            kps[f, :n[f]], desc[f, :n[f]], bounds, q["qx"], q["qy"], q["qr"], q["qmin"], q["qmax"], q["qdesc"], mode,
            nn_ratio, check_ori, u_right=None if u_right is None else u_right[f, :n[f]],
            feat_locked=None if feat_locked is None else feat_locked[f, :n[f]],
            q_ur=q["q_ur"] if u_right is not None else None, q_flags=q["flags"], q_angle=q["angle"], th_dist=th_dist)
        m = len(q["qx"])
        assert np.array_equal(mt[f, :m], rm), (f, np.nonzero(mt[f, :m] != rm)[0][:10])
        assert np.array_equal(fq[f, :n[f]], rfq)
        assert nm[f] == rnm
        tot += rnm
    return tot, mt, fq


@pytest.mark.parametrize("orb_dist", [100, 64, 30])
def test_search_by_projection_keyframe_relocalisation(oracle_mod, orb_dist):
    """SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist) (ORBmatcher.cc:1889-2010): keyframe
    MapPoints projected into the current frame, band [l-1,l+1], every non-null mvpMapPoints entry blocks its
    feature, best <= ORBdist, rotation histogram.  th = 10 / 3 as Tracking::Relocalization uses them."""
    from bench import shifts_for
    B = 3
    frames = S.make_frames("C1", B)
    ex = ORBextractor(1000, max_batch=B, **ORB)
    out = _extract_batch(ex, frames, (0, 1000))
    kps = ORBextractor.kps_from_tensor(out["kps"])
    desc = out["desc"].cpu().numpy()
    n = out["n"].cpu().numpy()
    sf = ex.GetScaleFactors()
    sh = shifts_for(0, B)
    rng = np.random.default_rng(13)
    bounds = (0.0, 752.0, 0.0, 480.0)
    cap = out["kps"].shape[1]
    for th in (10, 3):
        ql = []
        for f in range(B):
            s = (f - 1) % B
            # dup=2: the second copy of every MapPoint finds its best feature taken; drop = in sAlreadyFound / bad
            q = _proj_queries(kps[s, :n[s]], desc[s, :n[s]], sh[f], th, sf, (-1, 1), rng, dup=2, drop=0.2)
            q["flags"] = (q["flags"] | 2).astype(np.uint8) & np.where(q["flags"] & 1, 3, 2).astype(np.uint8)
            ql.append(q)
        locked = (rng.random((B, cap)) < 0.25).astype(np.uint8)   # features already holding a MapPoint
        tot, mt, fq = _proj_run(oracle_mod, out, ql, bounds, 1, 0.9, True, feat_locked=locked, th_dist=orb_dist)
        assert tot > (100 if orb_dist == 30 else 600)
        for f in range(B):  # nothing is ever written to a feature that was occupied on entry
            assert np.all(fq[f, :n[f]][locked[f, :n[f]] != 0] == -1)


@pytest.mark.parametrize("ratio_hamming", [1.0, 0.75])
def test_search_by_projection_sim3_keyframe(oracle_mod, ratio_hamming):
    """SearchByProjection(KeyFrame*, Sim3, vpPoints, vpMatched, th, ratioHamming) (ORBmatcher.cc:427-532, 534-646):
    candidates on levels [l-1, l] only, features already in vpMatched are skipped, best <= TH_LOW*ratioHamming,
    no rotation check.  Served by the threshold-parameter entry point."""
    from bench import shifts_for
    B = 3
    frames = S.make_frames("C1", B)
    ex = ORBextractor(1000, max_batch=B, **ORB)
    out = _extract_batch(ex, frames, (0, 1000))
    kps = ORBextractor.kps_from_tensor(out["kps"])
    desc = out["desc"].cpu().numpy()
    n = out["n"].cpu().numpy()
    sf = ex.GetScaleFactors()
    sh = shifts_for(0, B)
    rng = np.random.default_rng(17)
    bounds = (0.0, 752.0, 0.0, 480.0)
    cap = out["kps"].shape[1]
    ql = []
    for f in range(B):
        s = (f - 1) % B
        ql.append(_proj_queries(kps[s, :n[s]], desc[s, :n[s]], sh[f], 4, sf, (-1, 0), rng, dup=2, drop=0.15))
    matched = (rng.random((B, cap)) < 0.3).astype(np.uint8)     # vpMatched[idx] != NULL on entry
    th_dist = int(np.floor(50 * ratio_hamming))                 # bestDist <= TH_LOW*ratioHamming, integer distances
    tot, mt, fq = _proj_run(oracle_mod, out, ql, bounds, 1, 0.9, False, feat_locked=matched, th_dist=th_dist)
    assert tot > 300


@pytest.mark.parametrize("check_ori", [True, False])
def test_search_by_projection_lastframe(oracle_mod, check_ori):
    """ORBmatcher.cc:1676-1887 (mono shape): last-frame keypoints projected by the known shift, th=15."""
    from bench import shifts_for, MATCH_TH
    B = 4
    frames = S.make_frames("C1", B)
    ex = ORBextractor(1000, max_batch=B, **ORB)
    out = _extract_batch(ex, frames, (0, 1000))
    kps = ORBextractor.kps_from_tensor(out["kps"])
    desc = out["desc"].cpu().numpy()
    n = out["n"].cpu().numpy()
    sf = ex.GetScaleFactors()
    sh = shifts_for(0, B)
    rng = np.random.default_rng(3)
    bounds = (0.0, 752.0, 0.0, 480.0)
    # plain; duplicated queries (every second copy finds its best locked -> rescans); mixed flags
    for dup, drop, unl in ((1, 0.0, 0.0), (2, 0.0, 0.0), (3, 0.1, 0.4)):
        ql = []
        for f in range(B):
            s = (f - 1) % B
            ql.append(_proj_queries(kps[s, :n[s]], desc[s, :n[s]], sh[f], MATCH_TH, sf, (-1, 1), rng, dup, drop, unl))
        tot, mt, fq = _proj_run(oracle_mod, out, ql, bounds, 1, 0.9, check_ori)
        assert tot > B * 500 * (1 if dup == 1 else 1.2)


def test_search_by_projection_mappoints(oracle_mod):
    """ORBmatcher.cc:43-142 (C3 shape): 'map points' = keypoints of the 3 other frames, levels [l-1,l], ratio 0.8,
    some features already matched on entry, stereo consistency against mvuRight."""
    from bench import shifts_for
    B = 4
    frames = S.make_frames("C3", B)
    ex = ORBextractor(1000, max_batch=B, **ORB)
    out = _extract_batch(ex, frames, (0, 0))
    kps = ORBextractor.kps_from_tensor(out["kps"])
    desc = out["desc"].cpu().numpy()
    n = out["n"].cpu().numpy()
    cap = out["kps"].shape[1]
    sf = ex.GetScaleFactors()
    rng = np.random.default_rng(9)
    bounds = (0.0, 640.0, 0.0, 480.0)
    pos = np.array([S.frame_shift(t) for t in range(B)], np.float32)
    ql = []
    for f in range(B):
        parts = []
        for s in range(B):
            if s == f:
                continue
            parts.append(_proj_queries(kps[s, :n[s]], desc[s, :n[s]], pos[s] - pos[f], 12.0, sf, (-1, 0), rng, 1, 0.05, 0.2))
        ql.append({k_: np.concatenate([p[k_] for p in parts]) for k_ in parts[0]})
    locked = (rng.random((B, cap)) < 0.15).astype(np.uint8)
    u_right = np.where(rng.random((B, cap)) < 0.7, kps["x"] - np.float32(24.0), np.float32(-1)).astype(np.float32)
    for nn in (0.8, 0.6):
        tot, mt, fq = _proj_run(oracle_mod, out, ql, bounds, 0, nn, False, u_right=u_right, feat_locked=locked)
        assert tot > B * 300
    tot2, _, _ = _proj_run(oracle_mod, out, ql, bounds, 0, 0.8, False)
    assert tot2 > tot


def test_search_by_projection_c3_survey_shape(oracle_mod):
    """SURVEY §8(d) C3 exactly: 640x480, K=1000; the 'local map' of frame t is the keypoints + descriptors of frames
    t-4..t-1 (M ~ 4000), projected by the known shifts; th=3, r=4.0 -> radius 12*scale[l]; levels [l-1,l]; best + second
    with levels; nnratio 0.8 (Tracking.cc:3390-3393 -> ORBmatcher.cc:43-142)."""
    B = 6
    frames = S.make_frames("C3", B)
    ex = ORBextractor(1000, max_batch=B, **ORB)
    out = _extract_batch(ex, frames, (0, 0))
    kps = ORBextractor.kps_from_tensor(out["kps"])
    desc = out["desc"].cpu().numpy()
    n = out["n"].cpu().numpy()
    sf = ex.GetScaleFactors()
    rng = np.random.default_rng(19)
    bounds = (0.0, 640.0, 0.0, 480.0)
    pos = np.array([S.frame_shift(t) for t in range(B)], np.float32)
    ql = []
    for f in range(B):
        parts = [_proj_queries(kps[s % B, :n[s % B]], desc[s % B, :n[s % B]], pos[s % B] - pos[f], 4.0 * 3.0, sf, (-1, 0), rng,
                               1, 0.0, 0.0) for s in range(f - 4, f)]
        ql.append({k_: np.concatenate([p[k_] for p in parts]) for k_ in parts[0]})
        assert 3900 < len(ql[-1]["qx"]) < 4200
    tot, mt, fq = _proj_run(oracle_mod, out, ql, bounds, 0, 0.8, False)
    assert tot > B * 500


def test_search_area_c5_frame_to_frame(oracle_mod):
    """BASELINE config C5 matcher leg: 3840x2160, K=5000, frame t-1 keypoints shifted into frame t, window
    15*scale[octave], octave band [o-1,o+1], 64x48 grid over the 4K bounds -> fused GetFeaturesInArea + scoring."""
    from bench import shifts_for
    B = 2
    frames = S.make_frames("C5", B)
    ex = ORBextractor(5000, max_batch=B, **ORB)
    out = _extract_batch(ex, frames, (0, 1000))
    sf = ex.GetScaleFactors()
    bounds = (0.0, 3840.0, 0.0, 2160.0)
    shift = shifts_for(0, B)
    src = ((np.arange(B) - 1) % B).astype(np.int32)
    grid = ORBmatcher.grid_build(out["kps"], out["n"], bounds)
    qb = ORBmatcher.queries_shift(out["kps"], out["desc"], out["n"], _t(src), _t(shift), 15.0, sf, (-1, 1))
    kps = ORBextractor.kps_from_tensor(out["kps"])
    desc = out["desc"].cpu().numpy()
    n = out["n"].cpu().numpy()
    res = results_to_numpy(ORBmatcher.search_area(out["kps"], out["desc"], out["n"], grid, bounds, qb["qx"], qb["qy"],
                                                  qb["qr"], qb["qmin"], qb["qmax"], qb["qdesc"], qb["nq"], TIE_FIRST_MIN))
    for f in range(B):
        s_ = src[f]
        k_ref, k_q = kps[f, :n[f]], kps[s_, :n[s_]]
        assert n[f] >= 5000
        qx, qy = k_q["x"] + shift[f, 0], k_q["y"] + shift[f, 1]
        qr = (np.float32(15.0) * sf[k_q["octave"]]).astype(np.float32)
        off, idx = oracle_mod.grid_queries(k_ref, bounds, qx, qy, qr, k_q["octave"] - 1, k_q["octave"] + 1)
        ref = oracle_mod.argmin2_csr(desc[s_, :n[s_]], desc[f, :n[f]], off, idx, k_ref["octave"].astype(np.int8), TIE_FIRST_MIN)
        _same_results(res[f, :n[s_]], ref)
    assert (res[1, :n[0]]["best_dist"] <= 100).mean() > 0.6


def test_degenerate_inputs_new_entry_points(oracle_mod):
    """Flat frames (0 keypoints) and empty query sets through the stereo / projection / RGB-D entry points."""
    import torch
    from orb_slam3_vio_fixes_b200 import _lib
    B = 2
    flat = np.full((B, 480, 752), 90, np.uint8)
    mixed = np.stack([S.make_frames("C2", 1)[0], flat[0]])
    exl = ORBextractor(1200, max_batch=B, **ORB)
    exr = ORBextractor(1200, max_batch=B, **ORB)
    ol = _extract_batch(exl, mixed, (0, 0))
    orr = _extract_batch(exr, flat, (0, 0))
    assert int(ol["n"][1]) == 0 and int(orr["n"].sum()) == 0
    sf = exl.GetScaleFactors()
    bi, bd = ORBmatcher.stereo_rowband(ol["kps"], ol["desc"], ol["n"], orr["kps"], orr["desc"], orr["n"], sf, 458.0)
    res = ORBmatcher.stereo_refine(exl, exr, ol["kps"], ol["n"], orr["kps"], bi, bd, 47.9, 458.0)
    n0 = int(ol["n"][0])
    assert (bi[0, :n0] == -1).all() and (res["u_right"][0, :n0] == -1).all() and res["kept"].tolist() == [0, 0]
    # projection search: no queries, then queries against a frame without features
    bounds = (0.0, 752.0, 0.0, 480.0)
    grid = ORBmatcher.grid_build(ol["kps"], ol["n"], bounds)
    Q = 64
    dev = ol["kps"].device
    q = {"qx": torch.full((B, Q), 300.0, device=dev), "qy": torch.full((B, Q), 200.0, device=dev),
         "qr": torch.full((B, Q), 20.0, device=dev), "qmin": torch.zeros((B, Q), dtype=torch.int8, device=dev),
         "qmax": torch.full((B, Q), -1, dtype=torch.int8, device=dev),
         "qdesc": torch.zeros((B, Q, 32), dtype=torch.uint8, device=dev),
         "angle": torch.zeros((B, Q), device=dev), "nq": torch.tensor([0, Q], dtype=torch.int32, device=dev)}
    for mode in (_lib.PROJ_MAPPOINTS, _lib.PROJ_LASTFRAME):
        r = ORBmatcher.search_by_projection(ol["kps"], ol["desc"], ol["n"], grid, bounds, q, mode, 0.8, True)
        assert r["nmatches"].tolist() == [0, 0] and (r["match"][1] == -1).all() and (r["feat_query"][0, :n0] == -1).all()
    ur, dp = ORBmatcher.rgbd_depth(ol["kps"], ol["n"], torch.ones((B, 480, 752), device=dev), 40.0)
    assert (dp[0, :n0] == 1).all()
