"""SearchForTriangulation on a two-camera rig (ORBmatcher.cc:907-1146, camera selection :1036-1062) and its
geometric gate KannalaBrandt8::epipolarConstrain (KannalaBrandt8.cpp:216-220, 306-375).

The gate is floating-point triangulation (fp32 Eigen JacobiSVD + libm in the reference): tolerance contract.  The
product's gate (csrc/tri_core.h: fp32 rays, fp64 Jacobi null vector) must agree with an fp64 numpy evaluation for
every pair that is not within 1e-3 (relative) of a threshold; the loop logic around it is bit-exact against the
oracle, which takes the gate's outcome as an input."""
import ctypes as C

import numpy as np
import pytest

import rig_geometry as G
from conftest import has_cuda

AMBIGUOUS = 1e-3


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def scene(rng, n_pts, noise_px=0.7):
    """3-D points in front of both keyframes of the rig, observed by the four cameras with pixel noise."""
    P = np.stack([rng.uniform(-3, 3, n_pts), rng.uniform(-2, 2, n_pts), rng.uniform(1.0, 9, n_pts)], 1)
    K1, K2 = G.rig_poses(rng)
    cams = (G.CAM_L, G.CAM_R)
    obs = {}
    for name, Ts in (("1", K1), ("2", K2)):
        for side in (0, 1):
            Xc = P @ Ts[side][0].T + Ts[side][1]
            obs[name + "lr"[side]] = (G.kb8_project(cams[side], Xc) + rng.normal(0, noise_px, (n_pts, 2))).astype(np.float32)
    return P, K1, K2, obs


def test_gate_matches_fp64_outside_the_tolerance_band(harness):
    rng = np.random.default_rng(71)
    n = 40000
    P, K1, K2, obs = scene(rng, n, noise_px=1.2)
    sig = np.cumprod([1.0] + [1.2] * 7).astype(np.float32) ** 2
    s1, s2 = sig[rng.integers(0, 8, n)], sig[rng.integers(0, 8, n)]
    harness.h_kb8_epipolar.argtypes = [C.c_void_p] * 4 + [C.c_int] + [C.c_void_p] * 4
    tot_bad = tot_amb = tot_ok = 0
    for a, ca, Ta in (("1l", G.CAM_L, K1[0]), ("1r", G.CAM_R, K1[1])):
        for b, cb, Tb in (("2l", G.CAM_L, K2[0]), ("2r", G.CAM_R, K2[1])):
            uv1, uv2 = obs[a].copy(), obs[b].copy()
            uv2[n // 2:] = uv2[n // 2:][rng.permutation(n - n // 2)]      # half of the pairs are wrong correspondences
            R12, t12 = G.rel_pose(Ta, Tb)
            pose = np.concatenate([R12.reshape(-1), t12]).astype(np.float32)
            R32, t32 = pose[:9].reshape(3, 3).astype(np.float64), pose[9:].astype(np.float64)
            ok, margin = G.epipolar_gate64(ca, cb, uv1.astype(np.float64), uv2.astype(np.float64), R32, t32, s1, s2)
            out = np.zeros(n, np.uint8)
            c1 = np.array(ca[1:], np.float32); c2 = np.array(cb[1:], np.float32)
            harness.h_kb8_epipolar(_p(c1), _p(c2), _p(uv1), _p(uv2), n, _p(pose), _p(s1), _p(s2), _p(out))
            diff = out.astype(bool) != ok
            assert not (diff & (margin >= AMBIGUOUS)).any(), (a, b, margin[diff])
            tot_bad += int(diff.sum()); tot_amb += int((margin < AMBIGUOUS).sum()); tot_ok += int(ok.sum())
    print(f"KB8 epipolarConstrain fp32/fp64: {tot_bad} of {4 * n} decisions differ, all inside the band "
          f"({tot_amb} pairs are within {AMBIGUOUS} of a threshold); {tot_ok} pairs pass")
    assert tot_ok > n and tot_bad <= 4 * n // 1000


@pytest.mark.gpu
@pytest.mark.skipif(not has_cuda(), reason="needs CUDA")
@pytest.mark.parametrize("coarse", [False, True])
def test_gpu_search_for_triangulation_rig(oracle_mod, coarse):
    import torch
    from orb_slam3_vio_fixes_b200 import frame as F
    from orb_slam3_vio_fixes_b200 import synthetic as S
    from orb_slam3_vio_fixes_b200.matcher import ORBmatcher
    rng = np.random.default_rng(73)
    n_pts, twins = 700, 3
    # `twins` 3-D points share one texture (descriptor + a few bit flips): Hamming passes, geometry must decide
    base = rng.integers(0, 256, (n_pts // twins + 1, 32), dtype=np.uint8)
    P, K1, K2, obs = scene(rng, n_pts)
    tex = np.arange(n_pts) // twins

    def side_features(tag):
        """left|right concatenation of one keyframe; every camera sees a random 80 % of the points."""
        k_all, d_all, pid = [], [], []
        for side in "lr":
            seen = np.nonzero(rng.random(n_pts) < 0.8)[0]
            uv = obs[tag + side][seen]
            inimg = (uv[:, 0] > 5) & (uv[:, 0] < 507) & (uv[:, 1] > 5) & (uv[:, 1] < 507)
            seen, uv = seen[inimg], uv[inimg]
            k = np.zeros(len(seen), oracle_mod.KP_DTYPE)
            k["x"], k["y"] = uv[:, 0], uv[:, 1]
            k["size"], k["response"], k["class_id"] = 31, 50, -1
            k["octave"] = rng.integers(0, 4, len(seen))
            k["angle"] = (tex[seen] * 7.0 + rng.normal(0, 3, len(seen))) % 360
            flips = (rng.random((len(seen), 256)) < 0.03)
            d = base[tex[seen]] ^ np.packbits(flips, axis=1, bitorder="little")
            k_all.append(k); d_all.append(d); pid.append(seen)
        return np.concatenate(k_all), np.concatenate(d_all), len(k_all[0]), np.concatenate(pid)

    k1, d1, nl1, pid1 = side_features("1")
    k2, d2, nl2, pid2 = side_features("2")
    voc = S.make_vocabulary(6, 3, seed=4, desc_pool=base)
    fv1 = oracle_mod.bow_transform(d1, voc, 2)["fv"]
    fv2 = oracle_mod.bow_transform(d2, voc, 2)["fv"]
    sig = (np.cumprod([1.0] + [1.2] * 7).astype(np.float32) ** 2).astype(np.float32)
    # the gate for every (idx1, idx2) in fp64, per camera combination
    n1, n2 = len(k1), len(k2)
    gate = np.zeros((n1, n2), np.uint8)
    margin = np.full((n1, n2), np.inf)
    poses = np.zeros((4, 12), np.float32)
    cams = (G.CAM_L, G.CAM_R)
    for r1 in (0, 1):
        for r2 in (0, 1):
            R12, t12 = G.rel_pose(K1[r1], K2[r2])
            poses[2 * r1 + r2] = np.concatenate([R12.reshape(-1), t12])
            R32, t32 = poses[2 * r1 + r2, :9].reshape(3, 3).astype(np.float64), poses[2 * r1 + r2, 9:].astype(np.float64)
            i1 = np.arange(nl1) if r1 == 0 else np.arange(nl1, n1)
            i2 = np.arange(nl2) if r2 == 0 else np.arange(nl2, n2)
            a, b = np.meshgrid(i1, i2, indexing="ij")
            a, b = a.reshape(-1), b.reshape(-1)
            uv1 = np.stack([k1["x"][a], k1["y"][a]], 1).astype(np.float64)
            uv2 = np.stack([k2["x"][b], k2["y"][b]], 1).astype(np.float64)
            ok, m = G.epipolar_gate64(cams[r1], cams[r2], uv1, uv2, R32, t32, sig[k1["octave"][a]], sig[k2["octave"][b]])
            gate[a, b] = ok
            margin[a, b] = m
    has1 = (rng.random(n1) < 0.1).astype(np.uint8)
    has2 = (rng.random(n2) < 0.1).astype(np.uint8)
    ref12, ref_nm = oracle_mod.search_for_triangulation_rig(k1, d1, fv1, k2, d2, fv2, gate, coarse=coarse,
                                                            check_orientation=False, has_mp1=has1, has_mp2=has2)
    cap = max(n1, n2)
    ns = max(len(fv1[0]), len(fv2[0]))

    def fvt(fv):
        node = np.zeros((1, ns), np.int32); off = np.zeros((1, ns + 1), np.int32); idx = np.zeros((1, cap), np.int32)
        node[0, :len(fv[0])] = fv[0]; off[0, :len(fv[1])] = fv[1]; idx[0, :len(fv[2])] = fv[2]
        t = lambda a: torch.from_numpy(a).cuda()
        return {"node": t(node), "off": t(off), "idx": t(idx), "nn": t(np.array([len(fv[0])], np.int32))}

    def pad(a, shape, dt):
        o = np.zeros(shape, dt); o[0, :len(a)] = a
        return torch.from_numpy(o.view(np.uint8).reshape(1, cap, -1) if dt == k1.dtype else o).cuda()

    t32 = lambda v: torch.tensor([v], dtype=torch.int32).cuda()
    cl = F.camera(1, *G.CAM_L[1:5], G.CAM_L[5:9]); cr = F.camera(1, *G.CAM_R[1:5], G.CAM_R[5:9])
    res = ORBmatcher.search_for_triangulation_rig(
        pad(k1, (1, cap), k1.dtype), pad(d1, (1, cap, 32), np.uint8), t32(n1), t32(nl1), fvt(fv1),
        pad(k2, (1, cap), k2.dtype), pad(d2, (1, cap, 32), np.uint8), t32(n2), t32(nl2), fvt(fv2),
        torch.from_numpy(poses[None]).cuda(), (cl, cr), (cl, cr), sig, sig, coarse=coarse, check_orientation=False,
        has_mp1=pad(has1, (1, cap), np.uint8), has_mp2=pad(has2, (1, cap), np.uint8))
    got = res["match12"][0, :n1].cpu().numpy()
    # a query is "clear" when no Hamming-eligible candidate of its node sits inside the tolerance band of the gate
    amb_q = np.zeros(n1, bool)
    if not coarse:
        d1b, d2b = np.unpackbits(d1, axis=1), np.unpackbits(d2, axis=1)
        ham = (d1b[:, None, :] != d2b[None, :, :]).sum(-1)
        amb_q = ((ham <= 50) & (margin < AMBIGUOUS)).any(1)
    clear = ~amb_q
    assert np.array_equal(got[clear], ref12[clear]), np.nonzero(got != ref12)[0][:10]
    matched = ref12 >= 0
    assert matched.sum() > 200 and amb_q.sum() <= n1 // 50
    if not coarse:
        # the gate did discriminate: most accepted matches are the same 3-D point, which Hamming alone cannot tell
        same = pid1[matched] == pid2[ref12[matched]]
        assert same.mean() > 0.9
        assert int(res["nmatches"][0]) == int((got >= 0).sum())
    else:
        assert int(res["nmatches"][0]) == ref_nm
    # only_stereo: bStereo1/2 are false on a rig -> nothing is matched (:973-977)
    res = ORBmatcher.search_for_triangulation_rig(
        pad(k1, (1, cap), k1.dtype), pad(d1, (1, cap, 32), np.uint8), t32(n1), t32(nl1), fvt(fv1),
        pad(k2, (1, cap), k2.dtype), pad(d2, (1, cap, 32), np.uint8), t32(n2), t32(nl2), fvt(fv2),
        torch.from_numpy(poses[None]).cuda(), (cl, cr), (cl, cr), sig, sig, only_stereo=True, coarse=coarse)
    assert int(res["nmatches"][0]) == 0 and (res["match12"][0, :n1].cpu().numpy() == -1).all()


@pytest.mark.gpu
@pytest.mark.skipif(not has_cuda(), reason="needs CUDA")
def test_gpu_stereo_fisheye_matches(oracle_mod):
    """Frame::ComputeStereoFishEyeMatches (Frame.cc:1126-1166): knnMatch(k=2) over the lapping-area keypoints, ratio
    0.7, TriangulateMatches.  The index logic is exact; depth / p3D within 1e-3 relative of the fp64 evaluation."""
    import torch
    from orb_slam3_vio_fixes_b200 import frame as F
    from orb_slam3_vio_fixes_b200.matcher import ORBmatcher
    rng = np.random.default_rng(79)
    n_pts = 900
    P, K1, K2, obs = scene(rng, n_pts, noise_px=0.5)
    base = rng.integers(0, 256, (n_pts, 32), dtype=np.uint8)
    sig = (np.cumprod([1.0] + [1.2] * 7).astype(np.float32) ** 2).astype(np.float32)

    def cam_features(tag):
        seen = rng.permutation(np.nonzero(rng.random(n_pts) < 0.85)[0])
        uv = obs[tag][seen]
        k = np.zeros(len(seen), oracle_mod.KP_DTYPE)
        k["x"], k["y"], k["size"], k["response"], k["class_id"] = uv[:, 0], uv[:, 1], 31, 50, -1
        k["octave"] = rng.integers(0, 4, len(seen))
        flips = rng.random((len(seen), 256)) < 0.04
        return k, base[seen] ^ np.packbits(flips, axis=1, bitorder="little"), seen

    kl, dl, pl = cam_features("1l")
    kr, dr, pr = cam_features("1r")
    nl, nr = len(kl), len(kr)
    mono_l, mono_r = 130, 90          # features below these indices lie outside the lapping area
    Rlr, tlr = G.rel_pose(K1[0], K1[1])
    Rlr32, tlr32 = Rlr.astype(np.float32), tlr.astype(np.float32)
    # reference evaluation: Hamming knn (ties to the lower train index), ratio in double, gate in fp64
    qi, ti = G.fisheye_ratio_candidates(dl, dr, mono_l, mono_r)
    uv1 = np.stack([kl["x"][qi], kl["y"][qi]], 1).astype(np.float64)
    uv2 = np.stack([kr["x"][ti], kr["y"][ti]], 1).astype(np.float64)
    ok, margin = G.epipolar_gate64(G.CAM_L, G.CAM_R, uv1, uv2, Rlr32.astype(np.float64), tlr32.astype(np.float64),
                                   sig[kl["octave"][qi]], sig[kr["octave"][ti]])
    cap = max(nl, nr) + 7
    pad = lambda a, dt, tail=(): torch.from_numpy(np.concatenate([a, np.zeros((cap - len(a),) + tail, dt)])[None]).cuda()
    kt = lambda k: torch.from_numpy(np.concatenate([k, np.zeros(cap - len(k), k.dtype)]).view(np.uint8).reshape(1, cap, 28)).cuda()
    t32 = lambda v: torch.tensor([v], dtype=torch.int32).cuda()
    cl = F.camera(1, *G.CAM_L[1:5], G.CAM_L[5:9]); cr = F.camera(1, *G.CAM_R[1:5], G.CAM_R[5:9])
    res = ORBmatcher.stereo_fisheye_matches(kt(kl), pad(dl, np.uint8, (32,)), t32(nl), t32(mono_l), kt(kr),
                                            pad(dr, np.uint8, (32,)), t32(nr), t32(mono_r), cl, cr, Rlr32, tlr32, sig)
    l2r = res["left_to_right"][0, :nl].cpu().numpy()
    r2l = res["right_to_left"][0, :nr].cpu().numpy()
    depth = res["depth"][0, :nl].cpu().numpy()
    ref_l2r = np.full(nl, -1, np.int64)
    ref_l2r[qi[ok]] = ti[ok]
    amb = np.zeros(nl, bool)
    amb[qi[margin < AMBIGUOUS]] = True
    assert np.array_equal(l2r[~amb], ref_l2r[~amb])
    assert (l2r[:mono_l] == -1).all() and amb.sum() <= 5 and (ref_l2r >= 0).sum() > 120
    # mvRightToLeftMatch: the last left index written to a right feature
    ref_r2l = np.full(nr, -1, np.int64)
    for i in np.nonzero(l2r >= 0)[0]:
        ref_r2l[l2r[i]] = i
    assert np.array_equal(r2l, ref_r2l)
    # most matches are the same 3-D point, depth = z of that point in the left camera (noise 0.5 px)
    m = np.nonzero(l2r >= 0)[0]
    same = pl[m] == pr[l2r[m]]
    assert same.mean() > 0.97
    z_true = (P[pl[m]] @ K1[0][0].T + K1[0][1])[:, 2]
    assert np.median(np.abs(depth[m][same] - z_true[same]) / z_true[same]) < 0.1
    assert (depth[l2r < 0] == -1).all() and (depth[m] > 1e-4).all()
    p3d = res["p3d"][0, :nl].cpu().numpy()
    assert np.allclose(p3d[m, 2], depth[m], rtol=0, atol=0)
