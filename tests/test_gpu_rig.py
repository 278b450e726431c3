"""GPU parity (through the C ABI) of the two-camera rig (fisheye, F.Nleft != -1) matcher functions against the
oracle: SearchByProjection x2 (ORBmatcher.cc:43-213, :1676-1887 with their right-camera blocks) and
SearchByBoW(KeyFrame*, Frame&) (:295-323, :356-380).  The rig is synthetic: the left / right eyes of the C2
stereo frames are the two cameras; 'MapPoints' are keypoints of the other frames moved by the known shift."""
import numpy as np
import pytest

from conftest import has_cuda
from orb_slam3_vio_fixes_b200 import _lib
from orb_slam3_vio_fixes_b200 import synthetic as S

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not has_cuda(), reason="needs CUDA")]

ORB = dict(scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7)
BOUNDS = (0.0, 752.0, 0.0, 480.0)


def _t(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.fixture(scope="module")
def rig():
    import torch
    from orb_slam3_vio_fixes_b200.extractor import ORBextractor
    B = 4
    fl, fr = S.make_frames("C2", B), S.make_frames("C2", B, right=True)
    exl, exr = ORBextractor(1200, max_batch=B, **ORB), ORBextractor(1200, max_batch=B, **ORB)
    ol = exl.extract_batch_device(_t(fl), (0, 0))
    orr = exr.extract_batch_device(_t(fr), (0, 0))
    torch.cuda.synchronize()
    host = lambda o: (ORBextractor.kps_from_tensor(o["kps"]), o["desc"].cpu().numpy(), o["n"].cpu().numpy())
    return dict(B=B, ol=ol, orr=orr, L=host(ol), R=host(orr), sf=exl.GetScaleFactors(), keep=(exl, exr))


def _queries(rig, f, rng, th, band, drop_l=0.1, drop_r=0.15, unlocked=0.2, dup=1):
    """MapPoints of frame f: the left keypoints of the other frames, projected into both cameras."""
    kl, dl, nl = rig["L"]
    parts = []
    pos = np.array([S.frame_shift(t) for t in range(rig["B"])], np.float32)
    for s in range(rig["B"]):
        if s == f:
            continue
        k = np.concatenate([kl[s, :nl[s]]] * dup)
        d = np.concatenate([dl[s, :nl[s]]] * dup)
        sh = pos[s] - pos[f]
        n = len(k)
        q = dict(qx=(k["x"] + sh[0]).astype(np.float32), qy=(k["y"] + sh[1]).astype(np.float32),
                 qr=(np.float32(th) * rig["sf"][k["octave"]]).astype(np.float32),
                 qmin=(k["octave"] + band[0]).astype(np.int8), qmax=(k["octave"] + band[1]).astype(np.int8),
                 qdesc=d, angle=k["angle"].astype(np.float32))
        q["qx_r"] = (q["qx"] - np.float32(24.0) + rng.normal(0, 1.5, n)).astype(np.float32)  # the right eye sees it 24 px left
        q["qy_r"] = (q["qy"] + rng.normal(0, 1.0, n)).astype(np.float32)
        q["qr_r"] = q["qr"].copy()
        q["qmin_r"], q["qmax_r"] = q["qmin"].copy(), q["qmax"].copy()
        fl = np.full(n, 7, np.uint8)
        fl[rng.random(n) < unlocked] &= 5        # no observations
        fl[rng.random(n) < drop_l] &= 6          # not in view of the left camera
        fl[rng.random(n) < drop_r] &= 3          # not in view of the right camera
        q["flags"] = fl
        parts.append(q)
    return {k_: np.concatenate([p[k_] for p in parts]) for k_ in parts[0]}


def _pairs(nl, nr, rng, frac=0.4):
    """mvLeftToRightMatch / mvRightToLeftMatch: a random partial bijection left <-> right."""
    m = int(min(nl, nr) * frac)
    li, ri = rng.permutation(nl)[:m], rng.permutation(nr)[:m]
    l2r, r2l = np.full(nl, -1, np.int32), np.full(nr, -1, np.int32)
    l2r[li] = ri
    r2l[ri] = li
    return l2r, r2l


def _run(rig, oracle_mod, mode, qs, nn_ratio, check_ori, locked=None, pairs=None):
    import torch
    from orb_slam3_vio_fixes_b200.matcher import ORBmatcher
    B = rig["B"]
    ol, orr = rig["ol"], rig["orr"]
    cap = ol["kps"].shape[1]
    Q = max(len(q["qx"]) for q in qs)
    dt = dict(qx=np.float32, qy=np.float32, qr=np.float32, qmin=np.int8, qmax=np.int8, qx_r=np.float32, qy_r=np.float32,
              qr_r=np.float32, qmin_r=np.int8, qmax_r=np.int8, angle=np.float32, flags=np.uint8)
    qt = {k_: np.zeros((B, Q), v) for k_, v in dt.items()}
    qt["qdesc"] = np.zeros((B, Q, 32), np.uint8)
    qt["nq"] = np.zeros(B, np.int32)
    for f, q in enumerate(qs):
        m = len(q["qx"])
        for k_ in list(dt) + ["qdesc"]:
            qt[k_][f, :m] = q[k_]
        qt["nq"][f] = m
    qt = {k_: _t(v) for k_, v in qt.items()}
    ll = lr = pl = pr = None
    if locked is not None:
        ll, lr = _t(locked[0]), _t(locked[1])
    if pairs is not None:
        pl, pr = np.full((B, cap), -1, np.int32), np.full((B, cap), -1, np.int32)
        for f, (a, b) in enumerate(pairs):
            pl[f, :len(a)] = a
            pr[f, :len(b)] = b
        pl, pr = _t(pl), _t(pr)
    gl = ORBmatcher.grid_build(ol["kps"], ol["n"], BOUNDS)
    gr = ORBmatcher.grid_build(orr["kps"], orr["n"], BOUNDS)
    left = ORBmatcher.rig_camera(ol["kps"], ol["desc"], ol["n"], gl, ll, pl)
    right = ORBmatcher.rig_camera(orr["kps"], orr["desc"], orr["n"], gr, lr, pr)
    res = ORBmatcher.search_by_projection_rig(mode, left, right, BOUNDS, qt, nn_ratio, check_ori)
    torch.cuda.synchronize()
    got = {k_: v.cpu().numpy() for k_, v in res.items()}
    (kl, dl, nl), (kr, dr, nr) = rig["L"], rig["R"]
    tot = 0
    for f, q in enumerate(qs):
        ref = oracle_mod.search_by_projection_rig(
            mode, kl[f, :nl[f]], dl[f, :nl[f]], kr[f, :nr[f]], dr[f, :nr[f]], BOUNDS, q, nn_ratio, check_ori,
            locked_l=None if locked is None else locked[0][f, :nl[f]],
            locked_r=None if locked is None else locked[1][f, :nr[f]],
            l2r=None if pairs is None else pairs[f][0], r2l=None if pairs is None else pairs[f][1])
        m = len(q["qx"])
        assert np.array_equal(got["match_l"][f, :m], ref["match_l"]), (f, np.nonzero(got["match_l"][f, :m] != ref["match_l"])[0][:8])
        assert np.array_equal(got["match_r"][f, :m], ref["match_r"]), (f, np.nonzero(got["match_r"][f, :m] != ref["match_r"])[0][:8])
        assert np.array_equal(got["feat_query_l"][f, :nl[f]], ref["feat_query_l"])
        assert np.array_equal(got["feat_query_r"][f, :nr[f]], ref["feat_query_r"])
        assert got["nmatches"][f] == ref["nmatches"]
        tot += ref["nmatches"]
    return tot, got


@pytest.mark.parametrize("nn_ratio", [0.8, 0.6])
def test_search_by_projection_rig_mappoints(rig, oracle_mod, nn_ratio):
    """ORBmatcher.cc:43-213 with Nleft != -1: pair propagation, locks set and cleared, the skipped right search."""
    rng = np.random.default_rng(41)
    B = rig["B"]
    cap = rig["ol"]["kps"].shape[1]
    nl, nr = rig["L"][2], rig["R"][2]
    qs = [_queries(rig, f, rng, 12.0, (-1, 0)) for f in range(B)]
    pairs = [_pairs(nl[f], nr[f], rng) for f in range(B)]
    locked = ((rng.random((B, cap)) < 0.15).astype(np.uint8), (rng.random((B, cap)) < 0.15).astype(np.uint8))
    tot, got = _run(rig, oracle_mod, _lib.PROJ_MAPPOINTS, qs, nn_ratio, False, locked, pairs)
    assert tot > B * 800
    assert (got["match_l"] >= 0).sum() > B * 200 and (got["match_r"] >= 0).sum() > B * 200
    # no pairs, no locks: the two cameras only interact through the skipped right search
    tot2, _ = _run(rig, oracle_mod, _lib.PROJ_MAPPOINTS, qs, nn_ratio, False)
    assert tot2 > B * 500


def test_search_by_projection_rig_mappoints_duplicates_rescan(rig, oracle_mod):
    """Every MapPoint three times, tight radius: later copies find their list entries locked and fall back to the
    warp rescan; unlocked copies clear entry locks through the pairs (the 'dirty' path)."""
    rng = np.random.default_rng(43)
    B = rig["B"]
    cap = rig["ol"]["kps"].shape[1]
    nl, nr = rig["L"][2], rig["R"][2]
    qs = [_queries(rig, f, rng, 20.0, (-1, 0), 0.05, 0.05, 0.5, dup=3) for f in range(B)]
    pairs = [_pairs(nl[f], nr[f], rng, 0.8) for f in range(B)]
    locked = ((rng.random((B, cap)) < 0.5).astype(np.uint8), (rng.random((B, cap)) < 0.5).astype(np.uint8))
    tot, _ = _run(rig, oracle_mod, _lib.PROJ_MAPPOINTS, qs, 0.9, False, locked, pairs)
    assert tot > B * 1000


@pytest.mark.parametrize("check_ori", [True, False])
def test_search_by_projection_rig_lastframe(rig, oracle_mod, check_ori):
    """ORBmatcher.cc:1676-1887 with Nleft != -1: band [o-1,o+1], best only, right searched only when the left window
    was not empty, one rotation histogram over both cameras."""
    rng = np.random.default_rng(47)
    B = rig["B"]
    qs = []
    for f in range(B):
        q = _queries(rig, f, rng, 15.0, (-1, 1), 0.1, 0.0, 0.2)
        q["flags"] |= 4
        # some last-frame keypoints carry a rotated angle: minor histogram bins that ComputeThreeMaxima erases
        rot = rng.random(len(q["angle"])) < 0.04
        q["angle"] = np.where(rot, (q["angle"] + np.float32(95.0)) % np.float32(360.0), q["angle"]).astype(np.float32)
        qs.append(q)
    tot, got = _run(rig, oracle_mod, _lib.PROJ_LASTFRAME, qs, 0.9, check_ori)
    assert tot > B * 1500
    cap_l = (got["match_l"] >= 0).sum()
    assert cap_l > B * 500 and (got["match_r"] >= 0).sum() > B * 500


def test_search_by_bow_rig(rig, oracle_mod):
    """SearchByBoW(KeyFrame*, Frame&) with Nleft != -1: keyframe = left features of frame s, frame = left|right
    features of frame f in one numbering."""
    import torch
    from orb_slam3_vio_fixes_b200.matcher import ORBmatcher
    (kl, dl, nl), (kr, dr, nr) = rig["L"], rig["R"]
    B = rig["B"]
    pool = np.concatenate([dl[f, :nl[f]] for f in range(B)])
    voc = S.make_vocabulary(10, 4, seed=3, desc_pool=pool)
    pairs = [(0, 1), (1, 2), (2, 3), (3, 0)]
    cap = 2 * dl.shape[1]
    P = len(pairs)
    k1 = np.zeros((P, cap), kl.dtype); d1 = np.zeros((P, cap, 32), np.uint8); n1 = np.zeros(P, np.int32)
    k2 = np.zeros((P, cap), kl.dtype); d2 = np.zeros((P, cap, 32), np.uint8); n2 = np.zeros(P, np.int32)
    nleft = np.zeros(P, np.int32)
    rng = np.random.default_rng(5)
    valid1 = (rng.random((P, cap)) < 0.85).astype(np.uint8)
    fvs1, fvs2 = [], []
    for p, (s, f) in enumerate(pairs):
        n1[p] = nl[s]; k1[p, :nl[s]] = kl[s, :nl[s]]; d1[p, :nl[s]] = dl[s, :nl[s]]
        n2[p] = nl[f] + nr[f]; nleft[p] = nl[f]
        k2[p, :nl[f]] = kl[f, :nl[f]]; k2[p, nl[f]:n2[p]] = kr[f, :nr[f]]
        d2[p, :nl[f]] = dl[f, :nl[f]]; d2[p, nl[f]:n2[p]] = dr[f, :nr[f]]
        fvs1.append(oracle_mod.bow_transform(d1[p, :n1[p]], voc, 2)["fv"])
        fvs2.append(oracle_mod.bow_transform(d2[p, :n2[p]], voc, 2)["fv"])
    ns = max(max(len(a[0]) for a in fvs1), max(len(a[0]) for a in fvs2))

    def fv_tensors(fvs):
        node = np.zeros((P, ns), np.int32); off = np.zeros((P, ns + 1), np.int32); idx = np.zeros((P, cap), np.int32)
        nn = np.zeros(P, np.int32)
        for p, (a, b, c) in enumerate(fvs):
            nn[p] = len(a); node[p, :len(a)] = a; off[p, :len(b)] = b; idx[p, :len(c)] = c
        return {"node": _t(node), "off": _t(off), "idx": _t(idx), "nn": _t(nn)}

    kt = lambda k: _t(k.view(np.uint8).reshape(P, cap, 28))
    for nn_ratio, ori in ((0.7, True), (0.9, False)):
        res = ORBmatcher.search_by_bow_rig(kt(k1), _t(d1), _t(n1), fv_tensors(fvs1), kt(k2), _t(d2), _t(n2), _t(nleft),
                                           fv_tensors(fvs2), nn_ratio, ori, valid1=_t(valid1))
        torch.cuda.synchronize()
        tot = 0
        for p in range(P):
            m12l, m12r, m21, nm = oracle_mod.search_by_bow_rig(k1[p, :n1[p]], d1[p, :n1[p]], fvs1[p], k2[p, :n2[p]],
                                                               d2[p, :n2[p]], nleft[p], fvs2[p], nn_ratio, ori,
                                                               valid1=valid1[p, :n1[p]])
            assert np.array_equal(res["match12_left"][p, :n1[p]].cpu().numpy(), m12l)
            assert np.array_equal(res["match12_right"][p, :n1[p]].cpu().numpy(), m12r)
            assert np.array_equal(res["match21"][p, :n2[p]].cpu().numpy(), m21)
            assert int(res["nmatches"][p]) == nm
            tot += nm
            assert (m12r >= 0).sum() > 50 and (m12l >= 0).sum() > 50
        assert tot > P * 150


import os  # noqa: E402


@pytest.mark.parametrize("seed", range(int(os.environ.get("ORBX_STRESS_SEEDS", "3"))))
def test_search_by_projection_rig_random_sweep(rig, oracle_mod, seed):
    """Randomised configurations of the rig projection search (both variants): lock density, pair fraction, duplicated
    and unobserved MapPoints (the rescan and 'entry lock lost' paths), radius, ratio.  ORBX_STRESS_SEEDS widens it."""
    rng = np.random.default_rng(1000 + seed)
    B = rig["B"]
    cap = rig["ol"]["kps"].shape[1]
    nl, nr = rig["L"][2], rig["R"][2]
    th = float(rng.choice([4.0, 8.0, 12.0, 20.0, 30.0]))
    dup = int(rng.choice([1, 1, 2, 3]))
    unl = float(rng.choice([0.0, 0.2, 0.6]))
    lock = float(rng.choice([0.0, 0.15, 0.5]))
    pf = float(rng.choice([0.0, 0.4, 0.9]))
    nn = float(rng.choice([0.6, 0.8, 0.95]))
    qs = [_queries(rig, f, rng, th, (-1, 0), float(rng.choice([0.0, 0.1, 0.4])), float(rng.choice([0.0, 0.15, 0.4])), unl,
                   dup=dup) for f in range(B)]
    pairs = [_pairs(nl[f], nr[f], rng, pf) for f in range(B)]
    locked = ((rng.random((B, cap)) < lock).astype(np.uint8), (rng.random((B, cap)) < lock).astype(np.uint8))
    tot, _ = _run(rig, oracle_mod, _lib.PROJ_MAPPOINTS, qs, nn, False, locked, pairs)
    assert tot > 0
    qs2 = []
    for f in range(B):
        q = _queries(rig, f, rng, th, (-1, 1), float(rng.choice([0.0, 0.2])), 0.0, unl, dup=dup)
        q["flags"] |= 4
        rot = rng.random(len(q["angle"])) < 0.05
        q["angle"] = np.where(rot, (q["angle"] + np.float32(95.0)) % np.float32(360.0), q["angle"]).astype(np.float32)
        qs2.append(q)
    tot2, _ = _run(rig, oracle_mod, _lib.PROJ_LASTFRAME, qs2, 0.9, bool(rng.integers(0, 2)), locked)
    assert tot2 > 0
