"""GPU parity (through the C ABI) of the vocabulary-node matching functions against the oracle:
ORBmatcher::SearchByBoW x2, SearchForTriangulation, DBoW2 transform, ComputeDistinctiveDescriptors."""
import numpy as np
import pytest

from orb_slam3_vio_fixes_b200 import _lib
from orb_slam3_vio_fixes_b200 import synthetic as S
from orb_slam3_vio_fixes_b200.matcher import ORBmatcher, ORBVocabulary

pytestmark = pytest.mark.gpu


def _t(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.fixture(scope="module")
def frames_feats(oracle_mod):
    """keypoints + descriptors of 6 consecutive C1 frames (CPU oracle extractor) and a vocabulary whose node
    centres are drawn from them."""
    frames = S.make_frames("C1", 6)
    orc = oracle_mod.Extractor(1000, 1.2, 8, 20, 7)
    feats = [orc.extract(f, (0, 0))[:2] for f in frames]
    pool = np.concatenate([d for _, d in feats])
    voc = oracle_mod.make_vocabulary(10, 4, seed=3, desc_pool=pool)
    sf = np.array([1.2 ** i for i in range(8)], np.float32)
    return feats, voc, sf


def _pad_batch(feats, cap):
    b = len(feats)
    kps = np.zeros((b, cap), _lib.KP_DTYPE)
    desc = np.zeros((b, cap, 32), np.uint8)
    n = np.zeros(b, np.int32)
    for i, (k, d) in enumerate(feats):
        kps[i, :len(k)] = k
        desc[i, :len(k)] = d
        n[i] = len(k)
    return kps, desc, n


def _fv_batch(fvs, cap, ns):
    b = len(fvs)
    fv = {"node": np.zeros((b, ns), np.int32), "off": np.zeros((b, ns + 1), np.int32),
          "idx": np.zeros((b, cap), np.int32), "nn": np.zeros(b, np.int32)}
    for i, (node, off, idx) in enumerate(fvs):
        fv["node"][i, :len(node)] = node
        fv["off"][i, :len(off)] = off
        fv["idx"][i, :len(idx)] = idx
        fv["nn"][i] = len(node)
    return {k: _t(v) for k, v in fv.items()}


def _kt(kps):
    return _t(kps.view(np.float32).reshape(kps.shape + (7,)))


def test_bow_transform_matches_oracle(oracle_mod, frames_feats):
    import torch
    feats, voc, _ = frames_feats
    cap = 1100
    kps, desc, n = _pad_batch(feats, cap)
    V = ORBVocabulary(voc["child_off"], voc["child_idx"], voc["node_desc"], voc["node_word"], voc["node_weight"],
                      voc["L"])
    for levelsup in (2, 4, 1):
        out = V.transform(_t(desc), _t(n), levelsup=levelsup)
        torch.cuda.synchronize()
        for f, (k, d) in enumerate(feats):
            ref = oracle_mod.bow_transform(d, voc, levelsup)
            m = len(k)
            assert np.array_equal(out["feat_word"][f, :m].cpu().numpy(), ref["feat_word"])
            assert np.array_equal(out["feat_node"][f, :m].cpu().numpy(), ref["feat_node"])
            assert np.array_equal(out["feat_weight"][f, :m].cpu().numpy(), ref["feat_weight"])
            nb = int(out["bow_n"][f])
            assert nb == len(ref["bow_word"])
            assert np.array_equal(out["bow_word"][f, :nb].cpu().numpy(), ref["bow_word"])
            assert np.array_equal(out["bow_val"][f, :nb].cpu().numpy(), ref["bow_val"])  # fp64, bit-exact
            nn = int(out["fv"]["nn"][f])
            node, off, idx = ref["fv"]
            assert nn == len(node)
            assert np.array_equal(out["fv"]["node"][f, :nn].cpu().numpy(), node)
            assert np.array_equal(out["fv"]["off"][f, :nn + 1].cpu().numpy(), off)
            assert np.array_equal(out["fv"]["idx"][f, :off[-1]].cpu().numpy(), idx)


def test_bow_transform_small_node_stride_reports_needed(oracle_mod, frames_feats):
    feats, voc, _ = frames_feats
    kps, desc, n = _pad_batch(feats[:1], 1100)
    V = ORBVocabulary(voc["child_off"], voc["child_idx"], voc["node_desc"], voc["node_word"], voc["node_weight"],
                      voc["L"])
    ref = oracle_mod.bow_transform(feats[0][1], voc, 2)
    out = V.transform(_t(desc), _t(n), levelsup=2, node_stride=8)
    assert int(out["fv"]["nn"][0]) == -len(ref["fv"][0])


@pytest.mark.parametrize("mode", [_lib.BOW_KF_FRAME, _lib.BOW_KF_KF])
@pytest.mark.parametrize("levelsup", [2, 3])
def test_search_by_bow_matches_oracle(oracle_mod, frames_feats, mode, levelsup):
    import torch
    feats, voc, _ = frames_feats
    rng = np.random.default_rng(5 + mode)
    cap, ns = 1100, 1100
    pairs = [(0, 1), (1, 2), (2, 4), (5, 3), (4, 4)]
    f1 = [feats[a] for a, _ in pairs]
    f2 = [feats[b] for _, b in pairs]
    fv1 = [oracle_mod.bow_transform(d, voc, levelsup)["fv"] for _, d in f1]
    fv2 = [oracle_mod.bow_transform(d, voc, levelsup)["fv"] for _, d in f2]
    k1, d1, n1 = _pad_batch(f1, cap)
    k2, d2, n2 = _pad_batch(f2, cap)
    v1 = (rng.random((len(pairs), cap)) < 0.8).astype(np.uint8)
    v2 = (rng.random((len(pairs), cap)) < 0.9).astype(np.uint8)
    for ratio, ori in ((0.6, True), (0.9, False), (0.75, True)):
        out = ORBmatcher.search_by_bow(mode, _kt(k1), _t(d1), _t(n1), _fv_batch(fv1, cap, ns), _kt(k2), _t(d2), _t(n2),
                                       _fv_batch(fv2, cap, ns), ratio, ori, _t(v1), _t(v2))
        torch.cuda.synchronize()
        total = 0
        for p in range(len(pairs)):
            m12, m21, nm = oracle_mod.search_by_bow(mode, f1[p][0], f1[p][1], fv1[p], f2[p][0], f2[p][1], fv2[p], ratio,
                                                    ori, v1[p, :n1[p]], v2[p, :n2[p]])
            assert int(out["nmatches"][p]) == nm
            assert np.array_equal(out["match12"][p, :n1[p]].cpu().numpy(), m12)
            assert np.array_equal(out["match21"][p, :n2[p]].cpu().numpy(), m21)
            total += nm
        assert total > 300  # the scene really produces matches


def test_search_by_bow_no_flags_and_ties(oracle_mod):
    """few distinct descriptors (many exact ties), big nodes (> 32 candidates), no validity flags"""
    import torch
    rng = np.random.default_rng(11)
    n, cap = 700, 704
    base = rng.integers(0, 256, (25, 32), dtype=np.uint8)

    def mk():
        d = base[rng.integers(0, 25, n)].copy()
        for i in range(n):
            for _ in range(rng.integers(0, 3)):
                d[i, rng.integers(0, 32)] ^= np.uint8(1 << rng.integers(0, 8))
        k = np.zeros(n, _lib.KP_DTYPE)
        k["angle"] = rng.uniform(0, 360, n).astype(np.float32)
        nodes = rng.integers(0, 6, n) * 17 + 100
        groups = {}
        for i, nd in enumerate(nodes):
            groups.setdefault(int(nd), []).append(i)
        nl = sorted(groups)
        off = np.cumsum([0] + [len(groups[x]) for x in nl]).astype(np.int32)
        idx = np.concatenate([groups[x] for x in nl]).astype(np.int32)
        return k, d, (np.array(nl, np.int32), off, idx)

    a, b = mk(), mk()
    for mode in (0, 1):
        out = ORBmatcher.search_by_bow(mode, _kt(a[0][None]), _t(a[1][None]), _t(np.array([n], np.int32)),
                                       _fv_batch([a[2]], n, 16), _kt(b[0][None]), _t(b[1][None]),
                                       _t(np.array([n], np.int32)), _fv_batch([b[2]], n, 16), 0.9, True)
        torch.cuda.synchronize()
        m12, m21, nm = oracle_mod.search_by_bow(mode, a[0], a[1], a[2], b[0], b[1], b[2], 0.9, True)
        assert int(out["nmatches"][0]) == nm and nm > 20
        assert np.array_equal(out["match12"][0].cpu().numpy(), m12)
        assert np.array_equal(out["match21"][0].cpu().numpy(), m21)


@pytest.mark.parametrize("coarse,only_stereo", [(False, False), (True, False), (False, True)])
def test_search_for_triangulation_matches_oracle(oracle_mod, frames_feats, coarse, only_stereo):
    import torch
    feats, voc, sf = frames_feats
    rng = np.random.default_rng(21)
    cap, ns = 1100, 1100
    pairs = [(0, 1), (1, 3), (2, 2), (5, 4)]
    f1 = [feats[a] for a, _ in pairs]
    f2 = [feats[b] for _, b in pairs]
    fv1 = [oracle_mod.bow_transform(d, voc, 2)["fv"] for _, d in f1]
    fv2 = [oracle_mod.bow_transform(d, voc, 2)["fv"] for _, d in f2]
    k1, d1, n1 = _pad_batch(f1, cap)
    k2, d2, n2 = _pad_batch(f2, cap)
    B = len(pairs)
    h1 = (rng.random((B, cap)) < 0.3).astype(np.uint8)
    h2 = (rng.random((B, cap)) < 0.3).astype(np.uint8)
    u1 = np.where(rng.random((B, cap)) < 0.5, rng.uniform(0, 700, (B, cap)), -1.0).astype(np.float32)
    u2 = np.where(rng.random((B, cap)) < 0.5, rng.uniform(0, 700, (B, cap)), -1.0).astype(np.float32)
    # camera motion: the synthetic sequence shifts by (2,1) px per frame (a sideways translation), so the epipolar
    # lines are close to the motion direction; perturbed a little so that the 3.84*sigma2 gate really cuts
    F = np.zeros((B, 9), np.float32)
    ep = np.zeros((B, 2), np.float32)
    for p, (a, b) in enumerate(pairs):
        t = np.array([2.0 * (b - a) + 0.3, 1.0 * (b - a) - 0.2, 0.01 * (p + 1)], np.float64)
        tx = np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]])
        F[p] = (tx / max(1e-9, np.linalg.norm(t))).astype(np.float32).reshape(-1)
        ep[p] = (376 + 50 * t[0], 240 + 50 * t[1])
    sig2 = (sf * sf).astype(np.float32)
    out = ORBmatcher.search_for_triangulation(_kt(k1), _t(d1), _t(n1), _fv_batch(fv1, cap, ns), _kt(k2), _t(d2), _t(n2),
                                              _fv_batch(fv2, cap, ns), _t(F), _t(ep), sf, sig2, only_stereo, coarse,
                                              True, _t(h1), _t(h2), _t(u1), _t(u2))
    torch.cuda.synchronize()
    total = 0
    for p in range(B):
        m12, nm = oracle_mod.search_for_triangulation(f1[p][0], f1[p][1], fv1[p], f2[p][0], f2[p][1], fv2[p], F[p], ep[p],
                                                      sf, sig2, only_stereo, coarse, True, h1[p, :n1[p]], h2[p, :n2[p]],
                                                      u1[p, :n1[p]], u2[p, :n2[p]])
        assert int(out["nmatches"][p]) == nm
        assert np.array_equal(out["match12"][p, :n1[p]].cpu().numpy(), m12)
        total += nm
    assert total > 50
    # monocular keyframes, nothing triangulated yet: every optional input NULL
    out = ORBmatcher.search_for_triangulation(_kt(k1), _t(d1), _t(n1), _fv_batch(fv1, cap, ns), _kt(k2), _t(d2), _t(n2),
                                              _fv_batch(fv2, cap, ns), _t(F), _t(ep), sf, sig2, False, coarse, False)
    torch.cuda.synchronize()
    for p in range(B):
        m12, nm = oracle_mod.search_for_triangulation(f1[p][0], f1[p][1], fv1[p], f2[p][0], f2[p][1], fv2[p], F[p], ep[p],
                                                      sf, sig2, False, coarse, False)
        assert int(out["nmatches"][p]) == nm
        assert np.array_equal(out["match12"][p, :n1[p]].cpu().numpy(), m12)


def test_distinctive_descriptors_match_oracle(oracle_mod, frames_feats):
    import torch
    feats, _, _ = frames_feats
    rng = np.random.default_rng(31)
    pool = np.concatenate([d for _, d in feats])
    sizes = [1, 2, 3, 4, 5, 8, 31, 32, 33, 64, 100] + list(rng.integers(1, 40, 200))
    off = np.zeros(len(sizes) + 1, np.int32)
    off[1:] = np.cumsum(sizes)
    idx = np.zeros(off[-1], np.int32)
    for m, s in enumerate(sizes):
        # observations of one MapPoint: noisy copies of one descriptor + a few outliers
        c = rng.integers(0, len(pool))
        idx[off[m]:off[m + 1]] = np.where(rng.random(s) < 0.8, c, rng.integers(0, len(pool), s))
    pool = pool.copy()
    flip = rng.integers(0, len(pool), 4000)
    pool[flip, rng.integers(0, 32, 4000)] ^= (1 << rng.integers(0, 8, 4000)).astype(np.uint8)
    best, med = ORBmatcher.distinctive_descriptors(_t(pool), _t(off), _t(idx))
    torch.cuda.synchronize()
    best = best.cpu().numpy()
    for m in range(len(sizes)):
        assert best[m] == oracle_mod.distinctive_descriptor(pool[idx[off[m]:off[m + 1]]]), m


def test_fuse_candidates_match_oracle(oracle_mod, frames_feats):
    """Candidate scoring of ORBmatcher::Fuse: MapPoints = keypoints of neighbouring frames projected with noise."""
    import torch
    from orb_slam3_vio_fixes_b200.matcher import results_to_numpy
    feats, _, sf = frames_feats
    rng = np.random.default_rng(41)
    cap = 1100
    B = 4
    kps, desc, n = _pad_batch(feats[:B], cap)
    bounds = (0.0, 752.0, 0.0, 480.0)
    inv = (1.0 / (sf * sf)).astype(np.float32)
    Q = 2200
    qx, qy, qr, qur = (np.zeros((B, Q), np.float32) for _ in range(4))
    ql = np.zeros((B, Q), np.int8)
    qd = np.zeros((B, Q, 32), np.uint8)
    nq = np.zeros(B, np.int32)
    ur = np.where(rng.random((B, cap)) < 0.5, kps["x"] - 24.0 + rng.normal(0, 0.5, (B, cap)), -1.0).astype(np.float32)
    for f in range(B):
        src = [feats[(f + 1) % len(feats)], feats[(f + 2) % len(feats)]]
        k = np.concatenate([s[0] for s in src])
        d = np.concatenate([s[1] for s in src])
        m = len(k)
        shift = np.repeat(np.array([[-2.0, -1.0], [-4.0, -2.0]], np.float32), [len(src[0][0]), len(src[1][0])], 0)
        qx[f, :m] = k["x"] + shift[:, 0] + rng.normal(0, 1.0, m)
        qy[f, :m] = k["y"] + shift[:, 1] + rng.normal(0, 1.0, m)
        ql[f, :m] = np.clip(k["octave"] + rng.integers(-1, 2, m), 0, 7)
        qr[f, :m] = 3.0 * sf[ql[f, :m]]
        qur[f, :m] = qx[f, :m] - 24.0
        qd[f, :m] = d
        nq[f] = m
    grid = ORBmatcher.grid_build(_kt(kps), _t(n), bounds)
    for with_stereo in (False, True):
        out = ORBmatcher.fuse_candidates(_kt(kps), _t(desc), _t(n), grid, bounds, inv, _t(qx), _t(qy), _t(qr), _t(ql),
                                         _t(qd), _t(nq), _t(ur) if with_stereo else None,
                                         _t(qur) if with_stereo else None)
        torch.cuda.synchronize()
        res = results_to_numpy(out)
        hits = 0
        for f in range(B):
            m = nq[f]
            bi, bd = oracle_mod.fuse_candidates(kps[f, :n[f]], desc[f, :n[f]], bounds, inv, qx[f, :m], qy[f, :m], qr[f, :m],
                                                ql[f, :m], qd[f, :m], ur[f, :n[f]] if with_stereo else None,
                                                qur[f, :m] if with_stereo else None)
            assert np.array_equal(res["best_idx"][f, :m], bi)
            assert np.array_equal(res["best_dist"][f, :m], bd)
            hits += int(np.sum(bd <= 50))
        assert hits > 500
    # SearchBySim3's scoring loops (ORBmatcher.cc:1543-1580, 1623-1660): same walk and level band, no chi-square
    # gate (an all-zero invLevelSigma2 disables it), accept best <= TH_HIGH, then mutual agreement on the host
    zero = np.zeros(8, np.float32)
    out = ORBmatcher.fuse_candidates(_kt(kps), _t(desc), _t(n), grid, bounds, zero, _t(qx), _t(qy), _t(qr), _t(ql), _t(qd),
                                     _t(nq))
    torch.cuda.synchronize()
    res = results_to_numpy(out)
    more = 0
    for f in range(B):
        m = nq[f]
        bi, bd = oracle_mod.fuse_candidates(kps[f, :n[f]], desc[f, :n[f]], bounds, zero, qx[f, :m], qy[f, :m], qr[f, :m],
                                            ql[f, :m], qd[f, :m])
        assert np.array_equal(res["best_idx"][f, :m], bi) and np.array_equal(res["best_dist"][f, :m], bd)
        more += int(np.sum(bd <= 100))
    assert more > hits


@pytest.mark.parametrize("check_ori", [True, False])
def test_search_for_initialization_matches_oracle(oracle_mod, frames_feats, check_ori):
    import torch
    feats, _, _ = frames_feats
    cap = 1100
    pairs = [(0, 1), (0, 3), (2, 5), (4, 4)]
    f1 = [feats[a] for a, _ in pairs]
    f2 = [feats[b] for _, b in pairs]
    k1, d1, n1 = _pad_batch(f1, cap)
    k2, d2, n2 = _pad_batch(f2, cap)
    bounds = (0.0, 752.0, 0.0, 480.0)
    pm = np.stack([k1["x"], k1["y"]], -1).astype(np.float32)   # vbPrevMatched starts at F1's own keypoints
    grid2 = ORBmatcher.grid_build(_kt(k2), _t(n2), bounds)
    for window, ratio in ((100, 0.9), (30, 0.7)):
        pm_t = _t(pm.copy())
        out = ORBmatcher.search_for_initialization(_kt(k1), _t(d1), _t(n1), _kt(k2), _t(d2), _t(n2), grid2, bounds, pm_t,
                                                   window, ratio, check_ori)
        torch.cuda.synchronize()
        tot = 0
        for p in range(len(pairs)):
            m12, nm, pm2 = oracle_mod.search_for_initialization(f1[p][0], f1[p][1], f2[p][0], f2[p][1], bounds,
                                                                pm[p, :n1[p]], window, ratio, check_ori)
            assert int(out["nmatches"][p]) == nm
            assert np.array_equal(out["match12"][p, :n1[p]].cpu().numpy(), m12)
            assert np.array_equal(pm_t[p, :n1[p]].cpu().numpy(), pm2)
            tot += nm
        assert tot > 150
