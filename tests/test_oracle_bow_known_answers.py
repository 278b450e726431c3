"""Known-answer tests of the oracle's vocabulary-node matching restatements (SearchByBoW x2,
SearchForTriangulation, DBoW2 transform, ComputeDistinctiveDescriptors) on hand-made cases whose results
follow from reading the reference loops (ORBmatcher.cc:223-425, 765-905, 907-1146; TemplatedVocabulary.h:
1127-1259; MapPoint.cc:367-401)."""
import numpy as np

from orb_slam3_vio_fixes_b200._lib import KP_DTYPE


def _kps(n, angle=0.0, octave=0, xy=None):
    k = np.zeros(n, KP_DTYPE)
    k["angle"] = angle
    k["octave"] = octave
    if xy is not None:
        k["x"], k["y"] = xy[:, 0], xy[:, 1]
    return k


def _desc_with_bits(nbits_list):
    d = np.zeros((len(nbits_list), 32), np.uint8)
    for i, nb in enumerate(nbits_list):
        bits = np.zeros(256, np.uint8)
        bits[:nb] = 1
        d[i] = np.packbits(bits)
    return d


def _fv(groups):
    """groups: {node: [feature indices]} -> CSR"""
    nodes = sorted(groups)
    off, idx = [0], []
    for nd in nodes:
        idx += list(groups[nd])
        off.append(len(idx))
    return np.array(nodes, np.int32), np.array(off, np.int32), np.array(idx, np.int32)


def test_bow_threshold_boundary_and_ratio(oracle_mod):
    # one node; query = zeros; candidates at distance 50 and 200: mode 0 accepts (<=50), mode 1 rejects (<50)
    d1 = _desc_with_bits([0])
    d2 = _desc_with_bits([50, 200])
    fv1, fv2 = _fv({7: [0]}), _fv({7: [0, 1]})
    for mode, expect in ((0, 0), (1, -1)):
        m12, m21, nm = oracle_mod.search_by_bow(mode, _kps(1), d1, fv1, _kps(2), d2, fv2, 0.6, False)
        assert m12[0] == expect and nm == (expect >= 0)
    # ratio: best 30, second 45 -> 30 < 0.6*45 = 27 false -> no match; with nnratio 0.7 -> 31.5 true
    d2 = _desc_with_bits([30, 45])
    assert oracle_mod.search_by_bow(0, _kps(1), d1, fv1, _kps(2), d2, fv2, 0.6, False)[2] == 0
    assert oracle_mod.search_by_bow(0, _kps(1), d1, fv1, _kps(2), d2, fv2, 0.7, False)[2] == 1


def test_bow_already_matched_skip_is_sequential(oracle_mod):
    # two identical queries in one node, one perfect candidate + one far candidate: the first query takes the
    # perfect one; the second then only sees the far candidate (bestDist2 = 256, ratio passes if <= TH_LOW)
    d1 = _desc_with_bits([0, 0])
    d2 = _desc_with_bits([0, 40])
    m12, m21, nm = oracle_mod.search_by_bow(0, _kps(2), d1, _fv({3: [0, 1]}), _kps(2), d2, _fv({3: [0, 1]}), 0.6, False)
    assert list(m12) == [0, 1] and list(m21) == [0, 1] and nm == 2
    # different nodes never meet
    m12, _, nm = oracle_mod.search_by_bow(0, _kps(2), d1, _fv({3: [0], 4: [1]}), _kps(2), d2, _fv({3: [1], 5: [0]}),
                                          0.6, False)
    assert list(m12) == [1, -1] and nm == 1
    # invalid side-1 features (no MapPoint) are skipped; invalid side-2 only matter in mode 1
    v1 = np.array([0, 1], np.uint8)
    m12, _, nm = oracle_mod.search_by_bow(0, _kps(2), d1, _fv({3: [0, 1]}), _kps(2), d2, _fv({3: [0, 1]}), 0.6, False,
                                          valid1=v1)
    assert list(m12) == [-1, 0] and nm == 1
    v2 = np.array([0, 1], np.uint8)
    m12, _, nm = oracle_mod.search_by_bow(1, _kps(2), d1, _fv({3: [0, 1]}), _kps(2), d2, _fv({3: [0, 1]}), 0.6, False,
                                          valid2=v2)
    assert list(m12) == [1, -1] and nm == 1


def test_bow_rotation_histogram_erases_minor_bins(oracle_mod):
    n = 24
    d = np.random.default_rng(0).integers(0, 256, (n, 32), dtype=np.uint8)
    k1, k2 = _kps(n, angle=10.0), _kps(n, angle=10.0)
    k1["angle"][0] = 200.0   # a lone match in another bin: 1 < 0.1*23 -> erased
    fv = _fv({i: [i] for i in range(n)})
    m12, m21, nm = oracle_mod.search_by_bow(1, k1, d, fv, k2, d, fv, 0.6, True)
    assert nm == n - 1 and m12[0] == -1 and m21[0] == -1 and list(m12[1:]) == list(range(1, n))
    m12, _, nm = oracle_mod.search_by_bow(1, k1, d, fv, k2, d, fv, 0.6, False)
    assert nm == n and m12[0] == 0


def test_triangulation_last_minimum_and_gates(oracle_mod):
    d1 = _desc_with_bits([0])
    d2 = _desc_with_bits([20, 20, 51, 10])
    xy2 = np.array([[100, 50], [120, 50], [130, 50], [300, 300]], np.float32)
    k1 = _kps(1, xy=np.array([[90, 50]], np.float32))
    k2 = _kps(4, xy=xy2)
    fv1, fv2 = _fv({1: [0]}), _fv({1: [0, 1, 2, 3]})
    sc, sg = np.ones(8, np.float32), np.ones(8, np.float32)
    ep = np.array([-1e4, -1e4], np.float32)
    # coarse: no epipolar test: min dist 10 (idx 3) wins; without idx 3 the LAST of the two 20s wins
    m12, nm = oracle_mod.search_for_triangulation(k1, d1, fv1, k2, d2, fv2, None, ep, sc, sg, coarse=True,
                                                  check_orientation=False)
    assert m12[0] == 3 and nm == 1
    m12, _ = oracle_mod.search_for_triangulation(k1, d1, fv1, k2[:3], d2[:3], _fv({1: [0, 1, 2]}), None, ep, sc, sg,
                                                 coarse=True, check_orientation=False)
    assert m12[0] == 1
    # epipolar line y = 50 (pure x translation: F = [t]x with t = (1,0,0)): idx 3 is far from the line
    F = np.array([0, 0, 0, 0, 0, -1, 0, 1, 0], np.float32)
    m12, _ = oracle_mod.search_for_triangulation(k1, d1, fv1, k2, d2, fv2, F, ep, sc, sg, check_orientation=False)
    assert m12[0] == 1
    # epipole gate: candidates within 10 px (100*scale) of the epipole are dropped for mono-mono pairs
    ep2 = np.array([121, 50], np.float32)
    m12, _ = oracle_mod.search_for_triangulation(k1, d1, fv1, k2, d2, fv2, F, ep2, sc, sg, check_orientation=False)
    assert m12[0] == 0
    # ... but not when one side is stereo
    m12, _ = oracle_mod.search_for_triangulation(k1, d1, fv1, k2, d2, fv2, F, ep2, sc, sg, check_orientation=False,
                                                 u_right1=np.array([5.0], np.float32))
    assert m12[0] == 1
    # features that already hold a MapPoint are skipped; only_stereo needs both sides stereo
    m12, _ = oracle_mod.search_for_triangulation(k1, d1, fv1, k2, d2, fv2, F, ep, sc, sg, check_orientation=False,
                                                 has_mp2=np.array([0, 1, 0, 0], np.uint8))
    assert m12[0] == 0
    m12, nm = oracle_mod.search_for_triangulation(k1, d1, fv1, k2, d2, fv2, F, ep, sc, sg, only_stereo=True,
                                                  check_orientation=False)
    assert m12[0] == -1 and nm == 0


def test_bow_transform_small_tree(oracle_mod):
    # depth-2 binary tree: root -> {1,2}; 1 -> {3,4}; 2 -> {5,6}
    voc = {"child_off": np.array([0, 2, 4, 6, 6, 6, 6, 6], np.int32), "child_idx": np.array([1, 2, 3, 4, 5, 6], np.int32),
           "node_desc": _desc_with_bits([0, 0, 256, 0, 64, 192, 256]),
           "node_word": np.array([-1, -1, -1, 0, 1, 2, 3], np.int32),
           "node_weight": np.array([0, 0, 0, 1.0, 2.0, 0.0, 4.0]), "L": 2}
    feats = _desc_with_bits([10, 60, 250, 128, 200, 5])
    r = oracle_mod.bow_transform(feats, voc, levelsup=1)
    # 128 bits: tie between node 1 and 2 -> first child wins (d < best_d); then |128-0| vs |128-64| -> node 4
    assert list(r["feat_word"]) == [0, 1, 3, 1, 2, 0]
    assert list(r["feat_node"]) == [1, 1, 2, 1, 2, 1]
    # word 2 has weight 0 (stopped): feature 4 is in neither vector
    assert list(r["bow_word"]) == [0, 1, 3]
    assert np.array_equal(r["bow_val"], np.array([2.0, 4.0, 4.0]) / 10.0)
    node, off, idx = r["fv"]
    assert list(node) == [1, 2] and list(off) == [0, 4, 5] and list(idx) == [0, 1, 3, 5, 2]


def test_distinctive_descriptor_median_rule(oracle_mod):
    # N=4: median index = int(0.5*3) = 1 -> second smallest of each row (the row includes the 0 self-distance)
    d = _desc_with_bits([0, 10, 30, 100])
    # rows: [0,10,30,100] -> 10 ; [10,0,20,90] -> 10 ; [30,20,0,70] -> 20 ; [100,90,70,0] -> 70 ; first min wins
    assert oracle_mod.distinctive_descriptor(d) == 0
    d = _desc_with_bits([0, 30, 40, 100])
    # rows: 30 ; [30,0,10,70] -> 10 ; [40,10,0,60] -> 10 ; 60 -> first of the two 10s
    assert oracle_mod.distinctive_descriptor(d) == 1
    assert oracle_mod.distinctive_descriptor(d[:1]) == 0


def test_fuse_candidates_gates(oracle_mod):
    # three level-1 keypoints near the projection (100,100): at 1 px, 2 px and 3 px; invSigma2[1] = 1/1.44
    xy = np.array([[101, 100], [102, 100], [103, 100], [100.5, 100]], np.float32)
    k = _kps(4, octave=1, xy=xy)
    k["octave"][3] = 3                      # closest, but outside [l-1, l]
    d = _desc_with_bits([30, 10, 5, 0])
    inv = (1.0 / (np.array([1.2 ** i for i in range(8)], np.float32) ** 2)).astype(np.float32)
    q = dict(qx=[100.0], qy=[100.0], qr=[10.0], qlevel=[1], qdesc=_desc_with_bits([0]))
    # mono: e2*inv = 1/1.44, 4/1.44 = 2.78, 9/1.44 = 6.25 > 5.99 -> third dropped -> best = index 1 (dist 10)
    bi, bd = oracle_mod.fuse_candidates(k, d, (0, 752, 0, 480), inv, **q)
    assert bi[0] == 1 and bd[0] == 10
    # stereo coordinate on index 1 with er = 3: (4+9)/1.44 = 9.03 > 7.8 -> dropped -> index 0
    ur = np.array([-1, 50, -1, -1], np.float32)
    bi, bd = oracle_mod.fuse_candidates(k, d, (0, 752, 0, 480), inv, u_right=ur, q_ur=[53.0], **q)
    assert bi[0] == 0 and bd[0] == 30
    # predicted level 2 accepts octave 1 and 2; level 3 accepts the octave-3 keypoint only
    q["qlevel"] = [3]
    bi, bd = oracle_mod.fuse_candidates(k, d, (0, 752, 0, 480), inv, **q)
    assert bi[0] == 3 and bd[0] == 0


def test_search_for_initialization_stealing(oracle_mod):
    # two F1 keypoints compete for one F2 feature: the second (closer descriptor) steals it, the first ends unmatched
    k1 = _kps(2, xy=np.array([[100, 100], [105, 100]], np.float32))
    k2 = _kps(1, xy=np.array([[102, 100]], np.float32))
    d1 = _desc_with_bits([20, 5])
    d2 = _desc_with_bits([0])
    pm = np.stack([k1["x"], k1["y"]], 1)
    m12, nm, pm2 = oracle_mod.search_for_initialization(k1, d1, k2, d2, (0, 752, 0, 480), pm, 100, 0.9, False)
    assert list(m12) == [-1, 0] and nm == 1
    assert np.array_equal(pm2, np.array([[100, 100], [102, 100]], np.float32))
    # reversed descriptor quality: the later, worse one does not steal (vMatchedDistance <= dist)
    m12, nm, _ = oracle_mod.search_for_initialization(k1, d1[::-1].copy(), k2, d2, (0, 752, 0, 480), pm, 100, 0.9, False)
    assert list(m12) == [0, -1] and nm == 1
    # only level-0 keypoints of F1 take part, only level-0 features of F2 are candidates
    k1b = k1.copy(); k1b["octave"][1] = 1
    m12, nm, _ = oracle_mod.search_for_initialization(k1b, d1, k2, d2, (0, 752, 0, 480), pm, 100, 0.9, False)
    assert list(m12) == [0, -1]
    k2b = k2.copy(); k2b["octave"][0] = 1
    assert oracle_mod.search_for_initialization(k1, d1, k2b, d2, (0, 752, 0, 480), pm, 100, 0.9, False)[1] == 0
