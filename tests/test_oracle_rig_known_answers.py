"""Hand-derived known answers for the oracle's restatements of the two-camera (fisheye rig) matcher branches
(oracle/frame_oracle.cpp).  Each case is small enough to follow src/ORBmatcher.cc by hand; the comments name the
reference lines that decide the outcome."""
import numpy as np
import pytest


def bits(k):
    """A 32-byte descriptor with the k lowest bits set: DescriptorDistance(bits(a), bits(b)) = |a - b|."""
    d = np.zeros(32, np.uint8)
    for i in range(k):
        d[i // 8] |= 1 << (i % 8)
    return d


def kp(oracle_mod, xy_oct_ang):
    k = np.zeros(len(xy_oct_ang), oracle_mod.KP_DTYPE)
    for i, (x, y, o, a) in enumerate(xy_oct_ang):
        k[i] = (x, y, 31, a, 50, o, -1)
    return k


BOUNDS = (0.0, 640.0, 0.0, 480.0)


def q_of(rows):
    """rows: (xl, yl, xr, yr, level, desc_bits, flags, angle); windows of radius 10, band [level-1, level]."""
    n = len(rows)
    q = dict(qx=np.array([r[0] for r in rows], np.float32), qy=np.array([r[1] for r in rows], np.float32),
             qr=np.full(n, 10, np.float32), qmin=np.array([r[4] - 1 for r in rows], np.int8),
             qmax=np.array([r[4] for r in rows], np.int8),
             qx_r=np.array([r[2] for r in rows], np.float32), qy_r=np.array([r[3] for r in rows], np.float32),
             qr_r=np.full(n, 10, np.float32), qmin_r=np.array([r[4] - 1 for r in rows], np.int8),
             qmax_r=np.array([r[4] for r in rows], np.int8),
             qdesc=np.stack([bits(r[5]) for r in rows]), flags=np.array([r[6] for r in rows], np.uint8),
             angle=np.array([r[7] for r in rows], np.float32))
    return q


def test_mappoints_left_ratio_failure_skips_right(oracle_mod):
    """Two left candidates on the same level at distances 10 and 11: 10 > 0.8*11 fails the ratio test and the
    `continue` at ORBmatcher.cc:126-127 leaves the loop body, so the right camera (:144-210) is never searched —
    although its only candidate is a perfect match."""
    kl = kp(oracle_mod, [(100, 100, 1, 0), (103, 100, 1, 0)])
    dl = np.stack([bits(10), bits(11)])
    kr = kp(oracle_mod, [(80, 100, 1, 0)])
    dr = np.stack([bits(0)])
    q = q_of([(100, 100, 80, 100, 1, 0, 7, 0)])
    o = oracle_mod.search_by_projection_rig(0, kl, dl, kr, dr, BOUNDS, q, nn_ratio=0.8)
    assert o["nmatches"] == 0 and o["match_l"][0] == -1 and o["match_r"][0] == -1
    # same point with the left query switched off (mbTrackInView false): now the right block runs and matches
    q["flags"][:] = 6
    o = oracle_mod.search_by_projection_rig(0, kl, dl, kr, dr, BOUNDS, q, nn_ratio=0.8)
    assert o["nmatches"] == 1 and o["match_r"][0] == 0 and o["feat_query_r"][0] == 0


def test_mappoints_pair_propagation_locks_and_unlocks(oracle_mod):
    """Query 0 (observations > 0) matches left feature 0, whose mvLeftToRightMatch is right feature 0: both are
    written (nmatches += 2, :129-139) and both are now locked (:88-90, :169-171).  Query 1 would match right feature
    0 best (distance 0) but finds it locked and takes right feature 1 (distance 5) instead.  Query 2 has no
    observations; it matches left feature 1 whose pair is right feature 1: the overwrite (:131-135) UNLOCKS right
    feature 1, so query 3 can take it again."""
    kl = kp(oracle_mod, [(100, 100, 1, 0), (300, 100, 1, 0)])
    dl = np.stack([bits(0), bits(40)])
    kr = kp(oracle_mod, [(80, 100, 1, 0), (85, 100, 2, 0)])
    dr = np.stack([bits(0), bits(5)])
    l2r = np.array([0, 1], np.int32)
    r2l = np.array([-1, -1], np.int32)
    q = q_of([(100, 100, 500, 400, 1, 0, 3, 0),      # left only -> left 0 (+ right 0 through the pair)
              (600, 400, 80, 100, 2, 0, 6, 0),       # right only: right 0 locked -> right 1 (level 2)
              (300, 100, 500, 400, 1, 40, 1, 0),     # left only, no observations -> left 1, pair right 1 unlocked
              (600, 400, 85, 100, 2, 5, 6, 0)])      # right only: right 1 is free again
    o = oracle_mod.search_by_projection_rig(0, kl, dl, kr, dr, BOUNDS, q, nn_ratio=0.8, l2r=l2r, r2l=r2l)
    assert o["match_l"].tolist() == [0, -1, 1, -1]
    assert o["match_r"].tolist() == [-1, 1, -1, 1]
    assert o["feat_query_l"].tolist() == [0, 2]
    assert o["feat_query_r"].tolist() == [0, 3]
    assert o["nmatches"] == 2 + 1 + 2 + 1


def test_mappoints_right_match_writes_left_pair_first(oracle_mod):
    """mvRightToLeftMatch propagation (:197-201): the paired LEFT feature is overwritten even though an earlier query
    holds it."""
    kl = kp(oracle_mod, [(100, 100, 1, 0)])
    dl = np.stack([bits(0)])
    kr = kp(oracle_mod, [(80, 100, 1, 0)])
    dr = np.stack([bits(30)])
    q = q_of([(100, 100, 500, 400, 1, 0, 3, 0),      # left 0
              (600, 400, 80, 100, 1, 30, 6, 0)])     # right 0, whose pair is left 0
    o = oracle_mod.search_by_projection_rig(0, kl, dl, kr, dr, BOUNDS, q, r2l=np.array([0], np.int32))
    assert o["feat_query_l"].tolist() == [1] and o["feat_query_r"].tolist() == [1] and o["nmatches"] == 3


def test_lastframe_empty_left_window_skips_right_and_joint_histogram(oracle_mod):
    """Last-frame variant: query 0 has no left candidate in its window (`continue` at :1737-1738), so the right
    camera is not searched.  Queries 1..12 match left and right with rotation 0 (bin 0, 24 entries); query 13 matches
    only on the right with rotation 90 deg (bin 3): 1 < 0.1*24, so ComputeThreeMaxima drops bin 3 (:2041-2049) and the
    match is erased from the ONE histogram both cameras share (:1864-1881)."""
    rows, lk, rk, ld, rd = [], [], [], [], []
    rows.append((50, 50, 30, 50, 1, 0, 3, 0))  # nothing near (50,50) on the left; a perfect match sits on the right
    rk.append((30, 50, 1, 0)); rd.append(bits(0))
    for i in range(12):
        x = 100 + 40 * i
        lk.append((x, 200, 1, 0)); ld.append(bits(60 + i))
        rk.append((x - 20, 200, 1, 0)); rd.append(bits(60 + i))
        rows.append((x, 200, x - 20, 200, 1, 60 + i, 3, 0))
    lk.append((600, 400, 1, 0)); ld.append(bits(200))          # left candidate too far in Hamming distance (> TH_HIGH)
    rk.append((580, 400, 1, 270)); rd.append(bits(90))
    rows.append((600, 400, 580, 400, 1, 90, 3, 0))             # rot = 0 - 270 + 360 = 90 -> bin 3
    kl, kr = kp(oracle_mod, lk), kp(oracle_mod, rk)
    q = q_of(rows)
    o = oracle_mod.search_by_projection_rig(1, kl, np.stack(ld), kr, np.stack(rd), BOUNDS, q, check_orientation=True)
    assert o["match_l"][0] == -1 and o["match_r"][0] == -1 and o["feat_query_r"][0] == -1
    assert o["match_l"][1:13].tolist() == list(range(12)) and o["match_r"][1:13].tolist() == list(range(1, 13))
    assert o["match_r"][13] == -1 and o["feat_query_r"][13] == -1
    assert o["nmatches"] == 24
    o2 = oracle_mod.search_by_projection_rig(1, kl, np.stack(ld), kr, np.stack(rd), BOUNDS, q, check_orientation=False)
    assert o2["match_r"][13] == 13 and o2["nmatches"] == 25


def test_bow_rig_left_and_right_rules(oracle_mod):
    """SearchByBoW(KeyFrame*, Frame&) with Nleft = 2 (frame features 0,1 left; 2,3 right), one vocabulary node:
      kf 0: left distances (3, 4) fail 3 < 0.6*4, yet the right best (2 <= TH_LOW) is still written (`|| true`, :358);
      kf 1: left best 57 > TH_LOW: nothing is written although the right best is 7 (:327 guards both blocks);
      kf 2: left (41, 42) fails the ratio too; right feature 2 is held by kf 0, so feature 3 (distance 9) is written."""
    k1 = kp(oracle_mod, [(0, 0, 0, 0)] * 3)
    d1 = np.stack([bits(100), bits(161), bits(145)])
    k2 = kp(oracle_mod, [(0, 0, 0, 0)] * 4)
    d2 = np.stack([bits(103), bits(104), bits(102), bits(154)])  # left: 103, 104; right: 102, 154
    fv1 = (np.array([7], np.int32), np.array([0, 3], np.int32), np.array([0, 1, 2], np.int32))
    fv2 = (np.array([7], np.int32), np.array([0, 4], np.int32), np.array([0, 1, 2, 3], np.int32))
    m12l, m12r, m21, nm = oracle_mod.search_by_bow_rig(k1, d1, fv1, k2, d2, 2, fv2, nn_ratio=0.6,
                                                       check_orientation=False)
    assert m12l.tolist() == [-1, -1, -1]          # kf2: 41 < 0.6*42 is false as well
    assert m12r.tolist() == [2, -1, 3]            # kf2 right: feature 2 is held by kf 0 -> feature 3 at |145-154| = 9
    assert m21.tolist() == [-1, -1, 0, 2] and nm == 2
    # with a looser ratio kf 0 also takes left feature 0 (3 < 0.9*4); for kf 2 feature 0 is then gone, feature 1 is
    # the only left candidate (41 < 0.9*256) and is written too
    m12l, m12r, m21, nm = oracle_mod.search_by_bow_rig(k1, d1, fv1, k2, d2, 2, fv2, nn_ratio=0.9,
                                                       check_orientation=False)
    assert m12l.tolist() == [0, -1, 1] and m12r.tolist() == [2, -1, 3] and m21.tolist() == [0, 2, 0, 2] and nm == 4
