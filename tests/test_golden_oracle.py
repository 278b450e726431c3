"""The oracle against the committed golden vectors (tests/golden, generated where the oracle and
the cv2-based pipeline agreed).  Runs without cv2 and without /root/reference."""
from pathlib import Path

import numpy as np
import pytest

G = Path(__file__).resolve().parent / "golden"


def _load(name):
    return np.load(G / name, allow_pickle=False)


@pytest.mark.parametrize("name", ["c1_mono.npz", "c3_rgbd.npz", "small_lap.npz"])
def test_extract_matches_golden(oracle_mod, name):
    g = _load(name)
    assert bool(g["pinned_by_cv2"])
    orc = oracle_mod.Extractor(int(g["nfeatures"]))
    lap = tuple(int(v) for v in g["lap"])
    for i, img in enumerate(g["frames"]):
        k, d, m = orc.extract(img, lap)
        assert m == int(g[f"mono{i}"])
        assert np.array_equal(k, g[f"kps{i}"]) and np.array_equal(d, g[f"desc{i}"])
        if i == 0:
            assert np.array_equal(orc.pyramid(3, True), g["pyr3_border"])
            assert np.array_equal(orc.pyramid(7), g["pyr7"])
            for level in (0, 4):
                c = orc.level_keypoints(level, 0)
                assert np.array_equal(np.stack([c["x"], c["y"], c["response"]], 1).astype(np.int32), g[f"cand{level}"])


def test_match_matches_golden(oracle_mod):
    from bench import shifts_for, MATCH_TH
    g = _load("c1_mono.npz")
    sh = shifts_for(0, 2)
    for f in range(2):
        s = (f - 1) % 2
        kq, kr = g[f"kps{s}"], g[f"kps{f}"]
        qr = (np.float32(MATCH_TH) * g["scale_factors"][kq["octave"]]).astype(np.float32)
        off, idx = oracle_mod.grid_queries(kr, (0.0, 752.0, 0.0, 480.0), kq["x"] + sh[f, 0], kq["y"] + sh[f, 1], qr,
                                           kq["octave"] - 1, kq["octave"] + 1)
        assert len(idx) == int(g[f"match{f}_ncand"])
        res = oracle_mod.argmin2_csr(g[f"desc{s}"], g[f"desc{f}"], off, idx, kr["octave"].astype(np.int8), 0)
        assert np.array_equal(res, g[f"match{f}"])


def test_stereo_matches_golden(oracle_mod):
    g = _load("c2_stereo.npz")
    ol, orr = oracle_mod.Extractor(1200), oracle_mod.Extractor(1200)
    kl, dl, _ = ol.extract(g["left"][0], (0, 0))
    kr, dr, _ = orr.extract(g["right"][0], (0, 0))
    assert np.array_equal(kl, g["kps_l"]) and np.array_equal(dr, g["desc_r"])
    pl = [ol.pyramid(l) for l in range(8)]
    pr = [orr.pyramid(l) for l in range(8)]
    bi, bd, ur, dp, kept = oracle_mod.stereo_matches(kl, dl, kr, dr, ol.scale_factors, ol.inv_scale_factors, pl, pr,
                                                     g["mb"], g["bf"])
    assert np.array_equal(bi, g["best_idx_r"]) and np.array_equal(bd, g["best_dist"])
    assert np.array_equal(ur, g["u_right"]) and np.array_equal(dp, g["depth"]) and kept == int(g["kept"])


def test_bow_functions_match_golden(oracle_mod):
    """tests/golden/bow_c1.npz (tests/golden/make_golden_bow.py): written only where the oracle and an independent
    pure-Python restatement of the reference loops agreed."""
    from orb_slam3_vio_fixes_b200 import synthetic as S
    g, b = _load("c1_mono.npz"), _load("bow_c1.npz")
    assert bool(b["pinned_by_python_restatement"])
    k1, d1, k2, d2 = g["kps0"], g["desc0"], g["kps1"], g["desc1"]
    voc = S.make_vocabulary(int(b["voc_k"]), int(b["voc_L"]), seed=int(b["voc_seed"]), desc_pool=np.concatenate([d1, d2]))
    fv = {}
    for levelsup in (2, 3):
        for tag, d in (("a", d1), ("b", d2)):
            r = oracle_mod.bow_transform(d, voc, levelsup)
            assert np.array_equal(r["feat_word"], b[f"word_{tag}_{levelsup}"])
            assert np.array_equal(r["feat_node"], b[f"node_{tag}_{levelsup}"])
            assert np.array_equal(r["bow_word"], b[f"bow_word_{tag}_{levelsup}"])
            assert np.array_equal(r["bow_val"], b[f"bow_val_{tag}_{levelsup}"])
            fv[(levelsup, tag)] = r["fv"]
    ca, cb = fv[(2, "a")], fv[(2, "b")]
    v1, v2, u1, u2 = b["valid_a"], b["valid_b"], b["uright_a"], b["uright_b"]
    for mode in (0, 1):
        for ratio, ori in ((0.6, True), (0.9, False)):
            key = f"bow{mode}_{int(ratio * 100)}_{int(ori)}"
            m12, m21, n = oracle_mod.search_by_bow(mode, k1, d1, ca, k2, d2, cb, ratio, ori, v1, v2)
            assert n == int(b[key + "_n"]) and np.array_equal(m12, b[key + "_m12"]) and np.array_equal(m21, b[key + "_m21"])
    for only_stereo, coarse in ((0, 0), (1, 0), (0, 1)):
        m12, n = oracle_mod.search_for_triangulation(k1, d1, ca, k2, d2, cb, b["F12"], b["ep"], b["scale"], b["sigma2"],
                                                     only_stereo, coarse, True, 1 - v1, 1 - v2, u1, u2)
        assert n == int(b[f"tri_{only_stereo}{coarse}_n"]) and np.array_equal(m12, b[f"tri_{only_stereo}{coarse}_m12"])
    off, idx = b["dd_off"], b["dd_idx"]
    for m in range(len(off) - 1):
        assert oracle_mod.distinctive_descriptor(d1[idx[off[m]:off[m + 1]]]) == int(b["dd_best"][m])
