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
