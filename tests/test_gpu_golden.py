"""The CUDA path (through the C ABI) against the committed golden vectors — no oracle involved."""
from pathlib import Path

import numpy as np
import pytest

from orb_slam3_vio_fixes_b200.extractor import ORBextractor
from orb_slam3_vio_fixes_b200.matcher import ORBmatcher, results_to_numpy

pytestmark = pytest.mark.gpu
G = Path(__file__).resolve().parent / "golden"


def _t(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.parametrize("name", ["c1_mono.npz", "c3_rgbd.npz", "small_lap.npz"])
def test_extract_matches_golden(name):
    g = np.load(G / name)
    ex = ORBextractor(int(g["nfeatures"]), 1.2, 8, 20, 7)
    lap = tuple(int(v) for v in g["lap"])
    for i, img in enumerate(g["frames"]):
        mono, k, d = ex(img, None, lap)
        assert mono == int(g[f"mono{i}"])
        assert np.array_equal(k, g[f"kps{i}"]) and np.array_equal(d, g[f"desc{i}"])
        if i == 0:
            assert np.array_equal(ex.GetPyramid(3, with_border=True), g["pyr3_border"])
            assert np.array_equal(ex.GetPyramid(7), g["pyr7"])
            for level in (0, 4):
                assert np.array_equal(ex.debug_level_keypoints(level, 0), g[f"cand{level}"])


def test_strict_fp_flag_changes_nothing_else():
    """ORBX_FLAG_STRICT_FP only affects the descriptor sampler (and there only on exact .5 ties)."""
    g = np.load(G / "c3_rgbd.npz")
    ex = ORBextractor(1000, 1.2, 8, 20, 7, strict_fp=True)
    mono, k, d = ex(g["frames"][0], None, (0, 0))
    assert np.array_equal(k, g["kps0"])
    assert np.unpackbits(d ^ g["desc0"]).sum() <= 8


def test_match_and_stereo_match_golden():
    import torch
    from bench import shifts_for, MATCH_TH
    g = np.load(G / "c1_mono.npz")
    ex = ORBextractor(1000, 1.2, 8, 20, 7, max_batch=2)
    out = ex.extract_batch_device(_t(g["frames"]), (0, 1000))
    bounds = (0.0, 752.0, 0.0, 480.0)
    grid = ORBmatcher.grid_build(out["kps"], out["n"], bounds)
    qb = ORBmatcher.queries_shift(out["kps"], out["desc"], out["n"], _t(np.array([1, 0], np.int32)),
                                  _t(shifts_for(0, 2)), MATCH_TH, g["scale_factors"], (-1, 1))
    res = results_to_numpy(ORBmatcher.search_area(out["kps"], out["desc"], out["n"], grid, bounds, qb["qx"], qb["qy"],
                                                  qb["qr"], qb["qmin"], qb["qmax"], qb["qdesc"], qb["nq"]))
    torch.cuda.synchronize()
    for f in range(2):
        ref = g[f"match{f}"]
        assert np.array_equal(res[f, :len(ref)], ref)
    s = np.load(G / "c2_stereo.npz")
    exl, exr = ORBextractor(1200, 1.2, 8, 20, 7), ORBextractor(1200, 1.2, 8, 20, 7)
    left_t, right_t = _t(s["left"]), _t(s["right"])  # level 0 of the device pyramids IS the caller's buffer: keep alive
    ol = exl.extract_batch_device(left_t, (0, 0))
    orr = exr.extract_batch_device(right_t, (0, 0))
    bi, bd = ORBmatcher.stereo_rowband(ol["kps"], ol["desc"], ol["n"], orr["kps"], orr["desc"], orr["n"],
                                       s["scale_factors"], np.float32(s["bf"]) / np.float32(s["mb"]))
    n = int(ol["n"][0])
    assert n == len(s["kps_l"])
    assert np.array_equal(bi[0, :n].cpu().numpy(), s["best_idx_r"]) and np.array_equal(bd[0, :n].cpu().numpy(), s["best_dist"])
    # sub-pixel stage + median cut (Frame.cc:896-981) on the device-resident pyramids
    res = ORBmatcher.stereo_refine(exl, exr, ol["kps"], ol["n"], orr["kps"], bi, bd, np.float32(s["bf"]),
                                   np.float32(s["bf"]) / np.float32(s["mb"]))
    assert np.array_equal(res["u_right"][0, :n].cpu().numpy().view(np.uint32), s["u_right"].view(np.uint32))
    assert np.array_equal(res["depth"][0, :n].cpu().numpy().view(np.uint32), s["depth"].view(np.uint32))
    assert int(res["kept"][0]) == int(s["kept"])


def test_bow_functions_match_golden():
    """CUDA path vs tests/golden/bow_c1.npz: DBoW2 transform (fp64 BowVector bits), SearchByBoW x2,
    SearchForTriangulation, ComputeDistinctiveDescriptors — no oracle involved."""
    import torch
    from orb_slam3_vio_fixes_b200 import synthetic as S
    from orb_slam3_vio_fixes_b200 import _lib
    from orb_slam3_vio_fixes_b200.matcher import ORBVocabulary
    g, b = np.load(G / "c1_mono.npz"), np.load(G / "bow_c1.npz")
    feats = [(g["kps0"], g["desc0"]), (g["kps1"], g["desc1"])]
    cap = max(len(k) for k, _ in feats) + 8
    voc = S.make_vocabulary(int(b["voc_k"]), int(b["voc_L"]), seed=int(b["voc_seed"]),
                            desc_pool=np.concatenate([feats[0][1], feats[1][1]]))
    V = ORBVocabulary(voc["child_off"], voc["child_idx"], voc["node_desc"], voc["node_word"], voc["node_weight"], voc["L"])
    kps = np.zeros((2, cap), _lib.KP_DTYPE)
    desc = np.zeros((2, cap, 32), np.uint8)
    n = np.array([len(k) for k, _ in feats], np.int32)
    flags = np.zeros((2, cap), np.uint8)
    ur = np.full((2, cap), -1, np.float32)
    for i, (k, d) in enumerate(feats):
        kps[i, :n[i]], desc[i, :n[i]] = k, d
    flags[0, :n[0]], flags[1, :n[1]] = b["valid_a"], b["valid_b"]
    ur[0, :n[0]], ur[1, :n[1]] = b["uright_a"], b["uright_b"]
    kt = _t(kps.view(np.float32).reshape(2, cap, 7))
    tr = {}
    for levelsup in (2, 3):
        tr[levelsup] = V.transform(_t(desc), _t(n), levelsup=levelsup)
        torch.cuda.synchronize()
        for i, tag in enumerate("ab"):
            o = tr[levelsup]
            assert np.array_equal(o["feat_word"][i, :n[i]].cpu().numpy(), b[f"word_{tag}_{levelsup}"])
            assert np.array_equal(o["feat_node"][i, :n[i]].cpu().numpy(), b[f"node_{tag}_{levelsup}"])
            nb = int(o["bow_n"][i])
            assert np.array_equal(o["bow_word"][i, :nb].cpu().numpy(), b[f"bow_word_{tag}_{levelsup}"])
            assert np.array_equal(o["bow_val"][i, :nb].cpu().numpy(), b[f"bow_val_{tag}_{levelsup}"])
    fv = tr[2]["fv"]
    side = lambda i: (kt[i:i + 1], _t(desc[i:i + 1]), _t(n[i:i + 1]), {k_: v[i:i + 1].contiguous() for k_, v in fv.items()})
    (ka, da, na, fa), (kb, db, nb_, fb) = side(0), side(1)
    for mode in (0, 1):
        for ratio, ori in ((0.6, True), (0.9, False)):
            key = f"bow{mode}_{int(ratio * 100)}_{int(ori)}"
            r = ORBmatcher.search_by_bow(mode, ka, da, na, fa, kb, db, nb_, fb, ratio, ori, _t(flags[0:1]), _t(flags[1:2]))
            assert int(r["nmatches"][0]) == int(b[key + "_n"])
            assert np.array_equal(r["match12"][0, :n[0]].cpu().numpy(), b[key + "_m12"])
            assert np.array_equal(r["match21"][0, :n[1]].cpu().numpy(), b[key + "_m21"])
    for only_stereo, coarse in ((0, 0), (1, 0), (0, 1)):
        r = ORBmatcher.search_for_triangulation(ka, da, na, fa, kb, db, nb_, fb, _t(b["F12"][None]), _t(b["ep"][None]),
                                                b["scale"], b["sigma2"], only_stereo, coarse, True, _t(1 - flags[0:1]),
                                                _t(1 - flags[1:2]), _t(ur[0:1]), _t(ur[1:2]))
        assert int(r["nmatches"][0]) == int(b[f"tri_{only_stereo}{coarse}_n"])
        assert np.array_equal(r["match12"][0, :n[0]].cpu().numpy(), b[f"tri_{only_stereo}{coarse}_m12"])
    best, _ = ORBmatcher.distinctive_descriptors(_t(feats[0][1]), _t(b["dd_off"]), _t(b["dd_idx"]))
    assert np.array_equal(best.cpu().numpy(), b["dd_best"])
