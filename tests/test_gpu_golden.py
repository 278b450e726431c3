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
