"""The bench workloads at BASELINE.json's FULL sizes (C1: 512 frames of 752x480 per step, C5: 32 frames of 3840x2160),
checked through properties that do not depend on the size, because the oracle needs ~20 ms per frame:

  * partition invariance — frames are independent units, so the keypoints / descriptors / match records of frame f must
    be the same bytes whether f is extracted inside the 512-frame step (two sub-batches on two internal streams) or in a
    16-frame call of its own; every frame of the step is covered (a checksum per frame, then a checksum of checksums);
  * idempotence — the same step twice gives the same bytes (scratch reuse, the static work lists, stream rotation);
  * a spot check of a few frames of the step against the oracle, and of the match records against the oracle's window
    search, so the invariants above are anchored to the reference's results and not only to each other."""
import hashlib

import numpy as np
import pytest

from conftest import has_cuda
from orb_slam3_vio_fixes_b200 import synthetic as S

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not has_cuda(), reason="needs CUDA")]

ORB = dict(scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7)


def _digest(*arrays):
    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.digest()


def _frame_digests(out, res=None):
    """One checksum per frame over exactly the valid records (count, mono index, keypoints, descriptors, matches)."""
    n = out["n"].cpu().numpy()
    mono = out["mono"].cpu().numpy()
    kps = out["kps"].cpu().numpy()
    desc = out["desc"].cpu().numpy()
    r = None if res is None else res.cpu().numpy()   # raw record bytes
    return [_digest(n[f:f + 1], mono[f:f + 1], kps[f, :n[f]], desc[f, :n[f]], *([] if r is None else [r[f, :n[f]]]))
            for f in range(len(n))], n


def _step(ex, frames_dev, lap, bounds, t0, qb=None):
    """One bench step: extract + frame-to-frame window match (bench.py's device-resident step)."""
    import torch
    from bench import shifts_for, MATCH_TH
    from orb_slam3_vio_fixes_b200.matcher import ORBmatcher
    B = frames_dev.shape[0]
    out = ex.extract_batch_device(frames_dev, lap)
    sf = ex.GetScaleFactors()
    shift = torch.from_numpy(shifts_for(t0, B)).cuda()
    src = torch.from_numpy(((np.arange(B) - 1) % B).astype(np.int32)).cuda()   # frame f is searched with f-1's keypoints
    grid = ORBmatcher.grid_build(out["kps"], out["n"], bounds)
    q = ORBmatcher.queries_shift(out["kps"], out["desc"], out["n"], src, shift, MATCH_TH, sf, (-1, 1), qb)
    res = ORBmatcher.search_area(out["kps"], out["desc"], out["n"], grid, bounds, q["qx"], q["qy"], q["qr"], q["qmin"],
                                 q["qmax"], q["qdesc"], q["nq"])
    torch.cuda.synchronize()
    return out, res


@pytest.mark.parametrize("cfg,B,sub,spot", [("C1", 512, 16, (0, 255, 256, 511)), ("C5", 32, 4, (17,))])
def test_full_step_partition_invariance_and_idempotence(oracle_mod, cfg, B, sub, spot):
    import torch
    from orb_slam3_vio_fixes_b200.extractor import ORBextractor
    w, h, nf, _, _ = S.CONFIGS[cfg]
    lap = (0, 1000)
    frames = S.make_frames(cfg, B)
    dev = torch.from_numpy(frames).cuda()
    ex = ORBextractor(nf, max_batch=B, **ORB)
    out, _ = _step(ex, dev, lap, (0.0, float(w), 0.0, float(h)), 0)
    whole, n = _frame_digests(out)
    assert (n > 0.8 * nf).all()

    # idempotence: the same step again, same bytes
    out2, _ = _step(ex, dev, lap, (0.0, float(w), 0.0, float(h)), 0)
    again, _ = _frame_digests(out2)
    assert again == whole

    # partition invariance: every frame again, `sub` frames per call on a second (smaller) handle
    exs = ORBextractor(nf, max_batch=sub, **ORB)
    parts = []
    for f0 in range(0, B, sub):
        o = exs.extract_batch_device(dev[f0:f0 + sub], lap)
        torch.cuda.synchronize()
        parts += _frame_digests(o)[0]
    assert _digest(*parts) == _digest(*whole), [f for f in range(B) if parts[f] != whole[f]][:8]

    # anchor: a few frames of the full step against the oracle
    orc = oracle_mod.Extractor(nf, 1.2, 8, 20, 7)
    kps = ORBextractor.kps_from_tensor(out["kps"])
    desc = out["desc"].cpu().numpy()
    for f in spot:
        ok, od, om = orc.extract(frames[f], lap)
        assert n[f] == len(ok) and out["mono"][f].item() == om
        assert np.array_equal(kps[f, :n[f]], ok) and np.array_equal(desc[f, :n[f]], od)


def test_full_step_match_records_partition_invariance(oracle_mod):
    """The C1 step's match stage at 512 frames: the records of frame f (searched with the keypoints of frame f-1) are
    the same bytes as in a 32-frame call that holds frames f-1 and f, and equal the oracle's window search."""
    import torch
    from bench import shifts_for, MATCH_TH
    from orb_slam3_vio_fixes_b200.extractor import ORBextractor
    from orb_slam3_vio_fixes_b200.matcher import results_to_numpy
    from test_gpu_match import _same_results
    B, sub = 512, 32
    bounds = (0.0, 752.0, 0.0, 480.0)
    frames = S.make_frames("C1", B)
    dev = torch.from_numpy(frames).cuda()
    ex = ORBextractor(1000, max_batch=B, **ORB)
    out, res = _step(ex, dev, (0, 0), bounds, 0)
    n = out["n"].cpu().numpy()
    r_all = results_to_numpy(res)
    exs = ORBextractor(1000, max_batch=sub, **ORB)
    checked = 0
    for f0 in range(0, B, sub - 1):                      # windows overlap by one frame: every f >= 1 has its f-1 inside
        f1 = min(f0 + sub, B)
        o, r = _step(exs, dev[f0:f1], (0, 0), bounds, f0)
        rr = results_to_numpy(r)
        for g in range(1, f1 - f0):
            f = f0 + g
            assert np.array_equal(rr[g, :n[f - 1]], r_all[f, :n[f - 1]]), f
            checked += 1
    assert checked >= B - 1
    # anchor two frames to the oracle's GetFeaturesInArea + best / second-best loop
    kps = ORBextractor.kps_from_tensor(out["kps"])
    desc = out["desc"].cpu().numpy()
    sf = ex.GetScaleFactors()
    sh = shifts_for(0, B)
    for f in (1, 300):
        kq, kr = kps[f - 1, :n[f - 1]], kps[f, :n[f]]
        qr = (np.float32(MATCH_TH) * sf[kq["octave"]]).astype(np.float32)
        off, idx = oracle_mod.grid_queries(kr, bounds, kq["x"] + sh[f, 0], kq["y"] + sh[f, 1], qr, kq["octave"] - 1,
                                           kq["octave"] + 1)
        want = oracle_mod.argmin2_csr(desc[f - 1, :n[f - 1]], desc[f, :n[f]], off, idx, kr["octave"].astype(np.int8), 0)
        _same_results(r_all[f, :n[f - 1]], want, with_second=True)


@pytest.mark.parametrize("cfg,B,sub", [("C2", 256, 16), ("C4", 128, 8)])
def test_full_stereo_step_partition_invariance(oracle_mod, cfg, B, sub):
    """The stereo bench step at its full size (C2: 256 pairs, C4: 128 pairs): both eyes + the whole
    ComputeStereoMatches.  mvuRight / mvDepth of pair p are the same bytes inside the full step and in a `sub`-pair
    call; one pair of the step is compared with the oracle's ComputeStereoMatches."""
    import torch
    from bench import STEREO
    from orb_slam3_vio_fixes_b200.extractor import ORBextractor
    from orb_slam3_vio_fixes_b200.matcher import ORBmatcher
    w, h, nf, _, _ = S.CONFIGS[cfg]
    fx, bf = STEREO[cfg]
    mbf = np.float32(bf)
    mb = np.float32(bf / fx)
    max_d = mbf / mb
    left, right = S.make_frames(cfg, B), S.make_frames(cfg, B, right=True)
    dl, dr = torch.from_numpy(left).cuda(), torch.from_numpy(right).cuda()

    def run(exl, exr, a, b):
        ol, orr = exl.extract_batch_device(a, (0, 0)), exr.extract_batch_device(b, (0, 0))
        sf = exl.GetScaleFactors()
        bi, bd = ORBmatcher.stereo_rowband(ol["kps"], ol["desc"], ol["n"], orr["kps"], orr["desc"], orr["n"], sf, max_d)
        r = ORBmatcher.stereo_refine(exl, exr, ol["kps"], ol["n"], orr["kps"], bi, bd, mbf, max_d)
        torch.cuda.synchronize()
        n = ol["n"].cpu().numpy()
        ur, dp = r["u_right"].cpu().numpy(), r["depth"].cpu().numpy()
        return [_digest(n[p:p + 1], ur[p, :n[p]], dp[p, :n[p]]) for p in range(len(n))], (ol, orr, ur, dp, n)

    exl, exr = ORBextractor(nf, max_batch=B, **ORB), ORBextractor(nf, max_batch=B, **ORB)
    whole, (ol, orr, ur, dp, n) = run(exl, exr, dl, dr)
    assert run(exl, exr, dl, dr)[0] == whole                                    # idempotence
    sl, sr = ORBextractor(nf, max_batch=sub, **ORB), ORBextractor(nf, max_batch=sub, **ORB)
    parts = []
    for p0 in range(0, B, sub):
        parts += run(sl, sr, dl[p0:p0 + sub], dr[p0:p0 + sub])[0]
    assert _digest(*parts) == _digest(*whole), [p for p in range(B) if parts[p] != whole[p]][:8]
    # anchor: one pair of the full step against the oracle
    p = B // 2 + 1
    el, er = oracle_mod.Extractor(nf, 1.2, 8, 20, 7), oracle_mod.Extractor(nf, 1.2, 8, 20, 7)
    kl, ddl, _ = el.extract(left[p], (0, 0))
    kr, ddr, _ = er.extract(right[p], (0, 0))
    _, _, our, odp, kept = oracle_mod.stereo_matches(kl, ddl, kr, ddr, el.scale_factors, el.inv_scale_factors,
                                                     [el.pyramid(l) for l in range(8)], [er.pyramid(l) for l in range(8)],
                                                     mb, mbf)
    assert n[p] == len(kl) and kept > 100
    assert np.array_equal(ur[p, :n[p]], our) and np.array_equal(dp[p, :n[p]], odp)
