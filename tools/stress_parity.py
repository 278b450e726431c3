"""Randomised parity sweep on the GPU box: extraction (fused and level-blur modes), stereo matching and the
whole-function projection search against the oracle over random image sizes, feature budgets and content.
  python tools/stress_parity.py [seconds]"""
import sys
import time
import numpy as np
import torch
sys.path.insert(0, '.')
import oracle
from orb_slam3_vio_fixes_b200 import synthetic as S, _lib
from orb_slam3_vio_fixes_b200.extractor import ORBextractor
from orb_slam3_vio_fixes_b200.matcher import ORBmatcher

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 120.0
rng = np.random.default_rng(2026)
t0 = time.time()
cases = frames_done = kps_total = stereo_pairs = proj_frames = 0
while time.time() - t0 < budget:
    w = int(rng.integers(160, 1400)); h = int(rng.integers(140, 900))
    if w < h * 0.8:  # portrait levels with nIni == 0 are rejected at create (UB in the reference)
        w, h = h, w
    nf = int(rng.choice([50, 200, 500, 1000, 2000, 4000]))
    ini, mn = (20, 7) if rng.random() < 0.7 else (int(rng.integers(8, 40)), int(rng.integers(2, 8)))
    lap = (0, 0) if rng.random() < 0.5 else (int(rng.integers(0, w // 2)), int(rng.integers(w // 2, w + 50)))
    mode = [None, "fused", "level"][int(rng.integers(0, 3))]
    B = 2
    canvas = S.make_canvas(w, h, int(rng.integers(1, 1 << 30)))
    kind = rng.random()
    imgs = np.stack([S.make_frame(canvas, w, h, 7, t) for t in range(B)])
    imgs_r = np.stack([S.make_frame(canvas, w, h, 7, t, right=True) for t in range(B)])
    if kind < 0.15:   # heavy noise: many weak corners, threshold fallback everywhere
        imgs = rng.integers(0, 256, imgs.shape, dtype=np.uint8)
    elif kind < 0.25:  # low contrast
        imgs = (imgs // 8 + 100).astype(np.uint8)
    try:
        ex = ORBextractor(nf, 1.2, 8, ini, mn, max_batch=B, blur_mode=mode).plan(w, h)
    except Exception as e:  # unsupported geometry is rejected identically for every content
        continue
    ft = torch.from_numpy(imgs).cuda()
    out = ex.extract_batch_device(ft, lap)
    kps = ORBextractor.kps_from_tensor(out["kps"]); desc = out["desc"].cpu().numpy(); n = out["n"].cpu().numpy()
    mono = out["mono"].cpu().numpy()
    orc = oracle.Extractor(nf, 1.2, 8, ini, mn)
    for f in range(B):
        ok, od, om = orc.extract(imgs[f], lap)
        assert n[f] == len(ok) and mono[f] == om, (w, h, nf, ini, mn, lap, mode, n[f], len(ok))
        assert np.array_equal(kps[f, :n[f]], ok), (w, h, nf, ini, mn, lap, mode)
        assert np.array_equal(desc[f, :n[f]], od), (w, h, nf, ini, mn, lap, mode)
        frames_done += 1; kps_total += int(n[f])
    # stereo on the synthetic pair (only meaningful for the structured content)
    if kind >= 0.25 and lap == (0, 0):
        exr = ORBextractor(nf, 1.2, 8, ini, mn, max_batch=B, blur_mode=mode).plan(w, h)
        rt = torch.from_numpy(imgs_r).cuda()
        orr = exr.extract_batch_device(rt, (0, 0))
        sf, isf = ex.GetScaleFactors(), ex.GetInverseScaleFactors()
        mbf = np.float32(rng.uniform(20, 400)); mb = np.float32(mbf / np.float32(rng.uniform(100, 700)))
        max_d = mbf / mb
        bi, bd = ORBmatcher.stereo_rowband(out["kps"], out["desc"], out["n"], orr["kps"], orr["desc"], orr["n"], sf, max_d)
        res = ORBmatcher.stereo_refine(ex, exr, out["kps"], out["n"], orr["kps"], bi, bd, mbf, max_d)
        kr = ORBextractor.kps_from_tensor(orr["kps"]); dr = orr["desc"].cpu().numpy(); nr = orr["n"].cpu().numpy()
        for f in range(B):
            pl = [ex.GetPyramid(l, frame=f) for l in range(8)]
            pr = [exr.GetPyramid(l, frame=f) for l in range(8)]
            rbi, rbd, ur, dp, kept = oracle.stereo_matches(kps[f, :n[f]], desc[f, :n[f]], kr[f, :nr[f]], dr[f, :nr[f]],
                                                           sf, isf, pl, pr, mb, mbf)
            assert np.array_equal(bi[f, :n[f]].cpu().numpy(), rbi) and np.array_equal(bd[f, :n[f]].cpu().numpy(), rbd)
            assert np.array_equal(res["u_right"][f, :n[f]].cpu().numpy().view(np.uint32), ur.view(np.uint32)), (w, h, nf)
            assert np.array_equal(res["depth"][f, :n[f]].cpu().numpy().view(np.uint32), dp.view(np.uint32))
            assert int(res["kept"][f]) == kept
            stereo_pairs += 1
    # projection search: frame 1 <- frame 0 keypoints under the known shift, random flags / mode
    if n.min() > 0:
        bounds = (0.0, float(w), 0.0, float(h))
        grid = ORBmatcher.grid_build(out["kps"], out["n"], bounds)
        sf = ex.GetScaleFactors()
        pmode = int(rng.integers(0, 2)); th = float(rng.choice([7.0, 15.0, 30.0]))
        Q = int(n.max()) * 2
        qt = {k_: torch.zeros((B, Q), dtype=dt) for k_, dt in (("qx", torch.float32), ("qy", torch.float32), ("qr", torch.float32),
              ("qmin", torch.int8), ("qmax", torch.int8), ("angle", torch.float32), ("flags", torch.uint8), ("q_ur", torch.float32))}
        qt["qdesc"] = torch.zeros((B, Q, 32), dtype=torch.uint8); qt["nq"] = torch.zeros(B, dtype=torch.int32)
        qn = []
        for f in range(B):
            s_ = (f - 1) % B
            k = np.concatenate([kps[s_, :n[s_]]] * 2); d = np.concatenate([desc[s_, :n[s_]]] * 2)
            a, b = S.frame_shift(s_), S.frame_shift(f)
            q = {"qx": k["x"] + np.float32(a[0] - b[0]), "qy": k["y"] + np.float32(a[1] - b[1]),
                 "qr": (np.float32(th) * sf[k["octave"]]).astype(np.float32),
                 "qmin": (k["octave"] - 1).astype(np.int8), "qmax": (k["octave"] + (0 if pmode == 0 else 1)).astype(np.int8),
                 "angle": k["angle"].copy(), "flags": rng.choice([0, 1, 3, 3, 3], len(k)).astype(np.uint8),
                 "q_ur": (k["x"] - np.float32(24)).astype(np.float32), "qdesc": d}
            qn.append(q)
            for key in q:
                qt[key][f, :len(k)] = torch.from_numpy(np.ascontiguousarray(q[key]))
            qt["nq"][f] = len(k)
        qt = {k_: v.cuda() for k_, v in qt.items()}
        locked = (rng.random((B, kps.shape[1])) < 0.1).astype(np.uint8)
        u_r = np.where(rng.random((B, kps.shape[1])) < 0.5, kps["x"] - np.float32(24), np.float32(-1)).astype(np.float32)
        nn = float(rng.choice([0.6, 0.8, 0.9]))
        r = ORBmatcher.search_by_projection(out["kps"], out["desc"], out["n"], grid, bounds, qt, pmode, nn, True,
                                            u_right=torch.from_numpy(u_r).cuda(), feat_locked=torch.from_numpy(locked).cuda())
        mt, fq, nm = r["match"].cpu().numpy(), r["feat_query"].cpu().numpy(), r["nmatches"].cpu().numpy()
        for f in range(B):
            q = qn[f]
            rm, rfq, rnm = oracle.search_by_projection(kps[f, :n[f]], desc[f, :n[f]], bounds, q["qx"], q["qy"], q["qr"], q["qmin"],
                                                       q["qmax"], q["qdesc"], pmode, nn, True, u_right=u_r[f, :n[f]],
                                                       feat_locked=locked[f, :n[f]], q_ur=q["q_ur"], q_flags=q["flags"], q_angle=q["angle"])
            assert np.array_equal(mt[f, :len(rm)], rm) and np.array_equal(fq[f, :n[f]], rfq) and nm[f] == rnm, (w, h, nf, pmode, th)
            proj_frames += 1
    cases += 1
print(f"stress parity OK: {cases} random configurations, {frames_done} frames / {kps_total} keypoints bit-exact, "
      f"{stereo_pairs} stereo pairs, {proj_frames} projection-search frames, {time.time() - t0:.0f} s")
