#!/usr/bin/env python3
"""Generates tests/golden/*.npz — input frames plus the expected outputs of the hot path.

Run in the BUILD container (needs cv2): every golden is written only after the C++ oracle and the
independent cv2-based pipeline (real OpenCV primitives, oracle/cv2_pipeline.py) AND the reference's own
ORBextractor.cc compiled here (oracle/_ref) running over the real cv2 primitives produced identical
keypoints and descriptors for it, which is recorded in the file (`pinned_by_cv2`,
`pinned_by_reference_build`).  The GPU box
needs neither cv2 nor /root/reference to check against these.
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

import oracle  # noqa: E402
from bench import shifts_for, MATCH_TH  # noqa: E402
from orb_slam3_vio_fixes_b200 import synthetic as S  # noqa: E402

OUT = Path(__file__).resolve().parent


def extract_case(name, frames, nf, lap, cv2_check=True):
    from oracle.cv2_pipeline import Cv2Extractor
    from oracle import ref as R
    orc = oracle.Extractor(nf)
    cvx = Cv2Extractor(nf)
    # the reference's own ORBextractor.cc (oracle/_ref) running over the real OpenCV binary
    R.build()
    R.use_cv2_primitives()
    refx = R.RefExtractor(nf)
    d = {"frames": frames, "nfeatures": nf, "lap": np.array(lap, np.int32), "pinned_by_cv2": np.array(cv2_check),
         "pinned_by_reference_build": np.array(True)}
    for i, img in enumerate(frames):
        k, de, m = orc.extract(img, lap)
        k3, d3, m3 = refx.extract(img, lap)
        assert m == m3 and np.array_equal(k, k3) and np.array_equal(de, d3), (name, i, "reference build")
        if cv2_check:
            k2, d2, m2 = cvx.extract(img, lap)
            assert m == m2 and np.array_equal(k, k2) and np.array_equal(de, d2), (name, i)
        d[f"kps{i}"], d[f"desc{i}"], d[f"mono{i}"] = k, de, np.array(m, np.int32)
        if i == 0:
            d["pyr3_border"] = orc.pyramid(3, True)
            d["pyr7"] = orc.pyramid(orc.nlevels - 1)
            for level in (0, 4):
                c = orc.level_keypoints(level, 0)
                d[f"cand{level}"] = np.stack([c["x"], c["y"], c["response"]], 1).astype(np.int32)
    d["scale_factors"] = orc.scale_factors
    return d, orc


def main():
    # C1 (mono quirk lapping {0,1000}), two consecutive frames + the frame-to-frame match
    frames = S.make_frames("C1", 2)
    d, orc = extract_case("C1", frames, 1000, (0, 1000))
    sh = shifts_for(0, 2)
    for f in range(2):
        s = (f - 1) % 2
        kq, kr = d[f"kps{s}"], d[f"kps{f}"]
        qx, qy = kq["x"] + sh[f, 0], kq["y"] + sh[f, 1]
        qr = (np.float32(MATCH_TH) * d["scale_factors"][kq["octave"]]).astype(np.float32)
        off, idx = oracle.grid_queries(kr, (0.0, 752.0, 0.0, 480.0), qx, qy, qr, kq["octave"] - 1, kq["octave"] + 1)
        d[f"match{f}"] = oracle.argmin2_csr(d[f"desc{s}"], d[f"desc{f}"], off, idx, kr["octave"].astype(np.int8), 0)
        d[f"match{f}_ncand"] = np.array(len(idx), np.int64)
    np.savez_compressed(OUT / "c1_mono.npz", **d)

    # C3 (RealSense shape, lapping {0,0})
    d, _ = extract_case("C3", S.make_frames("C3", 1), 1000, (0, 0))
    np.savez_compressed(OUT / "c3_rgbd.npz", **d)

    # small frame, odd size, partial lapping area
    d, _ = extract_case("small", S.make_frames("C1", 1, width=411, height=307), 300, (100, 250))
    np.savez_compressed(OUT / "small_lap.npz", **d)

    # C2 stereo pair: both eyes + Hamming stage of ComputeStereoMatches
    left, right = S.make_frames("C2", 1), S.make_frames("C2", 1, right=True)
    dl, ol = extract_case("C2L", left, 1200, (0, 0))
    dr, orr = extract_case("C2R", right, 1200, (0, 0))
    fx, bf = np.float32(458.654), np.float32(47.90639384423901)
    mb = np.float32(bf / fx)
    pl = [ol.pyramid(l) for l in range(8)]
    pr = [orr.pyramid(l) for l in range(8)]
    bi, bd, ur, dp, kept = oracle.stereo_matches(dl["kps0"], dl["desc0"], dr["kps0"], dr["desc0"], ol.scale_factors,
                                                 ol.inv_scale_factors, pl, pr, mb, bf)
    np.savez_compressed(OUT / "c2_stereo.npz", left=left, right=right, kps_l=dl["kps0"], desc_l=dl["desc0"],
                        kps_r=dr["kps0"], desc_r=dr["desc0"], best_idx_r=bi, best_dist=bd, u_right=ur, depth=dp,
                        kept=np.array(kept), mb=mb, bf=bf, scale_factors=ol.scale_factors,
                        pinned_by_cv2=np.array(True), pinned_by_reference_build=np.array(True))
    for p in sorted(OUT.glob("*.npz")):
        print(p.name, p.stat().st_size)


if __name__ == "__main__":
    main()
