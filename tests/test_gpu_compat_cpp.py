"""The C++ drop-in classes (include/orbx_compat) built against liborbx.so and run on the GPU."""
import subprocess
from pathlib import Path

import numpy as np
import pytest

from orb_slam3_vio_fixes_b200 import synthetic as S
from orb_slam3_vio_fixes_b200 import build as B

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parents[1]


def _build_exe(tmp, name="compat_main"):
    B.build()
    exe = tmp / name
    pkg = ROOT / "orb_slam3_vio_fixes_b200"
    subprocess.check_call(["nvcc", "-O2", "-std=c++17", "-DORBX_NO_OPENCV", f"-I{ROOT / 'include'}",
                           str(ROOT / "tests" / "cpp" / f"{name}.cpp"), "-o", str(exe), f"-L{pkg}", "-lorbx",
                           "-Xlinker", f"-rpath={pkg}"])
    return exe


@pytest.mark.parametrize("lap", [(0, 0), (0, 1000)])
def test_cpp_dropin_matches_oracle(oracle_mod, tmp_path, lap):
    exe = _build_exe(tmp_path)
    img = S.make_frames("C1", 1)[0]
    raw = tmp_path / "in.raw"
    img.tofile(raw)
    out = subprocess.check_output([str(exe), str(raw), "752", "480", "1000", str(lap[0]), str(lap[1]),
                                   str(tmp_path / "o")], text=True)
    mono, n, dd, levels, nsf = [int(v) for v in out.split()]
    orc = oracle_mod.Extractor(1000)
    okps, odesc, omono = orc.extract(img, lap)
    kps = np.fromfile(tmp_path / "o.kps", dtype=oracle_mod.KP_DTYPE)
    desc = np.fromfile(tmp_path / "o.desc", dtype=np.uint8).reshape(-1, 32)
    assert (mono, n, levels, nsf) == (omono, len(okps), 8, 8)
    assert np.array_equal(kps, okps) and np.array_equal(desc, odesc)
    pyr3 = np.fromfile(tmp_path / "o.pyr3", dtype=np.uint8)
    assert np.array_equal(pyr3, orc.pyramid(3, True).ravel())
    assert dd == oracle_mod.descriptor_distance(odesc[0], odesc[1])
    res = np.fromfile(tmp_path / "o.match", dtype=oracle_mod.RES_DTYPE)
    off = (np.arange(n + 1) * 8).astype(np.int32)
    idx = ((np.arange(n)[:, None] + np.arange(1, 9)[None, :]) % n).astype(np.int32).ravel()
    ref = oracle_mod.argmin2_csr(odesc, odesc, off, idx, okps["octave"].astype(np.int8), 0)
    assert np.array_equal(res, ref)


def test_cpp_stereo_dropin_matches_oracle(oracle_mod, tmp_path):
    """Frame::ComputeStereoMatches replaced by ORB_SLAM3::ComputeStereoMatches(left, right, mb, mbf, ...)."""
    exe = _build_exe(tmp_path, "compat_stereo")
    left, right = S.make_frames("C2", 1)[0], S.make_frames("C2", 1, right=True)[0]
    left.tofile(tmp_path / "l.raw")
    right.tofile(tmp_path / "r.raw")
    mbf = np.float32(47.90639384423901)
    mb = np.float32(mbf / np.float32(458.654))
    out = subprocess.check_output([str(exe), str(tmp_path / "l.raw"), str(tmp_path / "r.raw"), "752", "480", "1200",
                                   repr(float(mb)), repr(float(mbf)), str(tmp_path / "o")], text=True)
    nl, nr, kept = [int(v) for v in out.split()]
    el, er = oracle_mod.Extractor(1200), oracle_mod.Extractor(1200)
    kl, dl, _ = el.extract(left, (0, 0))
    kr, dr, _ = er.extract(right, (0, 0))
    pl = [el.pyramid(l) for l in range(8)]
    pr = [er.pyramid(l) for l in range(8)]
    bi, bd, ur, dp, rkept = oracle_mod.stereo_matches(kl, dl, kr, dr, el.scale_factors, el.inv_scale_factors, pl, pr, mb, mbf)
    assert (nl, nr, kept) == (len(kl), len(kr), rkept) and kept > 100
    assert np.array_equal(np.fromfile(tmp_path / "o.ur", dtype=np.uint32), ur.view(np.uint32))
    assert np.array_equal(np.fromfile(tmp_path / "o.depth", dtype=np.uint32), dp.view(np.uint32))


def test_cpp_bow_helpers_match_oracle(oracle_mod, tmp_path):
    """include/orbx_compat/BowMatcher.h (SearchByBoW x2, SearchForTriangulation on std:: containers) vs the oracle."""
    exe = _build_exe(tmp_path, "compat_bow")
    frames = S.make_frames("C1", 2)
    orc = oracle_mod.Extractor(1000)
    (k1, d1, _), (k2, d2, _) = [orc.extract(f, (0, 0)) for f in frames]
    voc = S.make_vocabulary(10, 4, seed=5, desc_pool=np.concatenate([d1, d2]))
    fv1 = oracle_mod.bow_transform(d1, voc, 2)["fv"]
    fv2 = oracle_mod.bow_transform(d2, voc, 2)["fv"]
    rng = np.random.default_rng(9)
    v1 = (rng.random(len(k1)) < 0.7).astype(np.uint8)
    v2 = (rng.random(len(k2)) < 0.7).astype(np.uint8)
    u1 = np.where(rng.random(len(k1)) < 0.5, 100.0, -1.0).astype(np.float32)
    u2 = np.where(rng.random(len(k2)) < 0.5, 100.0, -1.0).astype(np.float32)
    F = np.array([0, -0.01, 0.447, 0.01, 0, -0.894, -0.447, 0.894, 0], np.float32)
    ep = np.array([500.0, 300.0], np.float32)
    sf = np.array([1.2 ** i for i in range(8)], np.float32)
    sg = (sf * sf).astype(np.float32)
    with open(tmp_path / "in.bin", "wb") as f:
        f.write(np.array([len(k1), len(k2), len(fv1[0]), len(fv2[0])], np.int32).tobytes())
        for k, d, v, u, fv in ((k1, d1, v1, u1, fv1), (k2, d2, v2, u2, fv2)):
            idx = np.zeros(len(k), np.int32)
            idx[:len(fv[2])] = fv[2]
            for a in (k, d, v, u, fv[0], fv[1], idx):
                f.write(np.ascontiguousarray(a).tobytes())
        for a in (F, ep, sf, sg):
            f.write(a.tobytes())
    out = subprocess.check_output([str(exe), str(tmp_path / "in.bin"), str(tmp_path / "o")], text=True)
    n0, n1, n2, npairs = [int(v) for v in out.split()]
    m12, m21, r0 = oracle_mod.search_by_bow(0, k1, d1, fv1, k2, d2, fv2, 0.7, True, v1, v2)
    assert n0 == r0 > 50
    assert np.array_equal(np.fromfile(tmp_path / "o.bow0_m12", np.int32), m12)
    assert np.array_equal(np.fromfile(tmp_path / "o.bow0_m21", np.int32), m21)
    m12, _, r1 = oracle_mod.search_by_bow(1, k1, d1, fv1, k2, d2, fv2, 0.8, True, v1, v2)
    assert n1 == r1 > 50
    assert np.array_equal(np.fromfile(tmp_path / "o.bow1_m12", np.int32), m12)
    t12, r2 = oracle_mod.search_for_triangulation(k1, d1, fv1, k2, d2, fv2, F, ep, sf, sg, False, False, True,
                                                  1 - v1, 1 - v2, u1, u2)
    assert n2 == r2 == npairs
    pairs = np.fromfile(tmp_path / "o.tri_pairs", np.int32).reshape(-1, 2)
    ref = np.stack([np.nonzero(t12 >= 0)[0], t12[t12 >= 0]], 1)
    assert np.array_equal(pairs, ref)
