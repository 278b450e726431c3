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
