"""GPU parity: the CUDA extractor (through the C ABI) against the CPU oracle, stage by stage."""
import numpy as np
import pytest

from orb_slam3_vio_fixes_b200 import synthetic as S
from orb_slam3_vio_fixes_b200.extractor import ORBextractor

pytestmark = pytest.mark.gpu

ORB = dict(scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7)


def _compare_frame(oracle_mod, ex, orc, img, lap):
    mono, kps, desc = ex(img, None, lap)
    okps, odesc, omono = orc.extract(img, lap)
    for level in range(ex.nlevels):
        assert np.array_equal(ex.GetPyramid(level), orc.pyramid(level)), f"pyramid level {level}"
        cand = ex.debug_level_keypoints(level, 0)
        oc = orc.level_keypoints(level, 0)
        ocand = np.stack([oc["x"], oc["y"], oc["response"]], 1).astype(np.int32)
        assert np.array_equal(cand, ocand), f"FAST candidates level {level}: {len(cand)} vs {len(ocand)}"
        sel = ex.debug_level_keypoints(level, 1)
        os_ = orc.level_keypoints(level, 1)
        osel = np.stack([os_["x"], os_["y"], os_["response"]], 1).astype(np.int32)
        assert np.array_equal(sel, osel), f"octree level {level}: {len(sel)} vs {len(osel)}"
    assert mono == omono
    assert len(kps) == len(okps)
    for f in ("x", "y", "size", "response", "octave", "class_id"):
        assert np.array_equal(kps[f], okps[f]), f
    # angles: 1e-3 deg tolerance stated by north_star; expected (and asserted) bit-exact
    assert np.max(np.abs(kps["angle"] - okps["angle"]), initial=0) <= 1e-3
    assert np.array_equal(kps["angle"], okps["angle"])
    assert np.array_equal(desc, odesc)


@pytest.mark.parametrize("cfg,lap", [("C1", (0, 1000)), ("C3", (0, 0))])
def test_single_frame_stages(oracle_mod, cfg, lap):
    w, h, nf, seed, _ = S.CONFIGS[cfg]
    frames = S.make_frames(cfg, 2)
    ex = ORBextractor(nf, **ORB)
    orc = oracle_mod.Extractor(nf, 1.2, 8, 20, 7)
    for img in frames:
        _compare_frame(oracle_mod, ex, orc, img, lap)
    with_b = ex.GetPyramid(3, with_border=True)
    assert np.array_equal(with_b, orc.pyramid(3, True))


def test_pyramid_border_and_tables(oracle_mod):
    frames = S.make_frames("C1", 1)
    ex = ORBextractor(1000, **ORB)
    orc = oracle_mod.Extractor(1000, 1.2, 8, 20, 7)
    ex(frames[0], None, (0, 0))
    orc.extract(frames[0], (0, 0))
    for level in range(8):
        assert np.array_equal(ex.GetPyramid(level, with_border=True), orc.pyramid(level, True))
    assert np.array_equal(ex.GetScaleFactors(), orc.scale_factors)
    assert np.array_equal(ex.GetInverseScaleFactors(), orc.inv_scale_factors)
    assert np.array_equal(ex.GetScaleSigmaSquares(), orc.level_sigma2)
    assert np.array_equal(ex.GetInverseScaleSigmaSquares(), orc.inv_level_sigma2)
    assert np.array_equal(ex.features_per_level(), orc.features_per_level)


def test_batch_device_matches_single(oracle_mod):
    import torch
    frames = S.make_frames("C1", 6)
    ex = ORBextractor(1000, max_batch=6, **ORB)
    orc = oracle_mod.Extractor(1000, 1.2, 8, 20, 7)
    d = torch.from_numpy(frames).cuda()
    out = ex.extract_batch_device(d, (0, 1000))
    torch.cuda.synchronize()
    n = out["n"].cpu().numpy()
    kps = ORBextractor.kps_from_tensor(out["kps"])
    desc = out["desc"].cpu().numpy()
    for f in range(len(frames)):
        ok, od, om = orc.extract(frames[f], (0, 1000))
        assert n[f] == len(ok) and out["mono"][f].item() == om
        assert np.array_equal(kps[f, :n[f]], ok)
        assert np.array_equal(desc[f, :n[f]], od)
    # host-buffer batch entry point gives the same bytes
    k2, d2, n2, m2 = ex.extract_batch_host(frames, (0, 1000))
    assert np.array_equal(n2, n)
    for f in range(len(frames)):
        assert np.array_equal(k2[f, :n[f]], kps[f, :n[f]]) and np.array_equal(d2[f, :n[f]], desc[f, :n[f]])


def test_empty_image_and_flat_image():
    ex = ORBextractor(1000, **ORB)
    mono, kps, desc = ex(np.zeros((0, 0), np.uint8))
    assert mono == -1 and len(kps) == 0
    flat = np.full((480, 752), 77, np.uint8)
    mono, kps, desc = ex(flat, None, (0, 0))
    assert mono == 0 and len(kps) == 0 and desc.shape == (0, 32)


@pytest.mark.parametrize("ini,mn,contrast,size", [
    (20, 7, 0.25, (752, 480)),    # low contrast: most cells fall back to minThFAST
    (20, 7, 0.08, (640, 480)),    # almost flat: both thresholds fail in many cells
    (7, 20, 1.0, (400, 300)),     # minTh > iniTh (the retry can only return less)
    (40, 40, 1.0, (752, 480)),    # no retry
    (0, 0, 0.05, (320, 240)),     # threshold 0: everything passes the quick test
    (130, 127, 1.0, (752, 480)),  # thresholds above 126
    (255, 1, 1.0, (320, 240)),    # nothing at iniTh anywhere
    (20, 7, 1.0, (1241, 376)),    # KITTI shape (wide cells)
    (20, 7, 1.0, (501, 317)),     # odd sizes
])
def test_fast_thresholds_and_fallback(oracle_mod, ini, mn, contrast, size):
    w, h = size
    img = S.make_frames("C1", 1, width=w, height=h)[0].astype(np.float64)
    img = np.clip((img - 128.0) * contrast + 128.0, 0, 255).astype(np.uint8)
    ex = ORBextractor(800, 1.2, 8, ini, mn)
    orc = oracle_mod.Extractor(800, 1.2, 8, ini, mn)
    mono, kps, desc = ex(img, None, (0, 0))
    okps, odesc, omono = orc.extract(img, (0, 0))
    for level in range(ex.nlevels):
        cand = ex.debug_level_keypoints(level, 0)
        oc = orc.level_keypoints(level, 0)
        ocand = np.stack([oc["x"], oc["y"], oc["response"]], 1).astype(np.int32)
        assert np.array_equal(cand, ocand), f"FAST candidates level {level}: {len(cand)} vs {len(ocand)}"
    assert mono == omono and np.array_equal(kps, okps) and np.array_equal(desc, odesc)


def test_batch_device_unaligned_pitch(oracle_mod):
    """Caller-owned frames whose row pitch is not a multiple of 4 take the byte-wise staging paths."""
    import torch
    frames = S.make_frames("C3", 3, width=501, height=317)
    ex = ORBextractor(600, max_batch=3, **ORB).plan(501, 317)
    orc = oracle_mod.Extractor(600, 1.2, 8, 20, 7)
    out = ex.extract_batch_device(torch.from_numpy(frames).cuda(), (0, 0))
    torch.cuda.synchronize()
    n = out["n"].cpu().numpy()
    kps = ORBextractor.kps_from_tensor(out["kps"])
    desc = out["desc"].cpu().numpy()
    for f in range(len(frames)):
        ok, od, om = orc.extract(frames[f], (0, 0))
        assert n[f] == len(ok) and np.array_equal(kps[f, :n[f]], ok) and np.array_equal(desc[f, :n[f]], od)


@pytest.mark.parametrize("mode", ["fused", "level"])
@pytest.mark.parametrize("cfg,size", [("C1", None), ("C3", None), ("C4", None), ("C1", (501, 317))])
def test_descriptor_modes_match_oracle(oracle_mod, mode, cfg, size):
    """Both descriptor-stage modes (per-keypoint blur in shared memory / whole-level blur) give the
    oracle's keypoints, angles and descriptors, including tiles and patches at the level edges."""
    w, h, nf, seed, _ = S.CONFIGS[cfg]
    if size:
        w, h = size
    frames = S.make_frames(cfg, 2, width=w, height=h)
    ex = ORBextractor(nf, blur_mode=mode, **ORB)
    orc = oracle_mod.Extractor(nf, 1.2, 8, 20, 7)
    for img in frames:
        mono, kps, desc = ex(img, None, (0, 0))
        okps, odesc, omono = orc.extract(img, (0, 0))
        assert mono == omono and np.array_equal(kps, okps)
        assert np.array_equal(desc, odesc)
    for strict in (True,):
        ex2 = ORBextractor(nf, blur_mode=mode, strict_fp=strict, **ORB)
        orc2 = oracle_mod.Extractor(nf, 1.2, 8, 20, 7, flags=1)  # ORACLE_STRICT_FP
        mono, kps, desc = ex2(frames[0], None, (0, 0))
        okps, odesc, omono = orc2.extract(frames[0], (0, 0))
        assert np.array_equal(kps, okps) and np.array_equal(desc, odesc)


def test_batch_device_unaligned_pitch_level_blur(oracle_mod):
    import torch
    frames = S.make_frames("C3", 2, width=501, height=317)
    ex = ORBextractor(600, max_batch=2, blur_mode="level", **ORB).plan(501, 317)
    orc = oracle_mod.Extractor(600, 1.2, 8, 20, 7)
    out = ex.extract_batch_device(torch.from_numpy(frames).cuda(), (0, 0))
    torch.cuda.synchronize()
    n = out["n"].cpu().numpy()
    kps = ORBextractor.kps_from_tensor(out["kps"])
    desc = out["desc"].cpu().numpy()
    for f in range(len(frames)):
        ok, od, om = orc.extract(frames[f], (0, 0))
        assert n[f] == len(ok) and np.array_equal(kps[f, :n[f]], ok) and np.array_equal(desc[f, :n[f]], od)


@pytest.mark.parametrize("cfg,lap,nframes", [("C2", (0, 0), 2), ("C5", (0, 1000), 1)])
def test_full_size_configs(oracle_mod, cfg, lap, nframes):
    """BASELINE configs not covered above at their full sizes: C2 (K=1200/eye, both eyes) and the
    4K stress shape C5 (3840x2160, K=5000: 19.7 k FAST cells, ~1 k keypoints per octree level)."""
    w, h, nf, seed, stereo = S.CONFIGS[cfg]
    ex = ORBextractor(nf, **ORB)
    orc = oracle_mod.Extractor(nf, 1.2, 8, 20, 7)
    eyes = (False, True) if stereo else (False,)
    for right in eyes:
        for img in S.make_frames(cfg, nframes, right=right):
            _compare_frame(oracle_mod, ex, orc, img, lap)


@pytest.mark.parametrize("chunk,edge", [(4, 0), (4, 2), (5, 3), (64, 0)])
def test_host_batch_chunk_schedules_give_identical_bytes(monkeypatch, chunk, edge):
    """orbx_extract_batch_host: any chunk schedule (ORBX_HOST_CHUNK, shorter first/last chunk ORBX_HOST_EDGE, ragged
    tail) returns the bytes of the device-resident path."""
    import torch
    monkeypatch.setenv("ORBX_HOST_CHUNK", str(chunk))
    monkeypatch.setenv("ORBX_HOST_EDGE", str(edge))
    B = 23
    frames = S.make_frames("C3", B, width=400, height=300)
    ex = ORBextractor(500, max_batch=B, **ORB)
    out = ex.extract_batch_device(torch.from_numpy(frames).cuda(), (0, 0))
    torch.cuda.synchronize()
    n = out["n"].cpu().numpy()
    kps = ORBextractor.kps_from_tensor(out["kps"])
    desc = out["desc"].cpu().numpy()
    k2, d2, n2, m2 = ex.extract_batch_host(frames, (0, 0))
    assert np.array_equal(n2, n) and n.min() > 100
    assert np.array_equal(m2, out["mono"].cpu().numpy())
    for f in range(B):
        assert np.array_equal(k2[f, :n[f]], kps[f, :n[f]]) and np.array_equal(d2[f, :n[f]], desc[f, :n[f]])


@pytest.mark.parametrize("size", [(752, 480), (640, 480), (501, 317), (502, 300), (503, 301), (500, 299), (1241, 376)])
def test_blur_levels_match_oracle(oracle_mod, size):
    """Level-blur mode: every blurred pyramid level == GaussianBlur(7x7, sigma 2, REFLECT_101) of that level
    (ORBextractor.cc:1131-1133), all pixels including the 3-px reflected frame, for level widths of every residue
    mod 4 (the last word of a row is partial) and tile edges in both directions."""
    w, h = size
    frames = S.make_frames("C1", 2, width=w, height=h)
    ex = ORBextractor(800, max_batch=2, blur_mode="level", **ORB).plan(w, h)
    import torch
    ex.extract_batch_device(torch.from_numpy(frames).cuda(), (0, 0))
    torch.cuda.synchronize()
    orc = oracle_mod.Extractor(800, 1.2, 8, 20, 7)
    for f in (0, 1):
        orc.extract(frames[f], (0, 0))
        for level in range(8):
            got = ex.debug_blur_level(level, frame=f)
            ref = oracle_mod.gaussian_blur7(orc.pyramid(level))
            bad = np.argwhere(got != ref)
            assert len(bad) == 0, f"{size} frame {f} level {level} {got.shape}: {len(bad)} px differ, first {bad[:4].tolist()}"


def test_handles_of_different_sizes_coexist(oracle_mod):
    """the octree's dynamic-shared-memory opt-in belongs to (function, device), not to a handle.  The mono
    initialisation extractor (5*nFeatures, Tracking.cc:595-601) is created first, the regular one afterwards; the
    big handle must still run."""
    img = S.make_frames("C3", 1)[0]
    big = ORBextractor(5000, **ORB)
    mono_b, k_b, d_b = big(img, None, (0, 1000))
    small = ORBextractor(1000, **ORB)
    mono_s, k_s, d_s = small(img, None, (0, 0))
    mono_b2, k_b2, d_b2 = big(img, None, (0, 1000))
    assert mono_b2 == mono_b and np.array_equal(k_b2, k_b) and np.array_equal(d_b2, d_b)
    ok, od, om = oracle_mod.Extractor(5000, 1.2, 8, 20, 7).extract(img, (0, 1000))
    assert om == mono_b and np.array_equal(ok, k_b) and np.array_equal(od, d_b)
    ok, od, om = oracle_mod.Extractor(1000, 1.2, 8, 20, 7).extract(img, (0, 0))
    assert om == mono_s and np.array_equal(ok, k_s) and np.array_equal(od, d_s)


def test_device_sincosf_exhaustive(oracle_mod):
    """The nvcc build of orbx_sincosf == glibc sincosf for every float in [0, 6.3] (1.09 G bit patterns): chunked
    order-independent checksums of (sin, cos) bits, device vs libm."""
    import ctypes as C
    from orb_slam3_vio_fixes_b200 import _lib
    hi = int(np.float32(6.3).view(np.uint32))
    count, chunk = hi + 1, 1 << 22
    n = (count + chunk - 1) // chunk
    got = np.zeros(n, np.uint64)
    _lib.check(_lib.lib().orbx_debug_sincos_checksum(0, 0, count, chunk, got.ctypes.data, n), "sincos checksum")
    ref = oracle_mod.sincos_checksum(0, count, chunk)
    bad = np.nonzero(got != ref)[0]
    assert len(bad) == 0, f"{len(bad)} of {n} chunks differ, first chunk {bad[:3]}"


def test_two_handles_from_two_threads(oracle_mod):
    """Frame.cc:122-125 calls the left and right extractor from two std::threads: two handles driven concurrently from
    two threads of ONE process give the oracle's bytes.  With >= 2 GPUs the second handle lives on device 1 (SURVEY
    §8(e): one process, explicit cudaSetDevice, one handle per GPU); the per-device cache of the dynamic shared-memory
    opt-in (orbx_ensure_dyn_smem) is what this exercises there — 5000 features need > 48 KB in k_octree."""
    import threading
    import torch
    ndev = torch.cuda.device_count()
    frames = [S.make_frames("C2", 3), S.make_frames("C2", 3, right=True)]
    res, err = [None, None], [None, None]

    def work(i):
        try:
            ex = ORBextractor(5000 if i == 0 else 1200, device=(i if ndev >= 2 else 0), **ORB)
            out = []
            for rep in range(4):
                for f in frames[i]:
                    mono, k, d = ex(f, None, (0, 0))
                    out.append((mono, k.copy(), d.copy()))
            res[i] = out
        except Exception as e:  # noqa: BLE001
            err[i] = e

    th = [threading.Thread(target=work, args=(i,)) for i in range(2)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert err == [None, None], err
    for i in range(2):
        orc = oracle_mod.Extractor(5000 if i == 0 else 1200)
        want = [orc.extract(f, (0, 0)) for f in frames[i]]
        for j, (mono, k, d) in enumerate(res[i]):
            wk, wd, wm = want[j % 3]
            assert mono == wm and np.array_equal(k, wk) and np.array_equal(d, wd), (i, j)


def test_get_pyramid_all_equals_per_level_copies(oracle_mod):
    """orbx_get_pyramid_all (one device->host transfer for mvImagePyramid) == orbx_get_pyramid level by level == oracle,
    with and without the 19-px REFLECT_101 ring."""
    img = S.make_frames("C3", 1)[0]
    ex = ORBextractor(1000, **ORB)
    ex(img, None, (0, 0))
    orc = oracle_mod.Extractor(1000)
    orc.extract(img, (0, 0))
    for border in (True, False):
        levels = ex.GetPyramidAll(with_border=border)
        assert len(levels) == 8
        for l, a in enumerate(levels):
            assert np.array_equal(a, ex.GetPyramid(l, with_border=border))
            assert np.array_equal(a, orc.pyramid(l, border))
