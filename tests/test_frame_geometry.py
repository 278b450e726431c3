"""Frame::UndistortKeyPoints and Frame::isInFrustum.

CPU part: the oracle restatement against the real OpenCV binary (cv2.undistortPoints, when cv2 imports) and the
product's host/device-shared arithmetic (csrc/frame_core.h compiled for the host by tests/cpu_harness) against the
oracle.  GPU part: the C ABI (orbm_undistort_keypoints, orbm_is_in_frustum) against the oracle.
Reference: src/Frame.cc:512-586, 747-781, 1168-1242; src/MapPoint.cc:531-546."""
import ctypes as C

import numpy as np
import pytest

from conftest import has_cuda

K_EUROC = (458.654, 457.296, 367.215, 248.375)
D_EUROC = [-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05]          # Examples/Monocular/EuRoC.yaml
D_K3 = [-0.28, 0.07, 0.0002, 1.7e-5, 0.012]
D_RATIONAL = [0.1, -0.05, 1e-4, -2e-4, 0.01, 0.02, -0.01, 0.003]
D_PRISM = D_RATIONAL + [1e-3, -2e-3, 5e-4, 1e-4]


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _kps(n, seed, w=752, h=480):
    from oracle import KP_DTYPE
    rng = np.random.default_rng(seed)
    k = np.zeros(n, KP_DTYPE)
    k["x"] = (rng.random(n) * w).astype(np.float32)
    k["y"] = (rng.random(n) * h).astype(np.float32)
    k["size"] = 31
    k["angle"] = (rng.random(n) * 360).astype(np.float32)
    k["response"] = rng.integers(1, 200, n)
    k["octave"] = rng.integers(0, 8, n)
    k["class_id"] = -1
    return k


@pytest.mark.parametrize("dist", [D_EUROC, D_K3, D_RATIONAL, D_PRISM])
def test_oracle_undistort_is_cv2_undistortPoints(oracle_mod, dist):
    cv2 = pytest.importorskip("cv2")
    fx, fy, cx, cy = K_EUROC
    k = _kps(20000, 3)
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float32)
    pts = np.stack([k["x"], k["y"]], 1).reshape(-1, 1, 2)
    ref = cv2.undistortPoints(pts, K, np.array(dist, np.float32), None, K).reshape(-1, 2)
    out = oracle_mod.undistort_keypoints(k, np.float32(fx), np.float32(fy), np.float32(cx), np.float32(cy), dist)
    assert np.array_equal(out["x"], ref[:, 0]) and np.array_equal(out["y"], ref[:, 1])
    for f in ("size", "angle", "response", "octave", "class_id"):
        assert np.array_equal(out[f], k[f])


def test_oracle_undistort_zero_k1_is_a_copy(oracle_mod):
    k = _kps(100, 4)
    out = oracle_mod.undistort_keypoints(k, *K_EUROC, [0.0, 0.07, 0.001, 0.002])
    assert np.array_equal(out, k)  # Frame.cc:749-753: only mDistCoef[0] is looked at


@pytest.mark.parametrize("dist", [D_EUROC, D_K3, D_RATIONAL, D_PRISM])
def test_harness_undistort_matches_oracle(harness, oracle_mod, dist):
    k = _kps(20000, 5)
    xy = np.ascontiguousarray(np.stack([k["x"], k["y"]], 1))
    out = np.zeros_like(xy)
    d = np.array(dist, np.float32)
    fx, fy, cx, cy = K_EUROC
    harness.h_undistort.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.c_void_p, C.c_int,
                                    C.c_void_p]
    harness.h_undistort(_p(xy), len(k), fx, fy, cx, cy, _p(d), d.size, _p(out))
    ref = oracle_mod.undistort_keypoints(k, fx, fy, cx, cy, dist)
    assert np.array_equal(out[:, 0], ref["x"]) and np.array_equal(out[:, 1], ref["y"])


@pytest.mark.parametrize("dist", [D_EUROC, D_K3, D_RATIONAL, D_PRISM, [0.0, 0.1, 0.0, 0.0], []])
@pytest.mark.parametrize("size", [(752, 480), (1241, 376)])
def test_compute_image_bounds_is_cv2_on_the_corners(dist, size):
    """Frame::ComputeImageBounds (Frame.cc:782-808): the host entry point against cv2.undistortPoints on the four
    corners with the reference's min / max pairing; distortion off -> the image rectangle."""
    cv2 = pytest.importorskip("cv2")
    from orb_slam3_vio_fixes_b200 import frame as F
    w, h = size
    got = F.ComputeImageBounds(w, h, *K_EUROC, dist)
    if not dist or dist[0] == 0.0:
        assert got == (0.0, float(w), 0.0, float(h))
        return
    K = np.array([[K_EUROC[0], 0, K_EUROC[2]], [0, K_EUROC[1], K_EUROC[3]], [0, 0, 1]], np.float32)
    pts = np.array([[0, 0], [w, 0], [0, h], [w, h]], np.float32).reshape(-1, 1, 2)
    m = cv2.undistortPoints(pts, K, np.array(dist, np.float32), None, K).reshape(4, 2)
    want = (min(m[0, 0], m[2, 0]), max(m[1, 0], m[3, 0]), min(m[0, 1], m[1, 1]), max(m[2, 1], m[3, 1]))
    assert tuple(np.float32(v) for v in want) == got


def frustum_case(n, seed, model=0):
    """A camera looking down +z of a slightly rotated pose at a cloud of MapPoints around it."""
    rng = np.random.default_rng(seed)
    ax = rng.normal(size=3); ax /= np.linalg.norm(ax)
    ang = 0.2
    Kx = np.array([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]])
    R = np.eye(3) + np.sin(ang) * Kx + (1 - np.cos(ang)) * Kx @ Kx
    t = rng.normal(size=3) * 0.5
    c = -R.T @ t
    pose = np.concatenate([R.reshape(-1), t, c]).astype(np.float32)
    pos = (rng.normal(size=(n, 3)) * [6, 4, 6] + [0, 0, 8]).astype(np.float32)   # some behind / outside
    normal = rng.normal(size=(n, 3)); normal /= np.linalg.norm(normal, axis=1, keepdims=True)
    d = np.linalg.norm(pos - c, axis=1)
    # point the normals roughly back at the camera for most points (viewCos > 0.5), leave some off
    towards = (pos - c) / d[:, None]
    mix = rng.random(n)[:, None]
    normal = towards * (0.4 + 0.6 * mix) + normal * 0.5 * (1 - mix)
    normal = (normal / np.linalg.norm(normal, axis=1, keepdims=True)).astype(np.float32)
    max_d = (d * rng.uniform(0.7, 3.5, n)).astype(np.float32)
    min_d = (max_d / np.float32(1.2 ** 7)).astype(np.float32)
    if model == 0:
        cam = (0, 458.654, 457.296, 367.215, 248.375)
    else:
        cam = (1, 190.978, 190.973, 254.932, 256.897, 0.0034, 0.00071, -0.0020, 0.00020)  # TUM-VI 512 fisheye
    return pos, normal, min_d, max_d, pose, cam


@pytest.mark.parametrize("rig", [0, 1])
def test_harness_in_frustum_matches_oracle_pinhole(harness, oracle_mod, rig):
    """csrc/frame_core.h (explicitly rounded fp32, orbx_logf) == the oracle (plain C++, libm logf), bit for bit."""
    n = 50000
    pos, normal, mn, mx, pose, cam = frustum_case(n, 7)
    bounds = (0.0, 752.0, 0.0, 480.0)
    log_sf = float(np.log(np.float32(1.2)).astype(np.float32))
    ref = oracle_mod.is_in_frustum(pos, normal, mn, mx, pose, cam, 47.9, bounds, 0.5, log_sf, 8, fisheye_rig=rig)
    o = {k: np.zeros_like(v) for k, v in ref.items()}
    camf = np.array(list(cam[1:]) + [0.0] * (9 - len(cam)), np.float32)
    harness.h_in_frustum.argtypes = [C.c_void_p] * 4 + [C.c_int, C.c_void_p, C.c_int, C.c_void_p] + [C.c_float] * 7 + \
        [C.c_int] * 2 + [C.c_void_p] * 7
    harness.h_in_frustum(_p(pos), _p(normal), _p(mn), _p(mx), n, _p(pose), cam[0], _p(camf), 47.9, *bounds, 0.5, log_sf,
                         8, rig, _p(o["in_view"]), _p(o["proj_x"]), _p(o["proj_y"]), _p(o["proj_xr"]), _p(o["depth"]),
                         _p(o["view_cos"]), _p(o["level"]))
    assert 0.1 * n < ref["in_view"].sum() < 0.9 * n
    assert len(set(ref["level"][ref["in_view"] > 0].tolist())) == 8  # every level is predicted somewhere
    for k in ref:
        assert np.array_equal(o[k], ref[k]), k


def test_in_frustum_fp32_vs_fp64_report(oracle_mod):
    """How many isInFrustum results change when the same gates are evaluated in fp64: the
    in-view decision and the predicted level may flip only for points that sit on a gate / level boundary."""
    n = 200000
    pos, normal, mn, mx, pose, cam = frustum_case(n, 8)
    bounds = (0.0, 752.0, 0.0, 480.0)
    log_sf = float(np.log(np.float32(1.2)).astype(np.float32))
    a = oracle_mod.is_in_frustum(pos, normal, mn, mx, pose, cam, 47.9, bounds, 0.5, log_sf, 8)
    b = oracle_mod.is_in_frustum(pos, normal, mn, mx, pose, cam, 47.9, bounds, 0.5, log_sf, 8, f64=True)
    both = (a["in_view"] > 0) & (b["in_view"] > 0)
    flips = int((a["in_view"] != b["in_view"]).sum())
    lvl = int((a["level"][both] != b["level"][both]).sum())
    du = float(np.abs(a["proj_x"][both] - b["proj_x"][both]).max())
    print(f"isInFrustum fp32 vs fp64 on {n} points: in-view flips {flips}, level differs {lvl}, max |du| {du:.2e} px")
    assert flips <= n // 5000 and lvl <= n // 5000 and du < 1e-3


gpu = pytest.mark.gpu


@gpu
@pytest.mark.skipif(not has_cuda(), reason="needs CUDA")
@pytest.mark.parametrize("dist", [D_EUROC, D_K3, D_PRISM, [0.0, 0.1, 0.0, 0.0]])
def test_gpu_undistort_keypoints(oracle_mod, dist):
    import torch
    from orb_slam3_vio_fixes_b200 import frame as F
    B, stride = 3, 1500
    ns = np.array([1500, 0, 977], np.int32)
    k = np.stack([_kps(stride, 10 + b) for b in range(B)])
    kt = torch.from_numpy(k.view(np.uint8).reshape(B, stride, 28)).cuda()
    out = F.UndistortKeyPoints(kt, torch.from_numpy(ns).cuda(), *K_EUROC, dist)
    got = out.cpu().numpy().reshape(B, stride * 28).view(k.dtype)
    for b in range(B):
        ref = oracle_mod.undistort_keypoints(k[b, :ns[b]], *K_EUROC, dist)
        assert np.array_equal(got[b, :ns[b]], ref)


@gpu
@pytest.mark.skipif(not has_cuda(), reason="needs CUDA")
@pytest.mark.parametrize("rig", [False, True])
def test_gpu_is_in_frustum_pinhole_bit_exact(oracle_mod, rig):
    import torch
    from orb_slam3_vio_fixes_b200 import frame as F
    B, M = 2, 30000
    cases = [frustum_case(M, 20 + b) for b in range(B)]
    nmp = np.array([M, M - 777], np.int32)
    t = lambda i, dt: torch.from_numpy(np.stack([c[i] for c in cases]).astype(dt)).cuda()
    cam = cases[0][5]
    bounds = (0.0, 752.0, 0.0, 480.0)
    log_sf = float(np.log(np.float32(1.2)).astype(np.float32))
    sf = np.cumprod([1.0] + [1.2] * 7).astype(np.float32)
    o = F.isInFrustum(t(0, np.float32), t(1, np.float32), t(2, np.float32), t(3, np.float32), torch.from_numpy(nmp).cuda(),
                      t(4, np.float32), F.camera(*cam), 47.9, bounds, 0.5, log_sf, 8, fisheye_rig=rig, th=3.0,
                      scale_factors=sf)
    for b in range(B):
        n = nmp[b]
        pos, normal, mn, mx, pose, _ = cases[b]
        ref = oracle_mod.is_in_frustum(pos[:n], normal[:n], mn[:n], mx[:n], pose, cam, 47.9, bounds, 0.5, log_sf, 8,
                                       fisheye_rig=rig)
        for k in ref:
            assert np.array_equal(o[k][b, :n].cpu().numpy(), ref[k]), k
        iv = ref["in_view"] > 0
        rad = np.where(ref["view_cos"].astype(np.float64) > 0.998, np.float32(2.5), np.float32(4.0)) * np.float32(3.0)
        qr = np.where(iv, rad.astype(np.float32) * sf[np.maximum(ref["level"], 0)], np.float32(0))
        assert np.array_equal(o["qr"][b, :n].cpu().numpy(), qr.astype(np.float32))
        assert np.array_equal(o["qmax"][b, :n].cpu().numpy()[iv], ref["level"][iv])
        assert np.array_equal(o["qmin"][b, :n].cpu().numpy()[iv], ref["level"][iv] - 1)


@gpu
@pytest.mark.skipif(not has_cuda(), reason="needs CUDA")
def test_gpu_is_in_frustum_kannala_brandt(oracle_mod):
    """KB8 projection: atan2f / cosf / sinf are libm in the reference and CUDA's on the device (a few ulp apart), so
    this branch is held to a tolerance: |du|,|dv| <= 1e-3 px, identical decisions except on a gate boundary."""
    import torch
    from orb_slam3_vio_fixes_b200 import frame as F
    M = 50000
    pos, normal, mn, mx, pose, cam = frustum_case(M, 30, model=1)
    bounds = (0.0, 512.0, 0.0, 512.0)
    log_sf = float(np.log(np.float32(1.2)).astype(np.float32))
    tt = lambda a: torch.from_numpy(np.ascontiguousarray(a)[None]).cuda()
    o = F.isInFrustum(tt(pos), tt(normal), tt(mn), tt(mx), torch.tensor([M], dtype=torch.int32).cuda(), tt(pose),
                      F.camera(cam[0], *cam[1:5], cam[5:9]), 0.0, bounds, 0.5, log_sf, 8, fisheye_rig=True)
    ref = oracle_mod.is_in_frustum(pos, normal, mn, mx, pose, cam, 0.0, bounds, 0.5, log_sf, 8, fisheye_rig=True)
    iv = o["in_view"][0].cpu().numpy()
    assert int((iv != ref["in_view"]).sum()) <= 2
    both = (iv > 0) & (ref["in_view"] > 0)
    assert both.sum() > M // 10
    assert np.abs(o["proj_x"][0].cpu().numpy()[both] - ref["proj_x"][both]).max() <= 1e-3
    assert np.abs(o["proj_y"][0].cpu().numpy()[both] - ref["proj_y"][both]).max() <= 1e-3
    assert np.array_equal(o["level"][0].cpu().numpy()[both], ref["level"][both])
