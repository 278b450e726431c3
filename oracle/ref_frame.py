"""oracle/_ref/libref_frame.so — the hot-path functions of the REFERENCE's own src/Frame.cc compiled here, verbatim —
TEST INFRASTRUCTURE ONLY.

Frame.cc as a whole cannot be compiled in this image (constructors, IMU and BoW members need Eigen, Sophus, g2o,
boost, the real OpenCV).  `make -C oracle ref_frame` cuts Frame::ComputeStereoMatches (Frame.cc:811-981),
ComputeStereoFromRGBD (:984-1005), isInFrustum (:512-586) and isInFrustumChecks (:1168-1242) out of the reference
file, from the signature line to the closing brace, into a scratch file outside the repository and compiles that text
unmodified against the stand-ins of oracle/ref_shim_matcher/prelude.h (oracle/ref_shim_matcher/frame_ref_api.cpp).
tests/test_ref_frame.py compares the results with the oracle's restatements — that is what pins rows m11 and (f).4.

/root/reference does not exist on the GPU box: there the prebuilt .so (git-ignored, travels with the snapshot) is
loaded, never rebuilt."""
import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

from . import _p

_DIR = Path(__file__).resolve().parent
_SO = _DIR / "_ref" / "libref_frame.so"
REFERENCE = Path("/root/reference")
_lib = None


def available():
    return _SO.exists() or (REFERENCE / "src" / "Frame.cc").exists()


def build(force=False):
    if (REFERENCE / "src" / "Frame.cc").exists():
        subprocess.check_call(["make", "-C", str(_DIR), "ref_frame"] + (["-B"] if force else ["-s"]),
                              stdout=subprocess.DEVNULL)
    return _SO if _SO.exists() else None


def lib():
    global _lib
    if _lib is None:
        if not _SO.exists():
            build()
        if not _SO.exists():
            raise FileNotFoundError("oracle/_ref/libref_frame.so: not built and /root/reference absent")
        _lib = C.CDLL(str(_SO))
    return _lib


def _f32(a):
    return np.ascontiguousarray(a, np.float32)


def compute_stereo_matches(kps_l, desc_l, kps_r, desc_r, scale_factors, inv_scale_factors, pyr_l, pyr_r, mb, mbf):
    """Frame::ComputeStereoMatches.  Returns (mvuRight, mvDepth, number of keypoints with a depth)."""
    kl, kr = np.ascontiguousarray(kps_l), np.ascontiguousarray(kps_r)
    dl, dr = np.ascontiguousarray(desc_l, np.uint8), np.ascontiguousarray(desc_r, np.uint8)
    nlev = len(pyr_l)
    pl = [np.ascontiguousarray(p, np.uint8) for p in pyr_l]
    pr = [np.ascontiguousarray(p, np.uint8) for p in pyr_r]
    PL = (C.c_void_p * nlev)(*[p.ctypes.data for p in pl])
    PR = (C.c_void_p * nlev)(*[p.ctypes.data for p in pr])
    pw = np.array([p.shape[1] for p in pl], np.int32)
    ph = np.array([p.shape[0] for p in pl], np.int32)
    sf, isf = _f32(scale_factors), _f32(inv_scale_factors)
    ur, dp = np.zeros(kl.size, np.float32), np.zeros(kl.size, np.float32)
    L = lib()
    L.ref_compute_stereo_matches.restype = C.c_int
    L.ref_compute_stereo_matches.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                             C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                             C.c_float, C.c_float, C.c_void_p, C.c_void_p]
    kept = L.ref_compute_stereo_matches(_p(kl), _p(dl), kl.size, _p(kr), _p(dr), kr.size, _p(sf), _p(isf), nlev, PL, PR,
                                        _p(pw), _p(ph), float(mb), float(mbf), _p(ur), _p(dp))
    return ur, dp, kept


def rgbd_depth(kps, kps_un, depth_img, mbf):
    """Frame::ComputeStereoFromRGBD.  Returns (mvuRight, mvDepth)."""
    k, ku = np.ascontiguousarray(kps), np.ascontiguousarray(kps_un)
    d = _f32(depth_img)
    ur, dp = np.zeros(k.size, np.float32), np.zeros(k.size, np.float32)
    L = lib()
    L.ref_rgbd_depth.restype = None
    L.ref_rgbd_depth.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_void_p,
                                 C.c_void_p]
    L.ref_rgbd_depth(_p(k), _p(ku), k.size, _p(d), d.shape[1], d.shape[0], float(mbf), _p(ur), _p(dp))
    return ur, dp


def is_in_frustum(pos, normal, mf_min_dist, mf_max_dist, pose, cam, mbf, bounds, cos_limit, log_sf, n_levels,
                  cam_r=None, trl=None, tlr=None):
    """Frame::isInFrustum (+ isInFrustumChecks when cam_r / trl / tlr describe a two-camera rig).
    pose = Rcw[9], tcw[3], Ow[3], Rwc[9]; cam = (fx, fy, cx, cy); trl = Rrl[9], trl[3]; tlr = mTlr.translation().
    Returns dict of the MapPoint tracking fields; for a rig also right_pose = R[9], t[3], twc[3] as composed at
    Frame.cc:1176-1180."""
    pos, normal, mn, mx, pose = _f32(pos), _f32(normal), _f32(mf_min_dist), _f32(mf_max_dist), _f32(pose)
    n = mn.size
    rig = cam_r is not None
    camf, camr = _f32(cam), _f32(cam_r if rig else cam)
    trl_, tlr_ = _f32(trl if rig else np.zeros(12)), _f32(tlr if rig else np.zeros(3))
    b = _f32(bounds)
    o = dict(in_view=np.zeros(n, np.uint8), proj_x=np.zeros(n, np.float32), proj_y=np.zeros(n, np.float32),
             proj_xr=np.zeros(n, np.float32), depth=np.zeros(n, np.float32), view_cos=np.zeros(n, np.float32),
             level=np.zeros(n, np.int8), in_view_r=np.zeros(n, np.uint8), proj_xr_r=np.zeros(n, np.float32),
             proj_yr=np.zeros(n, np.float32), depth_r=np.zeros(n, np.float32), view_cos_r=np.zeros(n, np.float32),
             level_r=np.zeros(n, np.int8), right_pose=np.zeros(15, np.float32))
    L = lib()
    L.ref_is_in_frustum.restype = C.c_int
    L.ref_is_in_frustum.argtypes = [C.c_void_p] * 4 + [C.c_int] + [C.c_void_p] * 3 + [C.c_float, C.c_void_p, C.c_float,
                                                                                  C.c_float, C.c_int, C.c_int] + \
        [C.c_void_p] * 16
    o["count"] = L.ref_is_in_frustum(_p(pos), _p(normal), _p(mn), _p(mx), n, _p(pose), _p(camf), _p(camr), float(mbf), _p(b),
                                     float(cos_limit), float(log_sf), int(n_levels), int(rig), _p(trl_), _p(tlr_),
                                     _p(o["in_view"]), _p(o["proj_x"]), _p(o["proj_y"]), _p(o["proj_xr"]), _p(o["depth"]),
                                     _p(o["view_cos"]), _p(o["level"]), _p(o["in_view_r"]), _p(o["proj_xr_r"]),
                                     _p(o["proj_yr"]), _p(o["depth_r"]), _p(o["view_cos_r"]), _p(o["level_r"]),
                                     _p(o["right_pose"]))
    return o


def stereo_fisheye_matches(kps_l, desc_l, mono_l, kps_r, desc_r, mono_r, level_sigma2, tri_depth):
    """Frame::ComputeStereoFishEyeMatches with TriangulateMatches answered from tri_depth [n_l, n_r].
    Returns (mvLeftToRightMatch, mvRightToLeftMatch, mvDepth)."""
    kl, kr = np.ascontiguousarray(kps_l), np.ascontiguousarray(kps_r)
    dl, dr = np.ascontiguousarray(desc_l, np.uint8), np.ascontiguousarray(desc_r, np.uint8)
    sg, td = _f32(level_sigma2), _f32(tri_depth)
    assert td.shape == (kl.size, kr.size)
    l2r, r2l, dp = np.zeros(kl.size, np.int32), np.zeros(kr.size, np.int32), np.zeros(kl.size, np.float32)
    L = lib()
    L.ref_stereo_fisheye_matches.restype = None
    L.ref_stereo_fisheye_matches.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                             C.c_int, C.c_void_p, C.c_int] + [C.c_void_p] * 4
    L.ref_stereo_fisheye_matches(_p(kl), _p(dl), kl.size, int(mono_l), _p(kr), _p(dr), kr.size, int(mono_r), _p(sg), sg.size,
                                 _p(td), _p(l2r), _p(r2l), _p(dp))
    return l2r, r2l, dp


def grid_queries(kps, bounds, qx, qy, qr, qmin, qmax):
    """Frame::AssignFeaturesToGrid + GetFeaturesInArea (the reference's text) -> (cand_off [nq+1], cand_idx)."""
    k = np.ascontiguousarray(kps)
    qx, qy, qr = _f32(qx), _f32(qy), _f32(qr)
    qmin, qmax = np.ascontiguousarray(qmin, np.int32), np.ascontiguousarray(qmax, np.int32)
    b = _f32(bounds)
    nq = qx.size
    off = np.zeros(nq + 1, np.int32)
    cap = max(1, nq * 64)
    L = lib()
    L.ref_grid_queries.restype = C.c_int
    L.ref_grid_queries.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 6 + [C.c_int, C.c_void_p, C.c_void_p, C.c_int]
    while True:
        idx = np.zeros(cap, np.int32)
        tot = L.ref_grid_queries(_p(k), k.size, _p(b), _p(qx), _p(qy), _p(qr), _p(qmin), _p(qmax), nq, _p(off), _p(idx), cap)
        if tot <= cap:
            return off, idx[:tot].copy()
        cap = tot


def distinctive_descriptor(desc, first, second=None, kf_bad=None):
    """MapPoint::ComputeDistinctiveDescriptors.  desc [rows, 32]; observation o holds row first[o] (and second[o] when
    >= 0).  Returns the chosen 32-byte descriptor, or None when the MapPoint kept none."""
    d = np.ascontiguousarray(desc, np.uint8)
    f = np.ascontiguousarray(first, np.int32)
    s2 = np.full(f.size, -1, np.int32) if second is None else np.ascontiguousarray(second, np.int32)
    kb = np.zeros(f.size, np.uint8) if kf_bad is None else np.ascontiguousarray(kf_bad, np.uint8)
    out = np.zeros(32, np.uint8)
    L = lib()
    L.ref_distinctive_descriptor.restype = C.c_int
    L.ref_distinctive_descriptor.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    ok = L.ref_distinctive_descriptor(_p(d), d.shape[0], _p(f), _p(s2), _p(kb), f.size, _p(out))
    return out if ok else None


def camera_project(model, cam, xyz):
    """Pinhole::project (model 0, cam = fx, fy, cx, cy) / KannalaBrandt8::project (model 1, cam + k0..k3) of
    Eigen::Vector3f points, the reference's own text.  Returns uv [n, 2]."""
    c = _f32(list(cam) + [0.0] * (8 - len(cam)))
    P = _f32(xyz).reshape(-1, 3)
    uv = np.zeros((P.shape[0], 2), np.float32)
    L = lib()
    L.ref_camera_project.restype = None
    L.ref_camera_project.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    L.ref_camera_project(int(model), _p(c), _p(P), P.shape[0], _p(uv))
    return uv


def kb8_unproject(cam, uv):
    """KannalaBrandt8::unprojectEig (10 Newton steps, precision 1e-6), the reference's own text.  Returns rays [n, 3]."""
    c = _f32(cam)
    p = _f32(uv).reshape(-1, 2)
    ray = np.zeros((p.shape[0], 3), np.float32)
    L = lib()
    L.ref_kb8_unproject.restype = None
    L.ref_kb8_unproject.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    L.ref_kb8_unproject(_p(c), _p(p), p.shape[0], _p(ray))
    return ray
