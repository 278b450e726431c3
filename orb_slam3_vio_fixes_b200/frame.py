"""Host-side mirror of the per-point Frame geometry (reference src/Frame.cc:512-586, 747-781, 1168-1242;
src/MapPoint.cc:531-546) on the C ABI: UndistortKeyPoints and the isInFrustum batch projection.
Inputs and outputs are torch CUDA tensors; everything runs on the current stream.  No CPU fallback."""
import ctypes as C

import numpy as np

from . import _lib

PINHOLE, KANNALA_BRANDT8 = 0, 1


def _stream(dev):
    import torch
    return torch.cuda.current_stream(dev).cuda_stream


def camera(model, fx, fy, cx, cy, k=(0.0, 0.0, 0.0, 0.0)):
    cam = _lib.OrbmCamera()
    cam.model, cam.fx, cam.fy, cam.cx, cam.cy = int(model), fx, fy, cx, cy
    for i in range(4):
        cam.k[i] = k[i]
    return cam


def UndistortKeyPoints(kps, n, fx, fy, cx, cy, dist_coef, out=None):
    """Frame::UndistortKeyPoints for a batch: kps [B, stride, 28] u8 (cv::KeyPoint records), n [B] i32.
    dist_coef = mDistCoef (4, 5, 8 or 12 floats).  Returns mvKeysUn in the same layout."""
    import torch
    B, stride = kps.shape[0], kps.shape[1]
    if out is None:
        out = torch.empty_like(kps)
    d = np.ascontiguousarray(dist_coef, np.float32)
    rc = _lib.lib().orbm_undistort_keypoints(kps.data_ptr(), n.data_ptr(), stride, B, fx, fy, cx, cy,
                                             d.ctypes.data_as(C.c_void_p) if d.size else None, d.size,
                                             out.data_ptr(), _stream(kps.device))
    _lib.check(rc, "orbm_undistort_keypoints")
    return out


def ComputeImageBounds(width, height, fx, fy, cx, cy, dist_coef):
    """Frame::ComputeImageBounds (Frame.cc:782-808).  Returns (mnMinX, mnMaxX, mnMinY, mnMaxY) as float32 values."""
    d = np.ascontiguousarray(dist_coef, np.float32)
    out = np.zeros(4, np.float32)
    rc = _lib.lib().orbm_compute_image_bounds(int(width), int(height), fx, fy, cx, cy,
                                              d.ctypes.data_as(C.c_void_p) if d.size else None, d.size,
                                              out.ctypes.data_as(C.c_void_p))
    _lib.check(rc, "orbm_compute_image_bounds")
    return tuple(out)


def isInFrustum(pos, normal, min_dist, max_dist, nmp, pose, cam, mbf, bounds, viewing_cos_limit, log_scale_factor,
                n_levels, fisheye_rig=False, th=None, scale_factors=None):
    """Frame::isInFrustum (fisheye_rig=False) / isInFrustumChecks (True) over a batch of frames.
    pos, normal [B, M, 3] f32; min_dist, max_dist [B, M] f32 (raw mfMinDistance / mfMaxDistance); nmp [B] i32;
    pose [B, 15] f32 (R row-major, t, camera centre).  With th and scale_factors also returns the
    SearchByProjection window (qr, qmin, qmax).  Returns a dict of tensors."""
    import torch
    B, M = min_dist.shape
    dev = pos.device
    o = dict(in_view=torch.empty((B, M), dtype=torch.uint8, device=dev),
             proj_x=torch.empty((B, M), dtype=torch.float32, device=dev),
             proj_y=torch.empty((B, M), dtype=torch.float32, device=dev),
             proj_xr=torch.empty((B, M), dtype=torch.float32, device=dev),
             depth=torch.empty((B, M), dtype=torch.float32, device=dev),
             view_cos=torch.empty((B, M), dtype=torch.float32, device=dev),
             level=torch.empty((B, M), dtype=torch.int8, device=dev))
    qr = qmin = qmax = None
    sf = None
    if th is not None:
        sf = np.ascontiguousarray(scale_factors, np.float32)
        o["qr"] = torch.empty((B, M), dtype=torch.float32, device=dev)
        o["qmin"] = torch.empty((B, M), dtype=torch.int8, device=dev)
        o["qmax"] = torch.empty((B, M), dtype=torch.int8, device=dev)
        qr, qmin, qmax = o["qr"].data_ptr(), o["qmin"].data_ptr(), o["qmax"].data_ptr()
    rc = _lib.lib().orbm_is_in_frustum(
        pos.data_ptr(), normal.data_ptr(), min_dist.data_ptr(), max_dist.data_ptr(), nmp.data_ptr(), M, B,
        pose.data_ptr(), C.byref(cam), mbf, bounds[0], bounds[1], bounds[2], bounds[3], viewing_cos_limit,
        log_scale_factor, n_levels, int(bool(fisheye_rig)), o["in_view"].data_ptr(), o["proj_x"].data_ptr(),
        o["proj_y"].data_ptr(), o["proj_xr"].data_ptr(), o["depth"].data_ptr(), o["view_cos"].data_ptr(),
        o["level"].data_ptr(), float(th) if th is not None else 0.0,
        sf.ctypes.data_as(C.c_void_p) if sf is not None else None, qr, qmin, qmax, _stream(dev))
    _lib.check(rc, "orbm_is_in_frustum")
    return o
