"""CPU oracle for the ORB front-end hot path — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package.  The product package (orb_slam3_vio_fixes_b200) never does.

`oracle.lib()` returns a ctypes handle on liborb_oracle.so (built by `make -C oracle`),
`oracle.Extractor` wraps the restated ORBextractor, and the free functions wrap the
individually pinned primitives.  See orb_oracle.cpp for the reference file:line map.
"""
import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

_DIR = Path(__file__).resolve().parent
_SO = _DIR / "liborb_oracle.so"
_lib = None

# cv::KeyPoint layout, 28 bytes
KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
RES_DTYPE = np.dtype([("best_idx", "<i4"), ("best_dist", "<i2"), ("second_dist", "<i2"),
                      ("best_level", "i1"), ("second_level", "i1"), ("pad", "<i2")])
STRICT_FP = 1


def build(force=False):
    srcs = [_DIR / "orb_oracle.cpp", _DIR / "frame_oracle.cpp"]
    if force or not _SO.exists() or any(_SO.stat().st_mtime < s.stat().st_mtime for s in srcs):
        subprocess.check_call(["make", "-C", str(_DIR), "-B" if force else "-s"],
                              stdout=subprocess.DEVNULL)
    return _SO


def lib():
    global _lib
    if _lib is None:
        if not _SO.exists():
            build()
        L = C.CDLL(str(_SO))
        L.oracle_create.restype = C.c_void_p
        L.oracle_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int]
        L.oracle_destroy.argtypes = [C.c_void_p]
        L.oracle_extract.restype = C.c_int
        L.oracle_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int,
                                     C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_int)]
        L.oracle_tables.argtypes = [C.c_void_p] + [C.c_void_p] * 6
        L.oracle_level_size.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.oracle_get_pyramid.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_int]
        L.oracle_level_keypoints.restype = C.c_int
        L.oracle_level_keypoints.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.oracle_resize_linear.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int]
        L.oracle_gaussian_blur7.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.oracle_border101.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.oracle_fast_atan2.restype = C.c_float
        L.oracle_fast_atan2.argtypes = [C.c_float, C.c_float]
        L.oracle_fast_atan2_array.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.oracle_fast.restype = C.c_int
        L.oracle_fast.argtypes = [C.c_void_p] + [C.c_int] * 7 + [C.c_void_p, C.c_int]
        L.oracle_distribute.restype = C.c_int
        L.oracle_distribute.argtypes = [C.c_void_p] + [C.c_int] * 6 + [C.c_void_p]
        L.oracle_sincos_deg.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_long]
        L.oracle_sincos_bits.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]
        L.oracle_descriptor_distance.restype = C.c_int
        L.oracle_descriptor_distance.argtypes = [C.c_void_p, C.c_void_p]
        L.oracle_argmin2_csr.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_int, C.c_void_p]
        L.oracle_grid_queries.restype = C.c_int
        L.oracle_grid_queries.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float,
                                          C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                          C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        L.oracle_stereo_matches.restype = C.c_int
        L.oracle_stereo_matches.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                            C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                            C.c_void_p, C.c_void_p, C.c_float, C.c_float,
                                            C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.oracle_knn2_dense.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.oracle_bench_extract.restype = C.c_double
        L.oracle_bench_extract.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                           C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                           C.POINTER(C.c_int)]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class Extractor:
    """Restated ORB_SLAM3::ORBextractor (ctor args as ORBextractor.h:49-50)."""

    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, flags=0):
        self.L = lib()
        self.nfeatures, self.nlevels = nfeatures, nlevels
        self.h = self.L.oracle_create(nfeatures, scale_factor, nlevels, ini_th, min_th, flags)
        sf, inv, s2, is2 = (np.zeros(nlevels, np.float32) for _ in range(4))
        nf = np.zeros(nlevels, np.int32)
        um = np.zeros(16, np.int32)
        self.L.oracle_tables(self.h, _p(sf), _p(inv), _p(s2), _p(is2), _p(nf), _p(um))
        self.scale_factors, self.inv_scale_factors = sf, inv
        self.level_sigma2, self.inv_level_sigma2 = s2, is2
        self.features_per_level, self.umax = nf, um

    def __del__(self):
        if getattr(self, "h", None):
            self.L.oracle_destroy(self.h)
            self.h = None

    def extract(self, img, lap=(0, 0)):
        """operator(): returns (keypoints[KP_DTYPE], descriptors[n,32] u8, monoIndex)."""
        img = np.ascontiguousarray(img, np.uint8)
        hgt, w = img.shape
        # a level returns up to max(N+3, 4*nIni) keypoints, nIni = round(width/height) root nodes (ORBextractor.cc:559)
        cap = self.nfeatures + self.nlevels * (3 + 4 * (w // max(hgt, 1) + 2)) + 64
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = C.c_int(0)
        mono = self.L.oracle_extract(self.h, _p(img), w, hgt, w, lap[0], lap[1], _p(kps), _p(desc), cap,
                                     C.byref(n))
        assert mono != -2, "capacity"
        return kps[:n.value].copy(), desc[:n.value].copy(), mono

    def level_size(self, level):
        w, h = C.c_int(), C.c_int()
        assert self.L.oracle_level_size(self.h, level, C.byref(w), C.byref(h)) == 0
        return w.value, h.value

    def pyramid(self, level, with_border=False):
        w, h = self.level_size(level)
        if with_border:
            w, h = w + 38, h + 38
        out = np.zeros((h, w), np.uint8)
        self.L.oracle_get_pyramid(self.h, level, _p(out), w, int(with_border))
        return out

    def level_keypoints(self, level, stage):
        n = self.L.oracle_level_keypoints(self.h, level, stage, None, 0)
        out = np.zeros(n, KP_DTYPE)
        if n:
            self.L.oracle_level_keypoints(self.h, level, stage, _p(out), n)
        return out


def resize_linear(src, dw, dh):
    src = np.ascontiguousarray(src, np.uint8)
    out = np.zeros((dh, dw), np.uint8)
    lib().oracle_resize_linear(_p(src), src.shape[1], src.shape[0], _p(out), dw, dh)
    return out


def gaussian_blur7(src):
    src = np.ascontiguousarray(src, np.uint8)
    out = np.zeros_like(src)
    lib().oracle_gaussian_blur7(_p(src), src.shape[1], src.shape[0], _p(out))
    return out


def border101(src, b=19):
    src = np.ascontiguousarray(src, np.uint8)
    out = np.zeros((src.shape[0] + 2 * b, src.shape[1] + 2 * b), np.uint8)
    lib().oracle_border101(_p(src), src.shape[1], src.shape[0], b, _p(out))
    return out


def fast_atan2(y, x):
    y = np.ascontiguousarray(y, np.float32)
    x = np.ascontiguousarray(x, np.float32)
    out = np.zeros_like(y)
    lib().oracle_fast_atan2_array(_p(y), _p(x), _p(out), y.size)
    return out


def fast(img, roi, threshold):
    """cv::FAST(img[y0:y1, x0:x1], th, nms=True) → int array [n,3] (x, y, score), ROI-relative."""
    img = np.ascontiguousarray(img, np.uint8)
    x0, y0, x1, y1 = roi
    cap = max(1, (x1 - x0) * (y1 - y0))
    out = np.zeros((cap, 3), np.int32)
    n = lib().oracle_fast(_p(img), img.shape[1], img.shape[0], x0, y0, x1, y1, threshold, _p(out), cap)
    return out[:n].copy()


def distribute(xyr, min_x, max_x, min_y, max_y, n_target):
    xyr = np.ascontiguousarray(xyr, np.float32).reshape(-1, 3)
    out = np.zeros((xyr.shape[0] + 8, 3), np.float32)
    n = lib().oracle_distribute(_p(xyr), xyr.shape[0], min_x, max_x, min_y, max_y, n_target, _p(out))
    return out[:n].copy()


def sincos_deg(angle_deg):
    a = np.ascontiguousarray(angle_deg, np.float32)
    c = np.zeros_like(a)
    s = np.zeros_like(a)
    lib().oracle_sincos_deg(_p(a), _p(c), _p(s), a.size)
    return c, s


def sincos_bits(lo, count, stride=1):
    c = np.zeros(count, np.float32)
    s = np.zeros(count, np.float32)
    lib().oracle_sincos_bits(lo, count, stride, _p(c), _p(s))
    return c, s


def sincos_checksum(lo, count, chunk=1 << 22, threads=None):
    """Per-chunk checksums of glibc sincosf over `count` float bit patterns from `lo` (see orb_oracle.cpp)."""
    L = lib()
    L.oracle_sincos_checksum.argtypes = [C.c_uint32, C.c_uint64, C.c_uint32, C.c_void_p, C.c_int]
    out = np.zeros((count + chunk - 1) // chunk, np.uint64)
    L.oracle_sincos_checksum(lo, count, chunk, _p(out), threads or host_threads())
    return out


def descriptor_distance(a, b):
    a = np.ascontiguousarray(a, np.uint8)
    b = np.ascontiguousarray(b, np.uint8)
    return lib().oracle_descriptor_distance(_p(a), _p(b))


def argmin2_csr(q, r, cand_off, cand_idx, ref_level=None, tie_rule=0):
    q = np.ascontiguousarray(q, np.uint8)
    r = np.ascontiguousarray(r, np.uint8)
    cand_off = np.ascontiguousarray(cand_off, np.int32)
    cand_idx = np.ascontiguousarray(cand_idx, np.int32)
    lv = None if ref_level is None else np.ascontiguousarray(ref_level, np.int8)
    out = np.zeros(q.shape[0], RES_DTYPE)
    lib().oracle_argmin2_csr(_p(q), q.shape[0], _p(r), r.shape[0], _p(cand_off), _p(cand_idx),
                             None if lv is None else _p(lv), tie_rule, _p(out))
    return out


def grid_queries(kps, bounds, qx, qy, qr, qmin, qmax):
    """AssignFeaturesToGrid + GetFeaturesInArea → (cand_off[nq+1], cand_idx)."""
    kps = np.ascontiguousarray(kps)
    qx, qy, qr = (np.ascontiguousarray(v, np.float32) for v in (qx, qy, qr))
    qmin, qmax = (np.ascontiguousarray(v, np.int32) for v in (qmin, qmax))
    nq = qx.size
    off = np.zeros(nq + 1, np.int32)
    cap = max(1, nq * 64)
    while True:
        idx = np.zeros(cap, np.int32)
        tot = lib().oracle_grid_queries(_p(kps), kps.size, *[float(b) for b in bounds], _p(qx), _p(qy), _p(qr),
                                        _p(qmin), _p(qmax), nq, _p(off), _p(idx), cap)
        if tot <= cap:
            return off, idx[:tot].copy()
        cap = tot


def stereo_matches(kps_l, desc_l, kps_r, desc_r, scale_factors, inv_scale_factors, pyr_l, pyr_r, mb, mbf):
    nl, nr = kps_l.size, kps_r.size
    nlev = len(pyr_l)
    pl = [np.ascontiguousarray(p, np.uint8) for p in pyr_l]
    pr = [np.ascontiguousarray(p, np.uint8) for p in pyr_r]
    PL = (C.c_void_p * nlev)(*[p.ctypes.data for p in pl])
    PR = (C.c_void_p * nlev)(*[p.ctypes.data for p in pr])
    pw = np.array([p.shape[1] for p in pl], np.int32)
    ph = np.array([p.shape[0] for p in pl], np.int32)
    bi = np.zeros(nl, np.int32)
    bd = np.zeros(nl, np.int32)
    ur = np.zeros(nl, np.float32)
    dp = np.zeros(nl, np.float32)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    isf = np.ascontiguousarray(inv_scale_factors, np.float32)
    kept = lib().oracle_stereo_matches(_p(np.ascontiguousarray(kps_l)), _p(np.ascontiguousarray(desc_l)), nl,
                                       _p(np.ascontiguousarray(kps_r)), _p(np.ascontiguousarray(desc_r)), nr,
                                       _p(sf), _p(isf), nlev, PL, PR, _p(pw), _p(ph), mb, mbf,
                                       _p(bi), _p(bd), _p(ur), _p(dp))
    return bi, bd, ur, dp, kept


def search_by_projection(kps, desc, bounds, qx, qy, qr, qmin, qmax, qdesc, mode, nn_ratio=0.6,
                         check_orientation=True, u_right=None, feat_locked=None, q_ur=None, q_flags=None,
                         q_angle=None, th_dist=100):
    """ORBmatcher::SearchByProjection (mode 0: MapPoints variant, mode 1: last-frame variant), one frame.
    Returns (match [nq], feat_query [n], nmatches)."""
    kps = np.ascontiguousarray(kps)
    n, nq = kps.size, len(qx)
    f32 = lambda a: None if a is None else np.ascontiguousarray(a, np.float32)
    u8 = lambda a: None if a is None else np.ascontiguousarray(a, np.uint8)
    i8 = lambda a: np.ascontiguousarray(a, np.int8)
    args = dict(desc=u8(desc), u_right=f32(u_right), feat_locked=u8(feat_locked), qx=f32(qx), qy=f32(qy), qr=f32(qr),
                qmin=i8(qmin), qmax=i8(qmax), qdesc=u8(qdesc), q_ur=f32(q_ur), q_flags=u8(q_flags),
                q_angle=f32(q_angle))
    pp = lambda a: None if a is None else _p(a)
    match = np.zeros(nq, np.int32)
    fq = np.zeros(n, np.int32)
    L = lib()
    L.oracle_search_by_projection.restype = C.c_int
    L.oracle_search_by_projection.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_float,
                                              C.c_float] + [C.c_void_p] * 8 + [C.c_int] + [C.c_void_p] * 3 + \
                                             [C.c_int, C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    nm = L.oracle_search_by_projection(_p(kps), pp(args["desc"]), n, bounds[0], bounds[1], bounds[2], bounds[3],
                                       pp(args["u_right"]), pp(args["feat_locked"]), pp(args["qx"]), pp(args["qy"]),
                                       pp(args["qr"]), pp(args["qmin"]), pp(args["qmax"]), pp(args["qdesc"]), nq,
                                       pp(args["q_ur"]), pp(args["q_flags"]), pp(args["q_angle"]), int(mode),
                                       float(nn_ratio), int(bool(check_orientation)), int(th_dist), _p(match), _p(fq))
    return match, fq, nm


def fuse_candidates(kps, desc, bounds, inv_sigma2, qx, qy, qr, qlevel, qdesc, u_right=None, q_ur=None):
    """Candidate scoring of ORBmatcher::Fuse, one keyframe.  Returns (best_idx, best_dist)."""
    kps = np.ascontiguousarray(kps)
    f32 = lambda a: None if a is None else np.ascontiguousarray(a, np.float32)
    d, qd = np.ascontiguousarray(desc, np.uint8), np.ascontiguousarray(qdesc, np.uint8)
    ur, sg, x, y, r, qu = f32(u_right), f32(inv_sigma2), f32(qx), f32(qy), f32(qr), f32(q_ur)
    lv = np.ascontiguousarray(qlevel, np.int8)
    pp = lambda a: None if a is None else _p(a)
    nq = len(x)
    bi, bd = np.zeros(nq, np.int32), np.zeros(nq, np.int32)
    L = lib()
    L.oracle_fuse_candidates.restype = None
    L.oracle_fuse_candidates.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float] + \
                                        [C.c_void_p] * 8 + [C.c_int, C.c_void_p, C.c_void_p]
    L.oracle_fuse_candidates(_p(kps), _p(d), kps.size, bounds[0], bounds[1], bounds[2], bounds[3], pp(ur), _p(sg),
                             _p(x), _p(y), _p(r), pp(qu), _p(lv), _p(qd), nq, _p(bi), _p(bd))
    return bi, bd


def search_for_initialization(kps1, desc1, kps2, desc2, bounds, prev_matched, window_size=100, nn_ratio=0.9,
                              check_orientation=True):
    """ORBmatcher::SearchForInitialization.  Returns (match12, nmatches, updated prev_matched)."""
    kps1, kps2 = np.ascontiguousarray(kps1), np.ascontiguousarray(kps2)
    d1, d2 = np.ascontiguousarray(desc1, np.uint8), np.ascontiguousarray(desc2, np.uint8)
    pm = np.array(prev_matched, np.float32, copy=True).reshape(-1, 2)
    m12 = np.zeros(kps1.size, np.int32)
    L = lib()
    L.oracle_search_for_initialization.restype = C.c_int
    L.oracle_search_for_initialization.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                                   C.c_float, C.c_float, C.c_float, C.c_float, C.c_void_p, C.c_int,
                                                   C.c_float, C.c_int, C.c_void_p]
    nm = L.oracle_search_for_initialization(_p(kps1), _p(d1), kps1.size, _p(kps2), _p(d2), kps2.size, bounds[0],
                                            bounds[1], bounds[2], bounds[3], _p(pm), int(window_size),
                                            float(nn_ratio), int(bool(check_orientation)), _p(m12))
    return m12, nm, pm


def rgbd_depth(kps, kps_un, depth_img, mbf):
    """Frame::ComputeStereoFromRGBD.  Returns (uRight, depth)."""
    kps = np.ascontiguousarray(kps)
    kps_un = np.ascontiguousarray(kps_un)
    d = np.ascontiguousarray(depth_img, np.float32)
    ur = np.zeros(kps.size, np.float32)
    dp = np.zeros(kps.size, np.float32)
    L = lib()
    L.oracle_rgbd_depth.restype = None
    L.oracle_rgbd_depth.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_float,
                                    C.c_void_p, C.c_void_p]
    L.oracle_rgbd_depth(_p(kps), _p(kps_un), kps.size, _p(d), d.shape[1], d.shape[0], mbf, _p(ur), _p(dp))
    return ur, dp


def knn2_dense(q, r):
    q = np.ascontiguousarray(q, np.uint8)
    r = np.ascontiguousarray(r, np.uint8)
    idx = np.zeros((q.shape[0], 2), np.int32)
    dist = np.zeros((q.shape[0], 2), np.int32)
    lib().oracle_knn2_dense(_p(q), q.shape[0], _p(r), r.shape[0], _p(idx), _p(dist))
    return idx, dist


def _fv_args(fv):
    node = np.ascontiguousarray(fv[0], np.int32)
    off = np.ascontiguousarray(fv[1], np.int32)
    idx = np.ascontiguousarray(fv[2], np.int32)
    return node, off, idx


def search_by_bow(mode, kps1, desc1, fv1, kps2, desc2, fv2, nn_ratio=0.6, check_orientation=True, valid1=None,
                  valid2=None):
    """ORBmatcher::SearchByBoW (mode 0: KeyFrame x Frame :223-425, mode 1: KeyFrame x KeyFrame :765-905).
    fv = (node ids ascending, off, idx).  Returns (match12, match21, nmatches)."""
    kps1, kps2 = np.ascontiguousarray(kps1), np.ascontiguousarray(kps2)
    d1, d2 = np.ascontiguousarray(desc1, np.uint8), np.ascontiguousarray(desc2, np.uint8)
    a, b = _fv_args(fv1), _fv_args(fv2)
    v1 = None if valid1 is None else np.ascontiguousarray(valid1, np.uint8)
    v2 = None if valid2 is None else np.ascontiguousarray(valid2, np.uint8)
    pp = lambda x: None if x is None else _p(x)
    m12 = np.zeros(kps1.size, np.int32)
    m21 = np.zeros(kps2.size, np.int32)
    L = lib()
    L.oracle_search_by_bow.restype = C.c_int
    L.oracle_search_by_bow.argtypes = [C.c_int] + ([C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                                    C.c_void_p, C.c_void_p, C.c_int] * 2) + \
                                      [C.c_float, C.c_int, C.c_void_p, C.c_void_p]
    nm = L.oracle_search_by_bow(int(mode), _p(kps1), _p(d1), kps1.size, pp(v1), _p(a[0]), _p(a[1]), _p(a[2]),
                                a[0].size, _p(kps2), _p(d2), kps2.size, pp(v2), _p(b[0]), _p(b[1]), _p(b[2]),
                                b[0].size, float(nn_ratio), int(bool(check_orientation)), _p(m12), _p(m21))
    return m12, m21, nm


def search_for_triangulation(kps1, desc1, fv1, kps2, desc2, fv2, F12, ep, scale2, sigma2_2, only_stereo=False,
                             coarse=False, check_orientation=True, has_mp1=None, has_mp2=None, u_right1=None,
                             u_right2=None):
    """ORBmatcher::SearchForTriangulation (:907-1146), pinhole single camera.  Returns (match12, nmatches)."""
    kps1, kps2 = np.ascontiguousarray(kps1), np.ascontiguousarray(kps2)
    d1, d2 = np.ascontiguousarray(desc1, np.uint8), np.ascontiguousarray(desc2, np.uint8)
    a, b = _fv_args(fv1), _fv_args(fv2)
    u8 = lambda x: None if x is None else np.ascontiguousarray(x, np.uint8)
    f32 = lambda x: None if x is None else np.ascontiguousarray(x, np.float32)
    h1, h2, u1, u2 = u8(has_mp1), u8(has_mp2), f32(u_right1), f32(u_right2)
    F, e, sc, sg = f32(np.zeros(9) if F12 is None else F12), f32(ep), f32(scale2), f32(sigma2_2)
    pp = lambda x: None if x is None else _p(x)
    m12 = np.zeros(kps1.size, np.int32)
    L = lib()
    L.oracle_search_for_triangulation.restype = C.c_int
    side = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
    L.oracle_search_for_triangulation.argtypes = side * 2 + [C.c_void_p] * 4 + [C.c_int] * 3 + [C.c_void_p]
    nm = L.oracle_search_for_triangulation(_p(kps1), _p(d1), kps1.size, pp(h1), pp(u1), _p(a[0]), _p(a[1]), _p(a[2]),
                                           a[0].size, _p(kps2), _p(d2), kps2.size, pp(h2), pp(u2), _p(b[0]),
                                           _p(b[1]), _p(b[2]), b[0].size, _p(F), _p(e), _p(sc), _p(sg),
                                           int(bool(only_stereo)), int(bool(coarse)), int(bool(check_orientation)),
                                           _p(m12))
    return m12, nm


def bow_transform(desc, voc, levelsup=4):
    """DBoW2 transform.  voc = dict(child_off, child_idx, node_desc, node_word, node_weight, L).
    Returns dict(feat_word, feat_node, feat_weight, bow_word, bow_val, fv=(node, off, idx))."""
    d = np.ascontiguousarray(desc, np.uint8)
    n = d.shape[0]
    co = np.ascontiguousarray(voc["child_off"], np.int32)
    ci = np.ascontiguousarray(voc["child_idx"], np.int32)
    nd = np.ascontiguousarray(voc["node_desc"], np.uint8)
    nw = np.ascontiguousarray(voc["node_word"], np.int32)
    wt = np.ascontiguousarray(voc["node_weight"], np.float64)
    fw, fn, fwt = np.zeros(n, np.int32), np.zeros(n, np.int32), np.zeros(n, np.float64)
    bw, bv = np.zeros(n, np.int32), np.zeros(n, np.float64)
    node, off, idx = np.zeros(n, np.int32), np.zeros(n + 1, np.int32), np.zeros(n, np.int32)
    nn = C.c_int(0)
    L = lib()
    L.oracle_bow_transform.restype = C.c_int
    L.oracle_bow_transform.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 5 + [C.c_int, C.c_int] + \
                                      [C.c_void_p] * 8 + [C.POINTER(C.c_int)]
    k = L.oracle_bow_transform(_p(d), n, _p(co), _p(ci), _p(nd), _p(nw), _p(wt), int(voc["L"]), int(levelsup),
                               _p(fw), _p(fn), _p(fwt), _p(bw), _p(bv), _p(node), _p(off), _p(idx), C.byref(nn))
    nn = nn.value
    return {"feat_word": fw, "feat_node": fn, "feat_weight": fwt, "bow_word": bw[:k].copy(), "bow_val": bv[:k].copy(),
            "fv": (node[:nn].copy(), off[:nn + 1].copy(), idx[:off[nn]].copy())}


def distinctive_descriptor(desc):
    """MapPoint::ComputeDistinctiveDescriptors core: index of the least-median row."""
    d = np.ascontiguousarray(desc, np.uint8)
    L = lib()
    L.oracle_distinctive_descriptor.restype = C.c_int
    L.oracle_distinctive_descriptor.argtypes = [C.c_void_p, C.c_int]
    return L.oracle_distinctive_descriptor(_p(d), d.shape[0])


from orb_slam3_vio_fixes_b200.synthetic import make_vocabulary  # noqa: E402,F401  (synthetic input, not a checker)


def undistort_keypoints(kps, fx, fy, cx, cy, dist_coef):
    """Frame::UndistortKeyPoints (cv::undistortPoints, R = I, P = K), one frame.  Returns mvKeysUn."""
    kps = np.ascontiguousarray(kps)
    d = np.ascontiguousarray(dist_coef, np.float32)
    out = np.zeros_like(kps)
    L = lib()
    L.oracle_undistort_keypoints.restype = None
    L.oracle_undistort_keypoints.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float,
                                             C.c_void_p, C.c_int, C.c_void_p]
    L.oracle_undistort_keypoints(_p(kps), kps.size, fx, fy, cx, cy, _p(d), d.size, _p(out))
    return out


def is_in_frustum(pos, normal, min_dist, max_dist, pose, cam, mbf, bounds, cos_limit, log_sf, n_levels,
                  fisheye_rig=False, f64=False):
    """Frame::isInFrustum / isInFrustumChecks + PredictScale for one camera pose.
    pose = 15 floats (R row-major, t, centre); cam = (model, fx, fy, cx, cy, k0..k3).
    Returns dict(in_view, proj_x, proj_y, proj_xr, depth, view_cos, level)."""
    pos = np.ascontiguousarray(pos, np.float32); normal = np.ascontiguousarray(normal, np.float32)
    mn = np.ascontiguousarray(min_dist, np.float32); mx = np.ascontiguousarray(max_dist, np.float32)
    pose = np.ascontiguousarray(pose, np.float32)
    camf = np.ascontiguousarray(list(cam[1:]) + [0.0] * (9 - len(cam)), np.float32)
    n = mn.size
    o = dict(in_view=np.zeros(n, np.uint8), proj_x=np.zeros(n, np.float32), proj_y=np.zeros(n, np.float32),
             proj_xr=np.zeros(n, np.float32), depth=np.zeros(n, np.float32), view_cos=np.zeros(n, np.float32),
             level=np.zeros(n, np.int8))
    L = lib()
    L.oracle_is_in_frustum.restype = None
    L.oracle_is_in_frustum.argtypes = [C.c_void_p] * 4 + [C.c_int, C.c_void_p, C.c_int, C.c_void_p] + \
        [C.c_float] * 7 + [C.c_int] * 3 + [C.c_void_p] * 7
    L.oracle_is_in_frustum(_p(pos), _p(normal), _p(mn), _p(mx), n, _p(pose), int(cam[0]), _p(camf), mbf, bounds[0],
                           bounds[1], bounds[2], bounds[3], cos_limit, log_sf, n_levels, int(bool(fisheye_rig)),
                           int(bool(f64)), _p(o["in_view"]), _p(o["proj_x"]), _p(o["proj_y"]), _p(o["proj_xr"]),
                           _p(o["depth"]), _p(o["view_cos"]), _p(o["level"]))
    return o


def search_by_bow_rig(kps1, desc1, fv1, kps2, desc2, nleft2, fv2, nn_ratio=0.6, check_orientation=True, valid1=None):
    """SearchByBoW(KeyFrame*, Frame&) with F.Nleft != -1.  Returns (match12_left, match12_right, match21, nmatches)."""
    kps1, kps2 = np.ascontiguousarray(kps1), np.ascontiguousarray(kps2)
    d1, d2 = np.ascontiguousarray(desc1, np.uint8), np.ascontiguousarray(desc2, np.uint8)
    a, b = _fv_args(fv1), _fv_args(fv2)
    v1 = None if valid1 is None else np.ascontiguousarray(valid1, np.uint8)
    m12l, m12r = np.zeros(kps1.size, np.int32), np.zeros(kps1.size, np.int32)
    m21 = np.zeros(kps2.size, np.int32)
    L = lib()
    L.oracle_search_by_bow_fisheye.restype = C.c_int
    L.oracle_search_by_bow_fisheye.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                               C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                               C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_int,
                                               C.c_void_p, C.c_void_p, C.c_void_p]
    nm = L.oracle_search_by_bow_fisheye(_p(kps1), _p(d1), kps1.size, None if v1 is None else _p(v1), _p(a[0]), _p(a[1]),
                                        _p(a[2]), a[0].size, _p(kps2), _p(d2), kps2.size, int(nleft2), _p(b[0]), _p(b[1]),
                                        _p(b[2]), b[0].size, float(nn_ratio), int(bool(check_orientation)), _p(m12l),
                                        _p(m12r), _p(m21))
    return m12l, m12r, m21, nm


def search_by_projection_rig(mode, kps_l, desc_l, kps_r, desc_r, bounds, q, nn_ratio=0.6, check_orientation=True,
                             locked_l=None, locked_r=None, l2r=None, r2l=None):
    """SearchByProjection on a two-camera Frame, one frame (local numbering per camera).  q: dict of numpy arrays
    qx,qy,qr,qmin,qmax,qx_r,qy_r,qr_r,qmin_r,qmax_r,qdesc and optional flags, angle.
    Returns dict(match_l, match_r, feat_query_l, feat_query_r, nmatches)."""
    kps_l, kps_r = np.ascontiguousarray(kps_l), np.ascontiguousarray(kps_r)
    nl, nr, nq = kps_l.size, kps_r.size, len(q["qx"])
    f32 = lambda a: np.ascontiguousarray(a, np.float32)
    i8 = lambda a: np.ascontiguousarray(a, np.int8)
    u8 = lambda a: None if a is None else np.ascontiguousarray(a, np.uint8)
    i32 = lambda a, n: np.full(n, -1, np.int32) if a is None else np.ascontiguousarray(a, np.int32)
    pp = lambda a: None if a is None else _p(a)
    hold = dict(dl=u8(desc_l), dr=u8(desc_r), ll=u8(locked_l), lr=u8(locked_r), l2r=i32(l2r, nl), r2l=i32(r2l, nr),
                qx=f32(q["qx"]), qy=f32(q["qy"]), qr=f32(q["qr"]), qmin=i8(q["qmin"]), qmax=i8(q["qmax"]),
                qxr=f32(q["qx_r"]), qyr=f32(q["qy_r"]), qrr=f32(q["qr_r"]), qminr=i8(q["qmin_r"]), qmaxr=i8(q["qmax_r"]),
                qd=u8(q["qdesc"]), fl=u8(q.get("flags")), ang=None if q.get("angle") is None else f32(q["angle"]))
    o = dict(match_l=np.zeros(nq, np.int32), match_r=np.zeros(nq, np.int32), feat_query_l=np.zeros(nl, np.int32),
             feat_query_r=np.zeros(nr, np.int32))
    L = lib()
    L.oracle_search_by_projection_rig.restype = C.c_int
    L.oracle_search_by_projection_rig.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                                  C.c_int] + [C.c_float] * 4 + [C.c_void_p] * 15 + [C.c_int] + \
        [C.c_void_p] * 2 + [C.c_float, C.c_int] + [C.c_void_p] * 4
    o["nmatches"] = L.oracle_search_by_projection_rig(
        int(mode), _p(kps_l), pp(hold["dl"]), nl, _p(kps_r), pp(hold["dr"]), nr, bounds[0], bounds[1], bounds[2],
        bounds[3], pp(hold["ll"]), pp(hold["lr"]), _p(hold["l2r"]), _p(hold["r2l"]), _p(hold["qx"]), _p(hold["qy"]),
        _p(hold["qr"]), _p(hold["qmin"]), _p(hold["qmax"]), _p(hold["qxr"]), _p(hold["qyr"]), _p(hold["qrr"]),
        _p(hold["qminr"]), _p(hold["qmaxr"]), _p(hold["qd"]), nq, pp(hold["fl"]), pp(hold["ang"]), float(nn_ratio),
        int(bool(check_orientation)), _p(o["match_l"]), _p(o["match_r"]), _p(o["feat_query_l"]),
        _p(o["feat_query_r"]))
    return o


def search_for_triangulation_rig(kps1, desc1, fv1, kps2, desc2, fv2, gate, only_stereo=False, coarse=False,
                                 check_orientation=True, has_mp1=None, has_mp2=None):
    """SearchForTriangulation on a two-camera rig; gate [n1,n2] u8 = epipolarConstrain outcome per pair (evaluated by
    the caller).  Returns (match12, nmatches)."""
    kps1, kps2 = np.ascontiguousarray(kps1), np.ascontiguousarray(kps2)
    d1, d2 = np.ascontiguousarray(desc1, np.uint8), np.ascontiguousarray(desc2, np.uint8)
    a, b = _fv_args(fv1), _fv_args(fv2)
    g = np.ascontiguousarray(gate, np.uint8)
    h1 = None if has_mp1 is None else np.ascontiguousarray(has_mp1, np.uint8)
    h2 = None if has_mp2 is None else np.ascontiguousarray(has_mp2, np.uint8)
    pp = lambda x: None if x is None else _p(x)
    m12 = np.zeros(kps1.size, np.int32)
    L = lib()
    L.oracle_search_for_triangulation_rig.restype = C.c_int
    L.oracle_search_for_triangulation_rig.argtypes = ([C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                                       C.c_void_p, C.c_int] * 2) + [C.c_void_p, C.c_int, C.c_int, C.c_int,
                                                                                    C.c_void_p]
    nm = L.oracle_search_for_triangulation_rig(_p(kps1), _p(d1), kps1.size, pp(h1), _p(a[0]), _p(a[1]), _p(a[2]),
                                               a[0].size, _p(kps2), _p(d2), kps2.size, pp(h2), _p(b[0]), _p(b[1]),
                                               _p(b[2]), b[0].size, _p(g), int(bool(only_stereo)), int(bool(coarse)),
                                               int(bool(check_orientation)), _p(m12))
    return m12, nm


def bench_extract(frames, nfeatures, scale_factor, nlevels, ini_th, min_th, lap, threads, flags=0):
    """Times the oracle over frames[n,h,w]; returns (seconds, total keypoints)."""
    frames = np.ascontiguousarray(frames, np.uint8)
    n, h, w = frames.shape
    tot = C.c_int(0)
    sec = lib().oracle_bench_extract(nfeatures, scale_factor, nlevels, ini_th, min_th, flags, _p(frames), n, w, h,
                                     lap[0], lap[1], threads, C.byref(tot))
    return sec, tot.value


def host_threads():
    return len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)


def bench_pipeline(frames, nfeatures, scale_factor, nlevels, ini_th, min_th, lap, shift_xy, th, threads,
                   do_match=True, flags=0, stages=None):
    """CPU baseline of bench.py's step (extract all frames + frame-to-frame window matching).
    Returns (seconds, total keypoints, total matches).  `stages`: optional float64[8] that receives the per-stage
    seconds summed over the workers (pyramid, FAST, octree, orientation, blur, descriptors) and the wall seconds of
    the extract and match phases."""
    L = lib()
    L.oracle_bench_pipeline.restype = C.c_double
    L.oracle_bench_pipeline.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int,
                                        C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_float, C.c_int, C.c_int,
                                        C.POINTER(C.c_long), C.POINTER(C.c_long), C.c_void_p]
    frames = np.ascontiguousarray(frames, np.uint8)
    shift_xy = np.ascontiguousarray(shift_xy, np.float32)
    n, h, w = frames.shape
    tk, tm = C.c_long(0), C.c_long(0)
    sec = L.oracle_bench_pipeline(nfeatures, scale_factor, nlevels, ini_th, min_th, flags, _p(frames), n, w, h,
                                  lap[0], lap[1], _p(shift_xy), th, threads, int(do_match), C.byref(tk), C.byref(tm),
                                  stages.ctypes.data if stages is not None else None)
    return sec, tk.value, tm.value
