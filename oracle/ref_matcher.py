"""oracle/_ref/libref_orbmatcher.so — the REFERENCE's own src/ORBmatcher.cc compiled here, unmodified —
TEST INFRASTRUCTURE ONLY.

`make -C oracle ref_matcher` compiles /root/reference/src/ORBmatcher.cc (where it lies, with the reference's
include/ORBmatcher.h) in front of a force-included prelude (oracle/ref_shim_matcher/prelude.h) that declares
stand-ins for Eigen / Sophus / DBoW2::FeatureVector / Frame / KeyFrame / MapPoint with exactly the members the file
touches.  The functions below build those objects from flat arrays and call the reference's ORBmatcher methods;
tests/test_ref_matcher.py compares the results with the oracle's restatements (oracle/orb_oracle.cpp,
oracle/frame_oracle.cpp) on the same inputs — that is what pins matcher rows m2-m9, m12.

/root/reference does not exist on the GPU box: there the prebuilt .so (git-ignored, travels with the snapshot) is
loaded, never rebuilt."""
import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

from . import _p, _fv_args

_DIR = Path(__file__).resolve().parent
_SO = _DIR / "_ref" / "libref_orbmatcher.so"
REFERENCE = Path("/root/reference")
_lib = None


def available():
    return _SO.exists() or (REFERENCE / "src" / "ORBmatcher.cc").exists()


def build(force=False):
    if (REFERENCE / "src" / "ORBmatcher.cc").exists():
        subprocess.check_call(["make", "-C", str(_DIR), "ref_matcher"] + (["-B"] if force else ["-s"]),
                              stdout=subprocess.DEVNULL)
    return _SO if _SO.exists() else None


def lib():
    global _lib
    if _lib is None:
        if not _SO.exists():
            build()
        if not _SO.exists():
            raise FileNotFoundError("oracle/_ref/libref_orbmatcher.so: not built and /root/reference absent")
        _lib = C.CDLL(str(_SO))
    return _lib


def _f32(a):
    return np.ascontiguousarray(a, np.float32)


def _u8(a):
    return None if a is None else np.ascontiguousarray(a, np.uint8)


def _i32(a):
    return None if a is None else np.ascontiguousarray(a, np.int32)


def _pp(a):
    return None if a is None else _p(a)


def search_by_projection_mappoints(kps_l, desc_l, bounds, sf, P, nn_ratio, th, kps_r=None, desc_r=None, u_right=None,
                                   locked=None, l2r=None, r2l=None, far_points=False, th_far=50.0):
    """ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th, bFarPoints, thFarPoints).
    P: dict flags (bit0 inView, bit1 obs>0, bit2 inViewR, bit3 bad), proj [M,4], view_cos, view_cos_r, level,
    level_r, depth, desc.  Returns (holder [N], nmatches)."""
    kl = np.ascontiguousarray(kps_l)
    rig = kps_r is not None
    kr = np.ascontiguousarray(kps_r) if rig else None
    n = kl.size + (kr.size if rig else 0)
    holder = np.zeros(n, np.int32)
    a = dict(dl=_u8(desc_l), dr=_u8(desc_r), b=_f32(bounds), sf=_f32(sf), ur=None if u_right is None else _f32(u_right),
             lk=_u8(locked), l2r=_i32(l2r), r2l=_i32(r2l), fl=_u8(P["flags"]), pj=_f32(P["proj"]), vc=_f32(P["view_cos"]),
             vcr=_f32(P["view_cos_r"]), lv=_i32(P["level"]), lvr=_i32(P["level_r"]), dp=_f32(P["depth"]), qd=_u8(P["desc"]))
    L = lib()
    L.ref_search_by_projection_mappoints.restype = C.c_int
    L.ref_search_by_projection_mappoints.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                                     C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                                     C.c_void_p, C.c_void_p, C.c_int] + [C.c_void_p] * 8 + \
        [C.c_float, C.c_float, C.c_int, C.c_float, C.c_void_p]
    nm = L.ref_search_by_projection_mappoints(_p(kl), _p(a["dl"]), kl.size, _pp(kr), _pp(a["dr"]), kr.size if rig else -1,
                                              _p(a["b"]), _p(a["sf"]), a["sf"].size, _pp(a["ur"]), _pp(a["lk"]),
                                              _pp(a["l2r"]), _pp(a["r2l"]), a["fl"].size, _p(a["fl"]), _p(a["pj"]),
                                              _p(a["vc"]), _p(a["vcr"]), _p(a["lv"]), _p(a["lvr"]), _p(a["dp"]),
                                              _p(a["qd"]), float(nn_ratio), float(th), int(bool(far_points)),
                                              float(th_far), _p(holder))
    return holder, nm


def search_by_projection_lastframe(kps_l, desc_l, bounds, sf, last_flags, uv, last_octave, last_angle, qdesc, th, mono,
                                   check_ori, kps_r=None, desc_r=None, u_right=None, locked=None, mb=0.1, mbf=40.0,
                                   trl_x=0.0, tlc_z=0.0):
    """ORBmatcher::SearchByProjection(Frame& Current, const Frame& Last, th, bMono).  Returns (holder, nmatches)."""
    kl = np.ascontiguousarray(kps_l)
    rig = kps_r is not None
    kr = np.ascontiguousarray(kps_r) if rig else None
    n = kl.size + (kr.size if rig else 0)
    holder = np.zeros(n, np.int32)
    a = dict(dl=_u8(desc_l), dr=_u8(desc_r), b=_f32(bounds), sf=_f32(sf), ur=None if u_right is None else _f32(u_right),
             lk=_u8(locked), fl=_u8(last_flags), uv=_f32(uv), oc=_i32(last_octave), an=_f32(last_angle), qd=_u8(qdesc))
    L = lib()
    L.ref_search_by_projection_lastframe.restype = C.c_int
    L.ref_search_by_projection_lastframe.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                                     C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int] + \
        [C.c_void_p] * 5 + [C.c_float] * 5 + [C.c_int, C.c_int, C.c_void_p]
    nm = L.ref_search_by_projection_lastframe(_p(kl), _p(a["dl"]), kl.size, _pp(kr), _pp(a["dr"]), kr.size if rig else -1,
                                              _p(a["b"]), _p(a["sf"]), a["sf"].size, _pp(a["ur"]), _pp(a["lk"]),
                                              a["fl"].size, _p(a["fl"]), _p(a["uv"]), _p(a["oc"]), _p(a["an"]),
                                              _p(a["qd"]), float(mb), float(mbf), float(trl_x), float(tlc_z), float(th),
                                              int(bool(mono)), int(bool(check_ori)), _p(holder))
    return holder, nm


def search_by_bow_kf_frame(kps1, desc1, holder_flags1, fv1, kps2, desc2, fv2, sf, nn_ratio, check_ori, nleft2=-1):
    k1, k2 = np.ascontiguousarray(kps1), np.ascontiguousarray(kps2)
    a, b = _fv_args(fv1), _fv_args(fv2)
    d1, d2, hf, s = _u8(desc1), _u8(desc2), _u8(holder_flags1), _f32(sf)
    m21 = np.zeros(k2.size, np.int32)
    L = lib()
    L.ref_search_by_bow_kf_frame.restype = C.c_int
    L.ref_search_by_bow_kf_frame.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                             C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p,
                                             C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_float, C.c_int,
                                             C.c_void_p]
    nm = L.ref_search_by_bow_kf_frame(_p(k1), _p(d1), k1.size, _pp(hf), _p(a[0]), _p(a[1]), _p(a[2]), a[0].size, _p(k2),
                                      _p(d2), k2.size, int(nleft2), _p(b[0]), _p(b[1]), _p(b[2]), b[0].size, _p(s), s.size,
                                      float(nn_ratio), int(bool(check_ori)), _p(m21))
    return m21, nm


def search_by_bow_kf_kf(kps1, desc1, hf1, fv1, kps2, desc2, hf2, fv2, sf, nn_ratio, check_ori):
    k1, k2 = np.ascontiguousarray(kps1), np.ascontiguousarray(kps2)
    a, b = _fv_args(fv1), _fv_args(fv2)
    d1, d2, h1, h2, s = _u8(desc1), _u8(desc2), _u8(hf1), _u8(hf2), _f32(sf)
    m12 = np.zeros(k1.size, np.int32)
    L = lib()
    L.ref_search_by_bow_kf_kf.restype = C.c_int
    L.ref_search_by_bow_kf_kf.argtypes = ([C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                           C.c_int] * 2) + [C.c_void_p, C.c_int, C.c_float, C.c_int, C.c_void_p]
    nm = L.ref_search_by_bow_kf_kf(_p(k1), _p(d1), k1.size, _pp(h1), _p(a[0]), _p(a[1]), _p(a[2]), a[0].size, _p(k2), _p(d2),
                                   k2.size, _pp(h2), _p(b[0]), _p(b[1]), _p(b[2]), b[0].size, _p(s), s.size,
                                   float(nn_ratio), int(bool(check_ori)), _p(m12))
    return m12, nm


def search_for_triangulation(kps1, desc1, hf1, ur1, fv1, kps2, desc2, hf2, ur2, fv2, sf, sigma2, pose1, pose2, cam,
                             only_stereo, coarse, check_ori):
    """Returns (match12, nmatches, F12 [9], ep [2]) — F12 / ep as the reference's own geometry produced them."""
    k1, k2 = np.ascontiguousarray(kps1), np.ascontiguousarray(kps2)
    a, b = _fv_args(fv1), _fv_args(fv2)
    d1, d2, h1, h2 = _u8(desc1), _u8(desc2), _u8(hf1), _u8(hf2)
    u1, u2 = (None if ur1 is None else _f32(ur1)), (None if ur2 is None else _f32(ur2))
    s, g, p1, p2, c = _f32(sf), _f32(sigma2), _f32(pose1), _f32(pose2), _f32(cam)
    m12 = np.zeros(k1.size, np.int32)
    F, ep = np.zeros(9, np.float32), np.zeros(2, np.float32)
    L = lib()
    L.ref_search_for_triangulation.restype = C.c_int
    L.ref_search_for_triangulation.argtypes = ([C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                                C.c_void_p, C.c_void_p, C.c_int] * 2) + \
        [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p,
         C.c_void_p, C.c_void_p]
    nm = L.ref_search_for_triangulation(_p(k1), _p(d1), k1.size, _pp(h1), _pp(u1), _p(a[0]), _p(a[1]), _p(a[2]), a[0].size,
                                        _p(k2), _p(d2), k2.size, _pp(h2), _pp(u2), _p(b[0]), _p(b[1]), _p(b[2]), b[0].size,
                                        _p(s), _p(g), s.size, _p(p1), _p(p2), _p(c), int(bool(only_stereo)),
                                        int(bool(coarse)), int(bool(check_ori)), _p(m12), _p(F), _p(ep))
    return m12, nm, F, ep


def search_for_initialization(kps1, desc1, kps2, desc2, bounds, sf, prev_matched, window_size, nn_ratio, check_ori):
    k1, k2 = np.ascontiguousarray(kps1), np.ascontiguousarray(kps2)
    d1, d2, b, s = _u8(desc1), _u8(desc2), _f32(bounds), _f32(sf)
    pm = np.array(prev_matched, np.float32, copy=True).reshape(-1, 2)
    m12 = np.zeros(k1.size, np.int32)
    L = lib()
    L.ref_search_for_initialization.restype = C.c_int
    L.ref_search_for_initialization.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                                C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_float, C.c_int,
                                                C.c_void_p]
    nm = L.ref_search_for_initialization(_p(k1), _p(d1), k1.size, _p(k2), _p(d2), k2.size, _p(b), _p(s), s.size, _p(pm),
                                         int(window_size), float(nn_ratio), int(bool(check_ori)), _p(m12))
    return m12, nm, pm


def search_by_projection_reloc(kps_c, desc_c, bounds, sf, holds_c, kf_flags, uv, max_dist, min_dist, kf_angle, qdesc, th,
                               orb_dist, check_ori):
    kc = np.ascontiguousarray(kps_c)
    a = dict(d=_u8(desc_c), b=_f32(bounds), s=_f32(sf), h=_u8(holds_c), f=_u8(kf_flags), uv=_f32(uv), mx=_f32(max_dist),
             mn=_f32(min_dist), an=_f32(kf_angle), q=_u8(qdesc))
    holder = np.zeros(kc.size, np.int32)
    L = lib()
    L.ref_search_by_projection_reloc.restype = C.c_int
    L.ref_search_by_projection_reloc.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                                 C.c_void_p, C.c_int] + [C.c_void_p] * 6 + [C.c_float, C.c_int, C.c_int,
                                                                                            C.c_void_p]
    nm = L.ref_search_by_projection_reloc(_p(kc), _p(a["d"]), kc.size, _p(a["b"]), _p(a["s"]), a["s"].size, _pp(a["h"]),
                                          a["f"].size, _p(a["f"]), _p(a["uv"]), _p(a["mx"]), _p(a["mn"]), _p(a["an"]),
                                          _p(a["q"]), float(th), int(orb_dist), int(bool(check_ori)), _p(holder))
    return holder, nm


def fuse(kps, desc, holder_flags, u_right, bounds, sf, sigma2, bf, mp_flags, uv, max_dist, min_dist, qdesc, th):
    """ORBmatcher::Fuse(KeyFrame*, vector<MapPoint*>&, th).  Returns (best_idx [M], nFused)."""
    k = np.ascontiguousarray(kps)
    a = dict(d=_u8(desc), h=_u8(holder_flags), u=None if u_right is None else _f32(u_right), b=_f32(bounds), s=_f32(sf),
             g=_f32(sigma2), f=_u8(mp_flags), uv=_f32(uv), mx=_f32(max_dist), mn=_f32(min_dist), q=_u8(qdesc))
    best = np.zeros(a["f"].size, np.int32)
    L = lib()
    L.ref_fuse.restype = C.c_int
    L.ref_fuse.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                           C.c_int, C.c_float, C.c_int] + [C.c_void_p] * 5 + [C.c_float, C.c_void_p]
    nf = L.ref_fuse(_p(k), _p(a["d"]), k.size, _pp(a["h"]), _pp(a["u"]), _p(a["b"]), _p(a["s"]), _p(a["g"]), a["s"].size,
                    float(bf), a["f"].size, _p(a["f"]), _p(a["uv"]), _p(a["mx"]), _p(a["mn"]), _p(a["q"]), float(th),
                    _p(best))
    return best, nf


def fuse_sim3(kps, desc, holder_flags, held_bad, bounds, sf, mp_flags, own_feat, uv, max_dist, min_dist, qdesc, th):
    """ORBmatcher::Fuse(KeyFrame*, Sim3f&, vpPoints, th, vpReplacePoint) (ORBmatcher.cc:1340-1455).
    Returns (best_idx [M], replaced [M], nFused).  A point that replaced through a feature taken over by another query
    point i' reports -3 - i'."""
    k = np.ascontiguousarray(kps)
    a = dict(d=_u8(desc), h=_u8(holder_flags), hb=_u8(held_bad), b=_f32(bounds), s=_f32(sf), f=_u8(mp_flags),
             o=_i32(own_feat), uv=_f32(uv), mx=_f32(max_dist), mn=_f32(min_dist), q=_u8(qdesc))
    M = a["f"].size
    best, repl = np.zeros(M, np.int32), np.zeros(M, np.uint8)
    L = lib()
    L.ref_fuse_sim3.restype = C.c_int
    L.ref_fuse_sim3.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                C.c_int] + [C.c_void_p] * 6 + [C.c_float, C.c_void_p, C.c_void_p]
    nf = L.ref_fuse_sim3(_p(k), _p(a["d"]), k.size, _p(a["h"]), _p(a["hb"]), _p(a["b"]), _p(a["s"]), a["s"].size, M,
                         _p(a["f"]), _p(a["o"]), _p(a["uv"]), _p(a["mx"]), _p(a["mn"]), _p(a["q"]), float(th), _p(best),
                         _p(repl))
    return best, repl, nf


def search_by_projection_sim3(variant, kps, desc, bounds, sf, matched_init, mp_flags, uv, max_dist, min_dist, qdesc, th,
                              ratio_hamming):
    """ORBmatcher::SearchByProjection(KeyFrame*, Sim3, ...) (variant 0: :427-532, 1: :534-646).
    Returns (matched [N], nmatches)."""
    k = np.ascontiguousarray(kps)
    a = dict(d=_u8(desc), b=_f32(bounds), s=_f32(sf), mi=_i32(matched_init), f=_u8(mp_flags), uv=_f32(uv), mx=_f32(max_dist),
             mn=_f32(min_dist), q=_u8(qdesc))
    out = np.zeros(k.size, np.int32)
    L = lib()
    L.ref_search_by_projection_sim3.restype = C.c_int
    L.ref_search_by_projection_sim3.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                                C.c_void_p, C.c_int] + [C.c_void_p] * 5 + [C.c_int, C.c_float, C.c_void_p]
    nm = L.ref_search_by_projection_sim3(int(variant), _p(k), _p(a["d"]), k.size, _p(a["b"]), _p(a["s"]), a["s"].size,
                                         _p(a["mi"]), a["f"].size, _p(a["f"]), _p(a["uv"]), _p(a["mx"]), _p(a["mn"]),
                                         _p(a["q"]), int(th), float(ratio_hamming), _p(out))
    return out, nm


def search_by_sim3(kps1, desc1, hf1, uv1, maxd1, mind1, kps2, desc2, hf2, uv2, maxd2, mind2, bounds, sf, init12, th):
    """ORBmatcher::SearchBySim3 (ORBmatcher.cc:1457-1674).  Returns (match12 [n1], nFound)."""
    k1, k2 = np.ascontiguousarray(kps1), np.ascontiguousarray(kps2)
    a = [_u8(desc1), _u8(hf1), _f32(uv1), _f32(maxd1), _f32(mind1), _u8(desc2), _u8(hf2), _f32(uv2), _f32(maxd2),
         _f32(mind2), _f32(bounds), _f32(sf), _i32(init12)]
    out = np.zeros(k1.size, np.int32)
    L = lib()
    L.ref_search_by_sim3.restype = C.c_int
    L.ref_search_by_sim3.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_float, C.c_void_p]
    nf = L.ref_search_by_sim3(_p(k1), _p(a[0]), k1.size, _p(a[1]), _p(a[2]), _p(a[3]), _p(a[4]), _p(k2), _p(a[5]), k2.size,
                              _p(a[6]), _p(a[7]), _p(a[8]), _p(a[9]), _p(a[10]), _p(a[11]), a[11].size, _p(a[12]),
                              float(th), _p(out))
    return out, nf


def search_for_triangulation_rig(kps1, desc1, nleft1, hf1, fv1, kps2, desc2, nleft2, hf2, fv2, sf, pose1, pose2, pose_rl,
                                 gate, only_stereo, coarse, check_ori):
    """ORBmatcher::SearchForTriangulation on rig keyframes; the geometric gate comes from gate [n1,n2] u8 (table cameras
    that also verify the camera / pose combination they are called with).  Returns (match12, nmatches, bad_calls)."""
    k1, k2 = np.ascontiguousarray(kps1), np.ascontiguousarray(kps2)
    a, b = _fv_args(fv1), _fv_args(fv2)
    d1, d2, h1, h2, s = _u8(desc1), _u8(desc2), _u8(hf1), _u8(hf2), _f32(sf)
    p1, p2, prl, g = _f32(pose1), _f32(pose2), _f32(pose_rl), _u8(gate)
    m12 = np.zeros(k1.size, np.int32)
    bad = C.c_int(0)
    L = lib()
    L.ref_search_for_triangulation_rig.restype = C.c_int
    L.ref_search_for_triangulation_rig.argtypes = ([C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                                    C.c_void_p, C.c_void_p, C.c_int] * 2) + \
        [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p,
         C.POINTER(C.c_int)]
    nm = L.ref_search_for_triangulation_rig(_p(k1), _p(d1), k1.size, int(nleft1), _pp(h1), _p(a[0]), _p(a[1]), _p(a[2]),
                                            a[0].size, _p(k2), _p(d2), k2.size, int(nleft2), _pp(h2), _p(b[0]), _p(b[1]),
                                            _p(b[2]), b[0].size, _p(s), s.size, _p(p1), _p(p2), _p(prl), _p(g),
                                            int(bool(only_stereo)), int(bool(coarse)), int(bool(check_ori)), _p(m12),
                                            C.byref(bad))
    return m12, nm, bad.value
