"""oracle/_ref — the REFERENCE's own ORBextractor.cc compiled here, unmodified — TEST INFRASTRUCTURE ONLY.

`make -C oracle ref` compiles /root/reference/src/ORBextractor.cc (where it lies) over the OpenCV header
shim in oracle/ref_shim/ into oracle/_ref/libref_orbextractor.so.  The image primitives the reference
calls (cv::resize / copyMakeBorder / GaussianBlur / FAST / fastAtan2) go through a function table:

* default: the oracle's exported primitives (each pinned bit-exact against the OpenCV 4.13 binary by
  tests/test_oracle_vs_cv2.py) — works anywhere the prebuilt .so travels, needs no cv2;
* `use_cv2_primitives()`: ctypes callbacks into the real cv2 of this container, so the reference's own
  cell loop / octree / ordering / orientation / BRIEF sampler run over real OpenCV results.  This is
  "the reference itself run here" for ORBextractor (tests/test_ref_extractor.py,
  tests/golden/make_golden.py).

/root/reference does not exist on the GPU box: nothing imports this module at run time there except
to load the prebuilt .so (never to rebuild it).
"""
import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

from . import KP_DTYPE, _p, build as _build_oracle

_DIR = Path(__file__).resolve().parent
_SO = _DIR / "_ref" / "libref_orbextractor.so"
REFERENCE = Path("/root/reference")
_lib = None
_keep = []  # keeps ctypes callbacks alive


class Prims(C.Structure):
    _fields_ = [
        ("resize_linear", C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int)),
        ("border101", C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p)),
        ("gaussian_blur7", C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.c_int, C.c_void_p)),
        ("fast_atan2", C.CFUNCTYPE(C.c_float, C.c_float, C.c_float)),
        ("fast", C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                             C.c_void_p, C.c_int)),
    ]


def available():
    return _SO.exists() or (REFERENCE / "src" / "ORBextractor.cc").exists()


def build(force=False):
    """Compile oracle/_ref when the reference sources are present (this container); otherwise use the
    prebuilt file that travelled with the snapshot."""
    src = REFERENCE / "src" / "ORBextractor.cc"
    if src.exists():
        _build_oracle()
        subprocess.check_call(["make", "-C", str(_DIR), "ref"] + (["-B"] if force else ["-s"]),
                              stdout=subprocess.DEVNULL)
    return _SO if _SO.exists() else None


def lib():
    global _lib
    if _lib is None:
        if not _SO.exists():
            build()
        if not _SO.exists():
            raise FileNotFoundError("oracle/_ref/libref_orbextractor.so: not built and /root/reference absent")
        _lib = C.CDLL(str(_SO))
        _lib.ref_create.restype = C.c_void_p
        _lib.ref_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        _lib.ref_destroy.argtypes = [C.c_void_p]
        _lib.ref_extract_roi.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int,
                                         C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                         C.POINTER(C.c_int)]
        _lib.ref_levels.argtypes = [C.c_void_p]
        _lib.ref_tables.argtypes = [C.c_void_p] * 5
        _lib.ref_level_size.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        _lib.ref_get_pyramid.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_int]
        _lib.ref_quota_umax.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        _lib.ref_set_prims.argtypes = [C.c_void_p]
    return _lib


class RefExtractor:
    """ORB_SLAM3::ORBextractor as the reference compiles it (ORBextractor.h:49-59)."""

    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7):
        self.L = lib()
        self.nfeatures, self.nlevels = nfeatures, nlevels
        self.h = self.L.ref_create(nfeatures, scale_factor, nlevels, ini_th, min_th)
        sf, inv, s2, is2 = (np.zeros(nlevels, np.float32) for _ in range(4))
        self.L.ref_tables(self.h, _p(sf), _p(inv), _p(s2), _p(is2))
        nf, um = np.zeros(nlevels, np.int32), np.zeros(16, np.int32)
        self.L.ref_quota_umax(self.h, _p(nf), _p(um))
        self.scale_factors, self.inv_scale_factors = sf, inv
        self.level_sigma2, self.inv_level_sigma2 = s2, is2
        self.features_per_level, self.umax = nf, um

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ref_destroy(self.h)
            self.h = None

    def extract(self, img, lap=(0, 0), roi=None):
        """operator().  `roi=(x, y, w, h)` passes a sub-rectangle of `img` as the reference would receive a
        ROI cv::Mat (level 0's border then comes from the parent, ORBextractor.cc:1190).
        Returns (keypoints[KP_DTYPE], descriptors[n,32], monoIndex)."""
        img = np.ascontiguousarray(img, np.uint8)
        if img.size == 0:
            n = C.c_int(0)
            mono = self.L.ref_extract_roi(self.h, None, 0, 0, 0, 0, 0, 0, 0, lap[0], lap[1], None, None, 0,
                                          C.byref(n))
            return np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8), mono
        ph, pw = img.shape
        x, y, w, hgt = roi if roi is not None else (0, 0, pw, ph)
        cap = self.nfeatures + self.nlevels * (3 + 4 * (w // max(hgt, 1) + 2)) + 64
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = C.c_int(0)
        mono = self.L.ref_extract_roi(self.h, _p(img), pw, ph, pw, x, y, w, hgt, lap[0], lap[1], _p(kps),
                                      _p(desc), cap, C.byref(n))
        assert mono != -2, "capacity"
        return kps[:n.value].copy(), desc[:n.value].copy(), mono

    def level_size(self, level):
        w, h = C.c_int(), C.c_int()
        assert self.L.ref_level_size(self.h, level, C.byref(w), C.byref(h)) == 0
        return w.value, h.value

    def pyramid(self, level, with_border=False):
        w, h = self.level_size(level)
        if with_border:
            w, h = w + 38, h + 38
        out = np.zeros((h, w), np.uint8)
        assert self.L.ref_get_pyramid(self.h, level, _p(out), w, int(with_border)) == 0
        return out


def use_oracle_primitives():
    lib().ref_set_prims(None)
    _keep.clear()


def use_cv2_primitives():
    """Route the reference's cv:: calls to the real OpenCV binary (cv2) of this container."""
    import cv2
    cv2.setNumThreads(1)

    def _img(ptr, w, h):
        return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint8)), shape=(h, w))

    def resize_linear(src, sw, sh, dst, dw, dh):
        _img(dst, dw, dh)[:] = cv2.resize(_img(src, sw, sh), (dw, dh), interpolation=cv2.INTER_LINEAR)

    def border101(src, w, h, b, dst):
        _img(dst, w + 2 * b, h + 2 * b)[:] = cv2.copyMakeBorder(
            _img(src, w, h), b, b, b, b, cv2.BORDER_REFLECT_101 | cv2.BORDER_ISOLATED)

    def gaussian_blur7(src, w, h, dst):
        _img(dst, w, h)[:] = cv2.GaussianBlur(_img(src, w, h).copy(), (7, 7), 2, sigmaY=2,
                                              borderType=cv2.BORDER_REFLECT_101)

    def fast_atan2(y, x):
        return cv2.fastAtan2(y, x)

    dets = {}

    def fast(img, w, h, x0, y0, x1, y1, threshold, out, cap):
        det = dets.get(threshold)
        if det is None:
            det = dets[threshold] = cv2.FastFeatureDetector_create(
                threshold, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
        kps = det.detect(_img(img, w, h)[y0:y1, x0:x1])
        o = np.ctypeslib.as_array(C.cast(out, C.POINTER(C.c_int32)), shape=(cap, 3))
        for i, k in enumerate(kps[:cap]):
            assert k.size == 7.0 and k.angle == -1.0 and k.octave == 0 and k.class_id == -1
            o[i] = (int(k.pt[0]), int(k.pt[1]), int(k.response))
        return len(kps)

    p = Prims()
    fields = dict(resize_linear=resize_linear, border101=border101, gaussian_blur7=gaussian_blur7,
                  fast_atan2=fast_atan2, fast=fast)
    for (name, ftype) in Prims._fields_:
        cb = ftype(fields[name])
        _keep.append(cb)
        setattr(p, name, cb)
    _keep.append(p)
    lib().ref_set_prims(C.byref(p))
