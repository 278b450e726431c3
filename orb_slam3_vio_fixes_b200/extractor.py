"""Host-side mirror of ORB_SLAM3::ORBextractor (reference include/ORBextractor.h:43-112) on top
of the C ABI.  Same constructor arguments, same call semantics (`operator()` returns monoIndex
and fills keypoints + descriptors; empty image → -1), same getters; `mvImagePyramid` /
`GetPyramid` read the device pyramid back on demand.

The C++ drop-in with the reference's exact signatures is include/orbx_compat/ORBextractor.h;
this Python class is what the parity tests and bench.py drive.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import KP_DTYPE


class DevView:
    """A library-owned device buffer seen as (pointer, shape, device) — the three things the
    batched matcher wrappers read from a torch tensor."""

    def __init__(self, ptr, shape, device):
        self._ptr, self.shape, self._dev = int(ptr), tuple(shape), device

    def data_ptr(self):
        return self._ptr

    @property
    def device(self):
        import torch
        return torch.device(self._dev)


class ORBextractor:
    HARRIS_SCORE, FAST_SCORE = 0, 1

    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, *, max_batch=1, device=0,
                 strict_fp=False, stage_timing=False, blur_mode=None):
        self.nfeatures, self.scaleFactor, self.nlevels = int(nfeatures), float(scaleFactor), int(nlevels)
        self.iniThFAST, self.minThFAST = int(iniThFAST), int(minThFAST)
        self.max_batch, self.device = int(max_batch), int(device)
        # blur_mode: None = chosen per plan, "fused" = per-keypoint blur in shared memory,
        # "level" = whole levels blurred once (ORBX_FLAG_FUSED_BLUR / ORBX_FLAG_LEVEL_BLUR)
        assert blur_mode in (None, "fused", "level")
        self.flags = (_lib.FLAG_STRICT_FP if strict_fp else 0) | (_lib.FLAG_STAGE_TIMING if stage_timing else 0) | \
            {None: 0, "fused": _lib.FLAG_FUSED_BLUR, "level": _lib.FLAG_LEVEL_BLUR}[blur_mode]
        self._h = None
        self._size = None

    # ---- plan / handle -------------------------------------------------------------------
    def _ensure(self, width, height):
        if self._h is not None and self._size == (width, height):
            return
        self.close()
        L = _lib.lib()
        cfg = _lib.OrbxConfig(self.nfeatures, self.scaleFactor, self.nlevels, self.iniThFAST, self.minThFAST,
                              width, height, self.max_batch, self.device, self.flags)
        h = C.c_void_p()
        _lib.check(L.orbx_create(C.byref(cfg), C.byref(h)), "orbx_create")
        self._h, self._size = h, (width, height)
        self.max_keypoints = L.orbx_max_keypoints(h)

    def plan(self, width, height):
        """Creates the device plan for a frame size up front (otherwise done on first call)."""
        self._ensure(int(width), int(height))
        return self

    def close(self):
        if getattr(self, "_h", None) is not None:
            _lib.lib().orbx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- operator() ----------------------------------------------------------------------------
    def __call__(self, image, mask=None, vLappingArea=(0, 0)):
        """Returns (monoIndex, keypoints[KP_DTYPE], descriptors[n,32] uint8).  `mask` is ignored
        exactly like the reference (ORBextractor.h:55)."""
        if image is None or image.size == 0:
            return -1, np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        assert image.dtype == np.uint8 and image.ndim == 2, "CV_8UC1 expected (ORBextractor.cc:1094)"
        hgt, w = image.shape
        self._ensure(w, hgt)
        cap = self.max_keypoints
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n, mono = C.c_int(0), C.c_int(0)
        img = image if image.strides[1] == 1 else np.ascontiguousarray(image)
        rc = _lib.lib().orbx_extract(self._h, img.ctypes.data, w, hgt, img.strides[0], int(vLappingArea[0]),
                                     int(vLappingArea[1]), kps.ctypes.data, desc.ctypes.data, cap,
                                     C.byref(n), C.byref(mono))
        _lib.check(rc, "orbx_extract")
        return mono.value, kps[:n.value].copy(), desc[:n.value].copy()

    def extract_batch_host(self, frames, vLappingArea=(0, 0), out=None):
        """frames: uint8 [B,H,W] host array (pinned for full speed).  Returns (kps[B,cap], desc[B,cap,32],
        n[B], mono[B]) as numpy arrays; copies in and out are inside the call (chunked so that H2D,
        kernels and D2H overlap).  `out` = the same four arrays preallocated (e.g. views of pinned
        torch tensors; kps may be any C-contiguous [B,cap,28-byte] buffer)."""
        b, hgt, w = frames.shape
        self._ensure(w, hgt)
        cap = self.max_keypoints
        if out is None:
            kps = np.zeros((b, cap), KP_DTYPE)
            desc = np.zeros((b, cap, 32), np.uint8)
            n = np.zeros(b, np.int32)
            mono = np.zeros(b, np.int32)
        else:
            kps, desc, n, mono = out
            assert kps.nbytes == b * cap * 28 and desc.nbytes == b * cap * 32 and n.nbytes == 4 * b == mono.nbytes
            assert all(a.flags["C_CONTIGUOUS"] for a in out)
        frames = np.ascontiguousarray(frames)
        rc = _lib.lib().orbx_extract_batch_host(self._h, frames.ctypes.data, b, int(vLappingArea[0]),
                                                int(vLappingArea[1]), kps.ctypes.data, desc.ctypes.data, cap,
                                                n.ctypes.data, mono.ctypes.data)
        _lib.check(rc, "orbx_extract_batch_host")
        return kps, desc, n, mono

    def last_outputs_device(self, batch):
        """Device-resident copies of what the last extract_batch_host / __call__ returned, as
        pointer views (data_ptr / shape / device) the ORBmatcher batched cores accept."""
        ptrs = [C.c_void_p() for _ in range(4)]
        cap = C.c_int(0)
        _lib.check(_lib.lib().orbx_last_outputs_device(self._h, *[C.byref(p) for p in ptrs], C.byref(cap)),
                   "orbx_last_outputs_device")
        dev = f"cuda:{self.device}"
        return {"kps": DevView(ptrs[0].value, (batch, cap.value, 7), dev),
                "desc": DevView(ptrs[1].value, (batch, cap.value, 32), dev),
                "n": DevView(ptrs[2].value, (batch,), dev), "mono": DevView(ptrs[3].value, (batch,), dev)}

    def extract_batch_device(self, frames, vLappingArea=(0, 0), out=None, stream=None):
        """frames: torch uint8 CUDA tensor [B,H,W] (row stride = pitch).  Asynchronous on `stream`
        (default: torch's current stream).  Returns dict of CUDA tensors:
        kps [B,cap,7] float32 view of cv::KeyPoint records, desc [B,cap,32] uint8, n [B], mono [B]."""
        import torch
        assert frames.is_cuda and frames.dtype == torch.uint8 and frames.dim() == 3 and frames.stride(2) == 1
        b, hgt, w = frames.shape
        self._ensure(w, hgt)
        cap = self.max_keypoints
        if out is None:
            out = self.alloc_outputs(b, frames.device)
        st = stream if stream is not None else torch.cuda.current_stream(frames.device).cuda_stream
        rc = _lib.lib().orbx_extract_batch_device(self._h, frames.data_ptr(), frames.stride(1), frames.stride(0), b,
                                                  int(vLappingArea[0]), int(vLappingArea[1]), out["kps"].data_ptr(),
                                                  out["desc"].data_ptr(), cap, out["n"].data_ptr(),
                                                  out["mono"].data_ptr(), st)
        _lib.check(rc, "orbx_extract_batch_device")
        return out

    def alloc_outputs(self, batch, device):
        import torch
        cap = self.max_keypoints
        return {"kps": torch.zeros((batch, cap, 7), dtype=torch.float32, device=device),
                "desc": torch.zeros((batch, cap, 32), dtype=torch.uint8, device=device),
                "n": torch.zeros(batch, dtype=torch.int32, device=device),
                "mono": torch.zeros(batch, dtype=torch.int32, device=device)}

    @staticmethod
    def kps_from_tensor(t):
        """[.., 7] float32 tensor (device layout) → numpy KP_DTYPE records."""
        a = np.ascontiguousarray(t.detach().cpu().numpy())
        return a.view(KP_DTYPE).reshape(a.shape[:-1])

    # ---- getters (ORBextractor.h:61-81) ----------------------------------------------------------
    def GetLevels(self):
        return self.nlevels

    def GetScaleFactor(self):
        return np.float32(self.scaleFactor)

    def _tables(self):
        if self._h is None:
            raise RuntimeError("scale tables need a planned size: call plan(width, height) or extract once")
        t = [np.zeros(self.nlevels, np.float32) for _ in range(4)] + [np.zeros(self.nlevels, np.int32)]
        _lib.check(_lib.lib().orbx_get_scale_tables(self._h, *[a.ctypes.data for a in t]), "orbx_get_scale_tables")
        return t

    def GetScaleFactors(self):
        return self._tables()[0]

    def GetInverseScaleFactors(self):
        return self._tables()[1]

    def GetScaleSigmaSquares(self):
        return self._tables()[2]

    def GetInverseScaleSigmaSquares(self):
        return self._tables()[3]

    def features_per_level(self):
        return self._tables()[4]

    # ---- pyramid ("GetPyramid" / mvImagePyramid, ORBextractor.h:83) ------------------------------
    def level_size(self, level):
        w, h = C.c_int(), C.c_int()
        _lib.check(_lib.lib().orbx_level_size(self._h, level, C.byref(w), C.byref(h)), "orbx_level_size")
        return w.value, h.value

    def GetPyramid(self, level, with_border=False, frame=0):
        w, h = self.level_size(level)
        if with_border:
            w, h = w + 38, h + 38
        out = np.zeros((h, w), np.uint8)
        _lib.check(_lib.lib().orbx_get_pyramid(self._h, frame, level, out.ctypes.data, w, int(with_border)),
                   "orbx_get_pyramid")
        return out

    def GetPyramidAll(self, with_border=True, frame=0):
        """Every level of the last call in one device->host transfer (orbx_get_pyramid_all): list of level images
        (copies of the handle's pinned staging buffer)."""
        L = _lib.lib()
        base = C.c_void_p()
        ofs = (C.c_size_t * self.nlevels)()
        strd = (C.c_size_t * self.nlevels)()
        _lib.check(L.orbx_get_pyramid_all(self._h, frame, int(with_border), C.byref(base), ofs, strd), "orbx_get_pyramid_all")
        out = []
        for level in range(self.nlevels):
            w, h = self.level_size(level)
            if with_border:
                w, h = w + 38, h + 38
            buf = (C.c_uint8 * (h * strd[level])).from_address(base.value + ofs[level])
            out.append(np.frombuffer(buf, np.uint8).reshape(h, strd[level])[:, :w].copy())
        return out

    @property
    def mvImagePyramid(self):
        """Level images of the last call, each a view into its bordered parent like the reference's
        ROI Mats (so `img[-19:...]`-style parent access is mirrored by `.base`)."""
        out = []
        for level in range(self.nlevels):
            full = self.GetPyramid(level, with_border=True)
            out.append(full[19:-19, 19:-19])
        return out

    def stage_times(self):
        """Device milliseconds of each stage of the last call: dict stage -> ms (needs stage_timing=True)."""
        ms = np.zeros(len(_lib.STAGES), np.float32)
        _lib.check(_lib.lib().orbx_stage_times(self._h, ms.ctypes.data), "orbx_stage_times")
        return dict(zip(_lib.STAGES, ms.tolist()))

    def debug_octree(self, level, xys):
        """DistributeOctTree of `level` on explicit keys [n,3] int32 (x, y relative to the border, response)."""
        xys = np.ascontiguousarray(xys, np.int32).reshape(-1, 3)
        out = np.zeros((len(xys) + 64, 3), np.int32)
        n = C.c_int(0)
        _lib.check(_lib.lib().orbx_debug_octree(self._h, level, xys.ctypes.data, len(xys), out.ctypes.data, len(out),
                                                C.byref(n)), "orbx_debug_octree")
        return out[:n.value].copy()

    def debug_blur_level(self, level, frame=0):
        """The blurred level the descriptor stage sampled (level-blur mode), as an [h,w] array."""
        w, h = self.level_size(level)
        out = np.zeros((h, w), np.uint8)
        _lib.check(_lib.lib().orbx_debug_blur_level(self._h, frame, level, out.ctypes.data, w), "orbx_debug_blur_level")
        return out

    def debug_level_keypoints(self, level, stage, frame=0):
        cap = 1 << 16
        while True:
            buf = np.zeros((cap, 3), np.int32)
            n = C.c_int(0)
            _lib.check(_lib.lib().orbx_debug_level_keypoints(self._h, frame, level, stage, buf.ctypes.data, cap,
                                                             C.byref(n)), "orbx_debug_level_keypoints")
            if n.value <= cap:
                return buf[:n.value].copy()
            cap = n.value
