"""ctypes binding of liborbx.so — the C ABI declared in include/orbx.h.

There is no CPU fallback: if the library is missing or no CUDA device is present every
compute entry point fails loudly (ORBX_E_NODEVICE / OSError).
"""
import ctypes as C
from pathlib import Path

import numpy as np

PKG = Path(__file__).resolve().parent
SO = PKG / "liborbx.so"

ORBX_OK = 0
ERRORS = {-1: "ORBX_E_INVALID", -2: "ORBX_E_CUDA", -3: "ORBX_E_CAPACITY", -4: "ORBX_E_NOMEM",
          -5: "ORBX_E_EMPTY", -6: "ORBX_E_NODEVICE", -7: "ORBX_E_INTERNAL"}
FLAG_STRICT_FP = 1
FLAG_STAGE_TIMING = 2
FLAG_FUSED_BLUR = 4
FLAG_LEVEL_BLUR = 8
STAGES = ("pyramid", "fast", "octree", "order", "describe")
TIE_FIRST_MIN, TIE_LAST_MIN = 0, 1
STATIC_SCRATCH_BYTES = 4 << 20  # ORBM_STATIC_SCRATCH_BYTES (orbm_search_by_projection_static)
PROJ_MAPPOINTS, PROJ_LASTFRAME = 0, 1
BOW_KF_FRAME, BOW_KF_KF = 0, 1
MAX_LEVELS = 12
GRID_COLS, GRID_ROWS = 64, 48

# cv::KeyPoint layout (orbx_keypoint), 28 bytes
KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
# orbm_result, 12 bytes
RES_DTYPE = np.dtype([("best_idx", "<i4"), ("best_dist", "<i2"), ("second_dist", "<i2"),
                      ("best_level", "i1"), ("second_level", "i1"), ("pad", "<i2")])

# every symbol include/orbx.h declares (tests/test_abi.py checks the library exports them all)
SYMBOLS = [
    "orbx_create", "orbx_destroy", "orbx_max_keypoints", "orbx_extract", "orbx_extract_batch_device",
    "orbx_extract_batch_host", "orbx_last_outputs_device", "orbx_get_pyramid", "orbx_get_pyramid_all", "orbx_pyramid_device", "orbx_level_size",
    "orbx_get_scale_tables", "orbx_stage_times", "orbx_check_status", "orbx_debug_blur_level", "orbx_debug_sincos_checksum", "orbx_debug_level_keypoints", "orbx_debug_octree", "orbm_argmin2_csr", "orbm_grid_build",
    "orbm_search_area", "orbm_queries_shift", "orbm_search_by_projection", "orbm_search_by_projection_th", "orbm_search_by_projection_static", "orbm_compute_image_bounds", "orbm_stereo_rowband", "orbm_stereo_refine", "orbm_rgbd_depth", "orbm_knn2_dense",
    "orbm_fuse_candidates", "orbm_descriptor_distance_pairs", "orbm_search_for_initialization", "orbm_search_by_bow", "orbm_search_for_triangulation", "orbm_bow_transform", "orbm_distinctive_descriptors",
    "orbm_undistort_keypoints", "orbm_is_in_frustum",
    "orbm_search_by_projection_rig", "orbm_search_by_bow_rig", "orbm_search_for_triangulation_rig", "orbm_stereo_fisheye_matches",
    "orbm_vocabulary_parse_text", "orbm_vocabulary_save_text", "orbm_vocabulary_host_free", "orbm_vocabulary_upload",
    "orbm_vocabulary_release", "orbm_vocabulary_load_text",
    "orbx_version", "orbx_last_error",
]


class OrbxConfig(C.Structure):
    _fields_ = [("nfeatures", C.c_int32), ("scale_factor", C.c_float), ("nlevels", C.c_int32),
                ("ini_th_fast", C.c_int32), ("min_th_fast", C.c_int32), ("width", C.c_int32),
                ("height", C.c_int32), ("max_batch", C.c_int32), ("device", C.c_int32),
                ("flags", C.c_uint32)]


class OrbmFeatureVector(C.Structure):  # orbm_feature_vector
    _fields_ = [("d_node", C.c_void_p), ("d_off", C.c_void_p), ("d_idx", C.c_void_p), ("d_nn", C.c_void_p)]


class OrbmVocabulary(C.Structure):  # orbm_vocabulary
    _fields_ = [("d_child_off", C.c_void_p), ("d_child_idx", C.c_void_p), ("d_node_desc", C.c_void_p),
                ("d_node_word", C.c_void_p), ("d_node_weight", C.c_void_p), ("n_nodes", C.c_int32),
                ("L", C.c_int32)]


class OrbmVocabularyHost(C.Structure):  # orbm_vocabulary_host
    _fields_ = [("k", C.c_int32), ("L", C.c_int32), ("scoring", C.c_int32), ("weighting", C.c_int32),
                ("n_nodes", C.c_int32), ("n_words", C.c_int32), ("child_off", C.POINTER(C.c_int32)),
                ("child_idx", C.POINTER(C.c_int32)), ("node_desc", C.POINTER(C.c_uint8)),
                ("node_word", C.POINTER(C.c_int32)), ("node_weight", C.POINTER(C.c_double)),
                ("node_parent", C.POINTER(C.c_int32))]


class OrbmRigCamera(C.Structure):  # orbm_rig_camera
    _fields_ = [("d_kps", C.c_void_p), ("d_desc", C.c_void_p), ("d_n", C.c_void_p), ("kp_stride", C.c_int32),
                ("d_cell_off", C.c_void_p), ("d_cell_idx", C.c_void_p), ("d_feat_locked", C.c_void_p),
                ("d_partner", C.c_void_p)]


class OrbmCamera(C.Structure):  # orbm_camera
    _fields_ = [("model", C.c_int32), ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float),
                ("k", C.c_float * 4)]


class OrbxError(RuntimeError):
    def __init__(self, code, where):
        self.code = code
        msg = ""
        try:
            msg = lib().orbx_last_error().decode()
        except Exception:
            pass
        super().__init__(f"{where}: {ERRORS.get(code, code)} {msg}")


_lib = None


def lib():
    """Loads liborbx.so (building it first if the sources are newer).  Raises if absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not SO.exists():
        from . import build as _b
        _b.build()
    L = C.CDLL(str(SO))
    vp, i32, f32, sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
    L.orbx_create.argtypes = [C.POINTER(OrbxConfig), C.POINTER(vp)]
    L.orbx_destroy.argtypes = [vp]
    L.orbx_destroy.restype = None
    L.orbx_max_keypoints.argtypes = [vp]
    L.orbx_extract.argtypes = [vp, vp, i32, i32, sz, i32, i32, vp, vp, i32, C.POINTER(i32), C.POINTER(i32)]
    L.orbx_extract_batch_device.argtypes = [vp, vp, sz, sz, i32, i32, i32, vp, vp, i32, vp, vp, vp]
    L.orbx_extract_batch_host.argtypes = [vp, vp, i32, i32, i32, vp, vp, i32, vp, vp]
    L.orbx_last_outputs_device.argtypes = [vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(i32)]
    L.orbx_get_pyramid.argtypes = [vp, i32, i32, vp, sz, i32]
    L.orbx_get_pyramid_all.argtypes = [vp, i32, i32, C.POINTER(vp), vp, vp]
    L.orbx_pyramid_device.argtypes = [vp, i32, i32, C.POINTER(vp), C.POINTER(sz), C.POINTER(i32), C.POINTER(i32)]
    L.orbx_level_size.argtypes = [vp, i32, C.POINTER(i32), C.POINTER(i32)]
    L.orbx_get_scale_tables.argtypes = [vp, vp, vp, vp, vp, vp]
    L.orbx_stage_times.argtypes = [vp, vp]
    L.orbx_check_status.argtypes = [vp, vp]
    L.orbx_debug_sincos_checksum.argtypes = [i32, C.c_uint32, C.c_uint64, C.c_uint32, vp, i32]
    L.orbx_debug_blur_level.argtypes = [vp, i32, i32, vp, C.c_size_t]
    L.orbx_debug_level_keypoints.argtypes = [vp, i32, i32, i32, vp, i32, C.POINTER(i32)]
    L.orbx_debug_octree.argtypes = [vp, i32, vp, i32, vp, i32, C.POINTER(i32)]
    L.orbm_argmin2_csr.argtypes = [vp, i32, vp, i32, vp, vp, vp, i32, vp, vp]
    L.orbm_grid_build.argtypes = [vp, vp, i32, i32, f32, f32, f32, f32, vp, vp, vp]
    L.orbm_search_area.argtypes = [vp, vp, vp, i32, vp, vp, f32, f32, f32, f32, vp, vp, vp, vp, vp, vp, vp,
                                   i32, i32, i32, vp, vp]
    L.orbm_descriptor_distance_pairs.argtypes = [vp, vp, vp, vp, i32, vp, vp]
    L.orbm_fuse_candidates.argtypes = [vp, vp, vp, i32, vp, vp, f32, f32, f32, f32, vp, vp, i32, vp, vp, vp, vp, vp,
                                       vp, vp, i32, i32, vp, vp]
    L.orbm_queries_shift.argtypes = [vp, vp, vp, i32, i32, vp, vp, f32, vp, i32, i32, i32, vp, vp, vp, vp, vp, vp,
                                     vp, vp]
    L.orbm_stereo_rowband.argtypes = [vp, vp, vp, vp, vp, vp, i32, i32, vp, i32, f32, vp, vp, vp]
    L.orbm_search_by_projection.argtypes = [vp, vp, vp, i32, i32, vp, vp, f32, f32, f32, f32, vp, vp, vp, vp, vp, vp, vp,
                                            vp, vp, i32, vp, vp, vp, i32, f32, i32, vp, vp, vp, vp, vp]
    L.orbm_compute_image_bounds.argtypes = [i32, i32, f32, f32, f32, f32, vp, i32, vp]
    L.orbm_search_by_projection_static.argtypes = [vp, vp, vp, i32, i32, vp, vp, f32, f32, f32, f32, vp, vp, vp, vp, vp, vp,
                                                   vp, vp, vp, i32, vp, vp, vp, i32, f32, i32, vp, vp, vp, vp]
    L.orbm_search_by_projection_th.argtypes = [vp, vp, vp, i32, i32, vp, vp, f32, f32, f32, f32, vp, vp, vp, vp, vp, vp,
                                               vp, vp, vp, i32, vp, vp, vp, i32, f32, i32, i32, vp, vp, vp, vp, vp]
    L.orbm_stereo_refine.argtypes = [vp, vp, vp, vp, vp, i32, i32, vp, vp, f32, f32, vp, vp, vp, vp, vp]
    L.orbm_rgbd_depth.argtypes = [vp, vp, vp, i32, i32, vp, sz, sz, i32, i32, f32, vp, vp, vp]
    L.orbm_knn2_dense.argtypes = [vp, i32, vp, i32, vp, vp, vp]
    pfv, pvoc = C.POINTER(OrbmFeatureVector), C.POINTER(OrbmVocabulary)
    L.orbm_search_by_bow.argtypes = [i32, vp, vp, vp, vp, pfv, vp, vp, vp, vp, pfv, i32, i32, i32, f32, i32, vp, vp,
                                     vp, vp]
    L.orbm_search_for_triangulation.argtypes = [vp, vp, vp, vp, vp, pfv, vp, vp, vp, vp, vp, pfv, i32, i32, i32, vp,
                                                vp, vp, vp, i32, i32, i32, i32, vp, vp, vp]
    L.orbm_bow_transform.argtypes = [pvoc, vp, vp, i32, i32, i32, vp, vp, vp, pfv, i32, vp, vp, vp, vp]
    L.orbm_search_for_initialization.argtypes = [vp, vp, vp, vp, vp, vp, i32, i32, vp, vp, f32, f32, f32, f32, vp, i32,
                                                 f32, i32, vp, vp, vp, vp]
    L.orbm_distinctive_descriptors.argtypes = [vp, vp, vp, i32, vp, vp, vp]
    L.orbm_undistort_keypoints.argtypes = [vp, vp, i32, i32, f32, f32, f32, f32, vp, i32, vp, vp]
    L.orbm_is_in_frustum.argtypes = [vp, vp, vp, vp, vp, i32, i32, vp, C.POINTER(OrbmCamera), f32, f32, f32, f32, f32,
                                     f32, f32, i32, i32, vp, vp, vp, vp, vp, vp, vp, f32, vp, vp, vp, vp, vp]
    prc = C.POINTER(OrbmRigCamera)
    L.orbm_search_by_projection_rig.argtypes = [i32, prc, prc, i32, f32, f32, f32, f32] + [vp] * 12 + [i32, vp, vp, f32,
                                                                                                   i32] + [vp] * 7
    L.orbm_search_by_bow_rig.argtypes = [vp, vp, vp, vp, pfv, vp, vp, vp, vp, pfv, i32, i32, i32, f32, i32, vp, vp, vp,
                                         vp, vp]
    pcam = C.POINTER(OrbmCamera)
    L.orbm_search_for_triangulation_rig.argtypes = [vp, vp, vp, vp, vp, pfv, vp, vp, vp, vp, vp, pfv, i32, i32, i32, vp,
                                                    pcam, pcam, vp, vp, i32, i32, i32, i32, vp, vp, vp]
    L.orbm_stereo_fisheye_matches.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, pcam, pcam, vp, vp, vp, i32, vp,
                                              vp, vp, vp, vp]
    pvh = C.POINTER(OrbmVocabularyHost)
    L.orbm_vocabulary_parse_text.argtypes = [C.c_char_p, pvh]
    L.orbm_vocabulary_save_text.argtypes = [C.c_char_p, pvh]
    L.orbm_vocabulary_host_free.argtypes = [pvh]
    L.orbm_vocabulary_host_free.restype = None
    L.orbm_vocabulary_upload.argtypes = [pvh, pvoc]
    L.orbm_vocabulary_release.argtypes = [pvoc]
    L.orbm_vocabulary_release.restype = None
    L.orbm_vocabulary_load_text.argtypes = [C.c_char_p, pvoc, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    L.orbx_version.restype = C.c_char_p
    L.orbx_last_error.restype = C.c_char_p
    _lib = L
    return L


def check(rc, where):
    if rc != ORBX_OK:
        raise OrbxError(rc, where)
