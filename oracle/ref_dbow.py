"""oracle/_ref/libref_dbow.so — the REFERENCE's own Thirdparty/DBoW2 compiled here, unmodified — TEST INFRASTRUCTURE ONLY.

`make -C oracle ref_dbow` compiles BowVector.cpp, FeatureVector.cpp, FORB.cpp, ScoringObject.cpp, DUtils/Random.cpp,
DUtils/Timestamp.cpp and the TemplatedVocabulary.h template where they lie under /root/reference/Thirdparty/DBoW2, over a
header shim for OpenCV core (cv::Mat as a byte matrix; FileStorage / FileNode declared, never opened) and
boost::serialization (oracle/ref_shim_dbow).  tests/test_ref_dbow.py compares loadFromTextFile, transform (BowVector,
FeatureVector, the per-feature descent), saveToTextFile and score with the oracle's restatement and with the product's
host-side text parser — that is what pins row (f).3.

/root/reference does not exist on the GPU box: there the prebuilt .so (git-ignored, travels with the snapshot) is
loaded, never rebuilt."""
import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

from . import _p

_DIR = Path(__file__).resolve().parent
_SO = _DIR / "_ref" / "libref_dbow.so"
REFERENCE = Path("/root/reference")
_lib = None


def available():
    return _SO.exists() or (REFERENCE / "Thirdparty" / "DBoW2" / "DBoW2" / "TemplatedVocabulary.h").exists()


def build(force=False):
    if (REFERENCE / "Thirdparty" / "DBoW2" / "DBoW2" / "TemplatedVocabulary.h").exists():
        subprocess.check_call(["make", "-C", str(_DIR), "ref_dbow"] + (["-B"] if force else ["-s"]),
                              stdout=subprocess.DEVNULL)
    return _SO if _SO.exists() else None


def lib():
    global _lib
    if _lib is None:
        if not _SO.exists():
            build()
        if not _SO.exists():
            raise FileNotFoundError("oracle/_ref/libref_dbow.so: not built and /root/reference absent")
        L = C.CDLL(str(_SO))
        L.ref_voc_load_text.restype = C.c_void_p
        L.ref_voc_load_text.argtypes = [C.c_char_p]
        L.ref_voc_free.argtypes = [C.c_void_p]
        L.ref_voc_save_text.argtypes = [C.c_void_p, C.c_char_p]
        L.ref_voc_info.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_voc_nodes.argtypes = [C.c_void_p] * 8
        L.ref_voc_transform.restype = C.c_int
        L.ref_voc_transform.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 6
        L.ref_voc_transform_features.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 3
        L.ref_voc_score.restype = C.c_double
        L.ref_voc_score.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        _lib = L
    return _lib


class Vocabulary:
    """ORBVocabulary (TemplatedVocabulary<FORB::TDescriptor, FORB>) loaded by the reference's loadFromTextFile."""

    def __init__(self, path):
        self._h = lib().ref_voc_load_text(str(path).encode())
        if not self._h:
            raise ValueError("loadFromTextFile returned false")
        info = np.zeros(6, np.int32)
        lib().ref_voc_info(self._h, _p(info))
        self.k, self.L, self.scoring, self.weighting, self.n_nodes, self.n_words = (int(v) for v in info)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().ref_voc_free(self._h)
            self._h = None

    def nodes(self):
        n = self.n_nodes
        o = dict(node_parent=np.zeros(n, np.int32), node_word=np.zeros(n, np.int32), node_weight=np.zeros(n, np.float64),
                 node_desc=np.zeros((n, 32), np.uint8), child_off=np.zeros(n + 1, np.int32),
                 child_idx=np.zeros(max(n - 1, 1), np.int32), word_node=np.zeros(max(self.n_words, 1), np.int32))
        lib().ref_voc_nodes(self._h, _p(o["node_parent"]), _p(o["node_word"]), _p(o["node_weight"]), _p(o["node_desc"]),
                            _p(o["child_off"]), _p(o["child_idx"]), _p(o["word_node"]))
        o["child_idx"] = o["child_idx"][:o["child_off"][-1]]
        o["word_node"] = o["word_node"][:self.n_words]
        return o

    def save_text(self, path):
        lib().ref_voc_save_text(self._h, str(path).encode())

    def transform(self, desc, levelsup):
        d = np.ascontiguousarray(desc, np.uint8)
        n = d.shape[0]
        bw, bv = np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.float64)
        node, off, idx = np.zeros(max(n, 1), np.int32), np.zeros(n + 1, np.int32), np.zeros(max(n, 1), np.int32)
        nn = C.c_int(0)
        k = lib().ref_voc_transform(self._h, _p(d), n, int(levelsup), _p(bw), _p(bv), _p(node), _p(off), _p(idx),
                                    C.byref(nn))
        nn = nn.value
        fw, fwt, fn = np.zeros(n, np.int32), np.zeros(n, np.float64), np.zeros(n, np.int32)
        lib().ref_voc_transform_features(self._h, _p(d), n, int(levelsup), _p(fw), _p(fwt), _p(fn))
        return {"feat_word": fw, "feat_node": fn, "feat_weight": fwt, "bow_word": bw[:k].copy(), "bow_val": bv[:k].copy(),
                "fv": (node[:nn].copy(), off[:nn + 1].copy(), idx[:off[nn]].copy())}

    def score(self, desc1, desc2):
        a, b = np.ascontiguousarray(desc1, np.uint8), np.ascontiguousarray(desc2, np.uint8)
        return lib().ref_voc_score(self._h, _p(a), a.shape[0], _p(b), b.shape[0])
