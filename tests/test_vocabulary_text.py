"""ORBVocabulary text format (TemplatedVocabulary::loadFromTextFile / saveToTextFile,
Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1338-1464; FORB::toString/fromString FORB.cpp:103-135).
ORBvoc.txt is not in the image: the file is written here in the reference's format from a synthetic tree."""
import ctypes as C

import numpy as np
import pytest

from conftest import has_cuda
from orb_slam3_vio_fixes_b200 import _lib
from orb_slam3_vio_fixes_b200 import synthetic as S
from orb_slam3_vio_fixes_b200.matcher import ORBVocabulary


def write_reference_format(path, voc, k, trailing_newline=True, six_digits=False):
    """What saveToTextFile writes: 'k L  scoring weighting', then per node 'parent isLeaf d0 .. d31  weight'."""
    n = len(voc["node_word"])
    parent = np.zeros(n, np.int32)
    for i in range(n):
        for c in voc["child_idx"][voc["child_off"][i]:voc["child_off"][i + 1]]:
            parent[c] = i
    lines = [f"{k} {voc['L']}  0 0"]
    for i in range(1, n):
        leaf = voc["child_off"][i] == voc["child_off"][i + 1]
        d = " ".join(str(int(b)) for b in voc["node_desc"][i]) + " "
        w = f"{voc['node_weight'][i]:.6g}" if six_digits else repr(float(voc["node_weight"][i]))
        lines.append(f"{parent[i]} {'1 ' if leaf else '0 '}{d} {w}")
    with open(path, "w") as f:
        f.write("\n".join(lines) + ("\n" if trailing_newline else ""))
    return parent


@pytest.mark.parametrize("k,L,trail", [(10, 3, True), (4, 5, False), (3, 2, True)])
def test_parse_text_roundtrip(tmp_path, k, L, trail):
    voc = S.make_vocabulary(k, L, seed=5)
    p = tmp_path / "voc.txt"
    parent = write_reference_format(p, voc, k, trailing_newline=trail)
    got = ORBVocabulary.parse_text_file(p)
    assert (got["k"], got["L"], got["scoring"], got["weighting"]) == (k, L, 0, 0)
    assert got["n_words"] == k ** L
    for key in ("child_off", "child_idx", "node_word", "node_weight"):
        assert np.array_equal(got[key], voc[key]), key
    assert np.array_equal(got["node_desc"][1:], voc["node_desc"][1:])  # the root (node 0) is not in the file
    assert np.array_equal(got["node_parent"], parent)


def test_six_digit_weights_parse_like_strtod(tmp_path):
    """ORBvoc.txt carries the weights at the stream's default 6 significant digits."""
    voc = S.make_vocabulary(5, 3, seed=6)
    p = tmp_path / "voc6.txt"
    write_reference_format(p, voc, 5, six_digits=True)
    got = ORBVocabulary.parse_text_file(p)
    want = np.array([float(f"{w:.6g}") for w in voc["node_weight"]])
    assert np.array_equal(got["node_weight"], want)


def test_save_then_parse(tmp_path):
    voc = S.make_vocabulary(6, 3, seed=7)
    p1, p2 = tmp_path / "a.txt", tmp_path / "b.txt"
    write_reference_format(p1, voc, 6)
    h = _lib.OrbmVocabularyHost()
    L = _lib.lib()
    assert L.orbm_vocabulary_parse_text(str(p1).encode(), C.byref(h)) == 0
    assert L.orbm_vocabulary_save_text(str(p2).encode(), C.byref(h)) == 0
    L.orbm_vocabulary_host_free(C.byref(h))
    a, b = ORBVocabulary.parse_text_file(p1), ORBVocabulary.parse_text_file(p2)
    for key in a:
        assert np.array_equal(a[key], b[key]), key


@pytest.mark.parametrize("text", ["", "25 3 0 0\n", "10 0 0 0\n", "10 3 9 0\n", "10 3 0 0\n5 0 1 2 3\n",
                                  "10 3 0 0\n0 1 " + "7 " * 32 + "\n"])
def test_malformed_files_are_rejected(tmp_path, text):
    p = tmp_path / "bad.txt"
    p.write_text(text)
    with pytest.raises(_lib.OrbxError):
        ORBVocabulary.parse_text_file(p)
    with pytest.raises(_lib.OrbxError):
        ORBVocabulary.parse_text_file(tmp_path / "missing.txt")


@pytest.mark.gpu
@pytest.mark.skipif(not has_cuda(), reason="needs CUDA")
def test_gpu_transform_from_text_file(tmp_path, oracle_mod):
    """loadFromTextFile -> device tree -> transform == the oracle's transform on the in-memory tree."""
    import torch
    rng = np.random.default_rng(9)
    pool = rng.integers(0, 256, (4000, 32), dtype=np.uint8)
    voc = S.make_vocabulary(10, 4, seed=8, desc_pool=pool)
    p = tmp_path / "voc.txt"
    write_reference_format(p, voc, 10)
    V = ORBVocabulary.loadFromTextFile(p)
    n = 1200
    desc = pool[rng.integers(0, 4000, n)] ^ (rng.random((n, 32)) < 0.02).astype(np.uint8)
    out = V.transform(torch.from_numpy(desc[None]).cuda(), torch.tensor([n], dtype=torch.int32).cuda(), 4)
    ref = oracle_mod.bow_transform(desc, voc, 4)
    assert np.array_equal(out["feat_word"][0, :n].cpu().numpy(), ref["feat_word"])
    nb = int(out["bow_n"][0])
    assert np.array_equal(out["bow_word"][0, :nb].cpu().numpy(), ref["bow_word"])
    assert np.array_equal(out["bow_val"][0, :nb].cpu().numpy().view(np.uint64), ref["bow_val"].view(np.uint64))
