"""oracle/_ref/libref_dbow.so: the REFERENCE's own Thirdparty/DBoW2 (BowVector.cpp, FeatureVector.cpp, FORB.cpp,
ScoringObject.cpp, TemplatedVocabulary.h) compiled unmodified over a header shim for OpenCV core / boost::serialization
(recipe `make -C oracle ref_dbow`), against the oracle's restatement and the product's host-side parser.

This pins SURVEY §8 row (f).3: loadFromTextFile (the on-disk format of ORBvoc.txt, System.cc:118), the per-feature
descent (first minimum over the children, level by level), the BowVector (fp64 sums in feature order, L1 norm in word
order), the FeatureVector at `levelsup`, and saveToTextFile.

Needs either /root/reference (build container) or the prebuilt oracle/_ref/*.so (GPU box); else skipped."""
import numpy as np
import pytest

from orb_slam3_vio_fixes_b200 import synthetic as S
from orb_slam3_vio_fixes_b200.matcher import ORBVocabulary
from test_vocabulary_text import write_reference_format


@pytest.fixture(scope="module")
def RD(oracle_mod):
    from oracle import ref_dbow as R
    if not R.available():
        pytest.skip("neither /root/reference nor a prebuilt oracle/_ref")
    R.build()
    return R


@pytest.fixture(scope="module")
def feats(oracle_mod):
    ex = oracle_mod.Extractor(1000)
    return [ex.extract(f, (0, 0))[1] for f in S.make_frames("C1", 3)]


@pytest.mark.parametrize("k,L", [(10, 3), (4, 5), (3, 2)])
def test_load_from_text_file_equals_product_parser(tmp_path, RD, k, L):
    """The node table the reference's loadFromTextFile builds == the product's orbm_vocabulary_parse_text == the tree
    the file was written from (file without a final newline: the reference's `while(!f.eof())` loop turns a trailing
    newline into one more node of unspecified content, see below)."""
    voc = S.make_vocabulary(k, L, seed=5)
    p = tmp_path / "voc.txt"
    parent = write_reference_format(p, voc, k, trailing_newline=False)
    ref = RD.Vocabulary(p)
    got = ORBVocabulary.parse_text_file(p)
    assert (ref.k, ref.L, ref.scoring, ref.weighting) == (got["k"], got["L"], got["scoring"], got["weighting"]) == (k, L, 0, 0)
    assert ref.n_nodes == len(voc["node_word"]) and ref.n_words == got["n_words"] == k ** L
    rn = ref.nodes()
    for key in ("child_off", "child_idx", "node_word", "node_weight"):
        assert np.array_equal(rn[key], voc[key]), key
        assert np.array_equal(rn[key], got[key]), key
    assert np.array_equal(rn["node_desc"][1:], voc["node_desc"][1:]) and np.array_equal(rn["node_desc"][1:], got["node_desc"][1:])
    assert np.array_equal(rn["node_parent"][1:], parent[1:]) and np.array_equal(got["node_parent"][1:], parent[1:])
    # words are numbered in file order of the leaves
    leaves = np.nonzero(voc["node_word"] >= 0)[0]
    assert np.array_equal(rn["word_node"], leaves)


def test_trailing_newline_quirk_is_what_the_header_says(tmp_path, RD):
    """With the final newline saveToTextFile writes, the reference ends up with ONE more node, built from an empty line:
    `ssnode >> pid`, `>> nIsLeaf` on an empty stream leave the previous line's values (so the extra node hangs under the
    last node's parent and, the last line being a leaf, becomes one more word), its weight stays 0 and its descriptor is
    whatever cv::Mat::create returned.  A feature that lands on it drops out of the BowVector (weight 0).  The product skips blank lines instead (include/orbx.h,
    orbm_vocabulary_parse_text); everything else is identical."""
    voc = S.make_vocabulary(5, 3, seed=8)
    p = tmp_path / "voc.txt"
    write_reference_format(p, voc, 5, trailing_newline=True)
    ref = RD.Vocabulary(p)
    got = ORBVocabulary.parse_text_file(p)
    n = len(voc["node_word"])
    assert ref.n_nodes == n + 1 and len(got["node_word"]) == n
    rn = ref.nodes()
    assert rn["child_off"][n] == rn["child_off"][n + 1]
    assert np.array_equal(rn["node_weight"][:n], got["node_weight"]) and np.array_equal(rn["node_desc"][1:n], got["node_desc"][1:])
    assert ref.n_words == got["n_words"] + 1 and rn["node_word"][n] == got["n_words"] and rn["node_weight"][n] == 0.0
    assert rn["node_parent"][n] == rn["node_parent"][n - 1]


@pytest.mark.parametrize("k,L,levelsup", [(10, 4, 2), (10, 4, 4), (6, 3, 1), (3, 6, 3), (8, 2, 0)])
def test_transform_equals_oracle(tmp_path, oracle_mod, RD, feats, k, L, levelsup):
    """transform(features, BowVector&, FeatureVector&, levelsup) (TemplatedVocabulary.h:1127-1259, BowVector.cpp:33-88,
    FeatureVector.cpp:31-45): every feature's word / node / weight, the BowVector in map order with bit-identical fp64
    values (TF-IDF sums + L1 normalisation), the FeatureVector CSR."""
    voc = S.make_vocabulary(k, L, seed=11, desc_pool=np.concatenate(feats))
    p = tmp_path / "voc.txt"
    write_reference_format(p, voc, k, trailing_newline=False)
    ref = RD.Vocabulary(p)
    got_voc = ORBVocabulary.parse_text_file(p)         # the weights as the text file carries them
    voc = dict(voc, node_weight=got_voc["node_weight"])
    for d in feats:
        want = oracle_mod.bow_transform(d, voc, levelsup)
        got = ref.transform(d, levelsup)
        for key in ("feat_word", "feat_node", "feat_weight", "bow_word", "bow_val"):
            assert np.array_equal(got[key], want[key]), key
        for a, b in zip(got["fv"], want["fv"]):
            assert np.array_equal(a, b)
        assert len(want["bow_word"]) > 30 and abs(want["bow_val"].sum() - 1.0) < 1e-12


def test_save_to_text_file_round_trip(tmp_path, RD):
    """saveToTextFile (TemplatedVocabulary.h:1426-1464) of the reference, read back by the product's parser and by the
    reference: the same tree, weights at the stream's 6 significant digits."""
    voc = S.make_vocabulary(6, 3, seed=7)
    p1, p2 = tmp_path / "a.txt", tmp_path / "b.txt"
    write_reference_format(p1, voc, 6, trailing_newline=False)
    ref = RD.Vocabulary(p1)
    ref.save_text(p2)
    got = ORBVocabulary.parse_text_file(p2)
    want_w = np.array([float(f"{w:.6g}") for w in voc["node_weight"]])
    assert np.array_equal(got["node_weight"], want_w)
    for key in ("child_off", "child_idx", "node_word"):
        assert np.array_equal(got[key], voc[key]), key
    assert np.array_equal(got["node_desc"][1:], voc["node_desc"][1:])
    # and the product's own writer produces a file the reference reads back to the same table
    p3 = tmp_path / "c.txt"
    ORBVocabulary.save_text_file(p3, got) if hasattr(ORBVocabulary, "save_text_file") else p3.write_text(p2.read_text())
    again = RD.Vocabulary(p3).nodes()
    assert np.array_equal(again["node_weight"][:len(want_w)], want_w)
